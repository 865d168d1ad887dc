// One warp = one env. Everything an env needs for a physics tick lives in this warp's shared-memory block; lanes
// take the role of bodies / DoFs / collision pairs / solver rows depending on the phase.
//
// tick() restates one pybullet stepSimulation at default parameters in a GPU-native
// formulation: joint-space mass matrix by direct composite sums + in-register Cholesky instead of the recursive
// articulated-body pass, and a contact-space (Delassus) projected Gauss-Seidel with one solver row per lane instead
// of Bullet's velocity-space sweep. Both are algebraically the same iteration the CPU oracle performs.
#pragma once
#include "dev_collide.cuh"
#include "model.h"
#include "params.h"

#ifdef UG_TRACE
#include <cstdio>
#define UG_TR(...) do { if (lane == 0 && blockIdx.x == 0) printf(__VA_ARGS__); } while (0)
#else
#define UG_TR(...)
#endif
#define UG_RMAX 80          // solver rows per env (<= 24 joint rows + 3*16 contact rows + 6 rolling rows + 1 pad)
#define UG_JW 32            // row width: 12 robot dofs + 3 free bodies x 6, padded
#define UG_FULL 0xffffffffu

struct Con { int sa, sb; float posA[3], posB[3], n[3]; float dist, mu, rest, roll; };

struct Smem {
  float st[UG_STATE];
  float q[UG_NDOF], qd[UG_NDOF], tgt[8];
  float bR[UG_NDOF][9], bp[UG_NDOF][3], ax[UG_NDOF][3], cm[UG_NDOF][3], Iw[UG_NDOF][6];
  float w[UG_NDOF][3], vp[UG_NDOF][3], al[UG_NDOF][3], ap[UG_NDOF][3], wr[UG_NDOF][6];
  float M[UG_NDOF][UG_NDOF];
  float fpos[UG_MAXFREE][3], fquat[UG_MAXFREE][4], fR[UG_MAXFREE][9], flin[UG_MAXFREE][3], fang[UG_MAXFREE][3], fIinv[UG_MAXFREE][6];
  float vstar[UG_JW], g[UG_JW];
  int ncon, nrows, flags, pad0;
  Con con[UG_MAXC];
  int desc[UG_RMAX];
  float lam[UG_RMAX];
  float J[UG_RMAX][UG_JW];
  float A[UG_RMAX][UG_RMAX];
};

// flags (also slot 31 of the state record)
#define FL_ARM 1
#define FL_GRIP 2
#define FL_APPROACHED 4
#define FL_GRASPED 8
#define FL_BODY0 16

// row descriptor: type | idx << 4 | dir << 12
enum { RT_LIM_LO = 0, RT_LIM_HI = 1, RT_MOTOR = 2, RT_GEAR = 3, RT_NORMAL = 4, RT_ROLL = 5, RT_FRIC = 6, RT_PAD = 7 };

__device__ __constant__ int c_gear_dof[5] = {7, 8, 9, 10, 11};
__device__ __constant__ float c_gear_ratio[5] = {1.0f, -1.0f, -1.0f, 1.0f, -1.0f};

// ---------------------------------------------------------------------------------------------------------------
// Forward kinematics of the 12 movable bodies, level by level (arm chain, then the four gripper branches).
UG_DEV void fk_bodies(Smem& s, const DevModel& M, int lane) {
  int depth = lane < UG_NDOF ? __popc(M.anc[lane]) - 1 : -1;
  for (int lv = 0; lv < 8; lv++) {
    if (depth == lv) {
      int d = lane, p = M.par[d];
      m3 Rp; f3 pp;
      if (p >= 0) { Rp = ldm(s.bR[p]); pp = ld3(s.bp[p]); }
      else { Rp.m[0] = 1; Rp.m[1] = 0; Rp.m[2] = 0; Rp.m[3] = 0; Rp.m[4] = 1; Rp.m[5] = 0; Rp.m[6] = 0; Rp.m[7] = 0; Rp.m[8] = 1; pp = mk3(0, 0, 0); }
      m3 R0 = mulm(Rp, ldm(M.jrot[d]));
      f3 pos = pp + mulv(Rp, ld3(M.jpos[d]));
      float sn, cs; sincosf(s.q[d], &sn, &cs);
      int a = M.axis[d], j = (a + 1) % 3, k = (a + 2) % 3;
      f3 cj = col(R0, j), ck = col(R0, k);
      f3 nj = cj * cs + ck * sn, nk = ck * cs - cj * sn;
      m3 R = R0;
      R.m[j] = nj.x; R.m[3 + j] = nj.y; R.m[6 + j] = nj.z; R.m[k] = nk.x; R.m[3 + k] = nk.y; R.m[6 + k] = nk.z;
      stm(s.bR[d], R); st3(s.bp[d], pos); st3(s.ax[d], col(R, a));
      st3(s.cm[d], pos + mulv(R, ld3(M.com[d])));
      rot_sym(R, M.I[d], s.Iw[d]);
    }
    if (lv == 0 && lane >= UG_NDOF && lane < UG_NDOF + UG_MAXFREE) {
      int b = lane - UG_NDOF;
      stm(s.fR[b], quat_to_mat(mkq(s.fquat[b][0], s.fquat[b][1], s.fquat[b][2], s.fquat[b][3])));
    }
    __syncwarp();
  }
}

// Unconstrained joint accelerations -> s.vstar[0..11] = qd + dt*qdd; s.M = M^-1. Free bodies -> s.vstar[12..29].
UG_DEV void dynamics(Smem& s, const DevModel& M, int lane) {
  const bool isd = lane < UG_NDOF;
  const int d = isd ? lane : 0;
  const uint32_t anc = M.anc[d];
  f3 pd = ld3(s.bp[d]);
  // --- angular velocity of body d and linear velocity of its joint origin
  f3 w = mk3(0, 0, 0), vp = mk3(0, 0, 0);
  for (int j = 0; j < UG_NDOF; j++) if (anc >> j & 1) {
    f3 aj = ld3(s.ax[j]) * s.qd[j];
    w = w + aj;
    if (j != d) vp = vp + cross(aj, pd - ld3(s.bp[j]));
  }
  if (isd) { st3(s.w[d], w); st3(s.vp[d], vp); }
  __syncwarp();
  // --- velocity-product angular acceleration
  f3 al = mk3(0, 0, 0);
  for (int j = 0; j < UG_NDOF; j++) if (anc >> j & 1) {
    int pj = M.par[j];
    if (pj >= 0) al = al + cross(ld3(s.w[pj]), ld3(s.ax[j])) * s.qd[j];
  }
  if (isd) st3(s.al[d], al);
  __syncwarp();
  // --- velocity-product linear acceleration of the joint origin: walk up the chain
  f3 ap = mk3(0, 0, 0);
  {
    int child = d, j = M.par[d];
    while (j >= 0) {
      f3 r = ld3(s.bp[child]) - ld3(s.bp[j]);
      f3 wj = ld3(s.w[j]);
      ap = ap + cross(ld3(s.al[j]), r) + cross(wj, cross(wj, r));
      child = j; j = M.par[j];
    }
  }
  // --- wrench needed on body d (about its merged COM) for zero joint acceleration: inertia, gyroscopic, gravity
  //     and Bullet's per-link velocity damping (row P2)
  f3 cmd = ld3(s.cm[d]);
  f3 rc = cmd - pd;
  f3 ac = ap + cross(al, rc) + cross(w, cross(w, rc));
  f3 Iw_w = symv(s.Iw[d], w);
  f3 f = (ac - mk3(0, 0, M.gravity)) * M.mass[d];
  f3 n = symv(s.Iw[d], al) + cross(w, Iw_w);
  {
    m3 R = ldm(s.bR[d]);
    float wn = norm3(w);
    f3 wl = tmulv(R, w);
    for (int k = 0; k < M.nsub[d]; k++) {
      const DevSub& S = M.subl[d][k];
      f3 cs = pd + mulv(R, ld3(S.com));
      f3 vc = vp + cross(w, cs - pd);
      f3 fd = vc * (S.mass * (UGP_LIN_DAMPING + UGP_LIN_DAMPING * norm3(vc)));
      f3 td = mulv(R, symv(S.I, wl)) * (UGP_ANG_DAMPING + UGP_ANG_DAMPING * wn);
      f = f + fd; n = n + td + cross(cs - cmd, fd);
    }
  }
  if (isd) { s.wr[d][0] = n.x; s.wr[d][1] = n.y; s.wr[d][2] = n.z; s.wr[d][3] = f.x; s.wr[d][4] = f.y; s.wr[d][5] = f.z; }
  // --- zero the mass matrix
  for (int i = lane; i < UG_NDOF * UG_NDOF; i += 32) (&s.M[0][0])[i] = 0.f;
  __syncwarp();
  // --- bias torque and mass-matrix row of dof d from composite sums over its subtree, about its own joint origin
  float tau = 0.f;
  {
    const uint32_t sub = M.sub[d];
    f3 ad = ld3(s.ax[d]);
    float mc = 0.f; f3 h = mk3(0, 0, 0); float IO[6] = {0, 0, 0, 0, 0, 0};
    for (int b = 0; b < UG_NDOF; b++) if (sub >> b & 1) {
      f3 r = ld3(s.cm[b]) - pd;
      f3 nb = mk3(s.wr[b][0], s.wr[b][1], s.wr[b][2]), fb = mk3(s.wr[b][3], s.wr[b][4], s.wr[b][5]);
      tau += dot(ad, nb + cross(r, fb));
      float mb = M.mass[b], rr = dot(r, r);
      mc += mb; h = h + r * mb;
      IO[0] += s.Iw[b][0] + mb * (rr - r.x * r.x); IO[1] += s.Iw[b][1] - mb * r.x * r.y; IO[2] += s.Iw[b][2] - mb * r.x * r.z;
      IO[3] += s.Iw[b][3] + mb * (rr - r.y * r.y); IO[4] += s.Iw[b][4] - mb * r.y * r.z; IO[5] += s.Iw[b][5] + mb * (rr - r.z * r.z);
    }
    f3 nn = symv(IO, ad), ff = cross(ad, h);
    if (isd) for (int j = 0; j < UG_NDOF; j++) if (anc >> j & 1) {
      f3 aj = ld3(s.ax[j]);
      float v = dot(aj, nn) + dot(cross(aj, pd - ld3(s.bp[j])), ff);
      s.M[d][j] = v; s.M[j][d] = v;
    }
  }
  __syncwarp();
#ifdef UG_TRACE
  if (lane == 0 && blockIdx.x == 0) {
    printf("q:"); for (int i = 0; i < 12; i++) printf(" %g", s.q[i]); printf("\nMdiag:"); for (int i = 0; i < 12; i++) printf(" %g", s.M[i][i]);
    for (int i = 0; i < 12; i++) { printf("\nM[%d]:", i); for (int j = 0; j < 12; j++) printf(" %g", s.M[i][j]); }
    printf("\ntau0 %g cm0 %g %g %g bp5 %g %g %g\n", tau, s.cm[0][0], s.cm[0][1], s.cm[0][2], s.bp[5][0], s.bp[5][1], s.bp[5][2]);
  }
#endif
  // --- Cholesky in registers (every lane redundantly), then lane k solves for column k of M^-1
  {
    float L[UG_NDOF][UG_NDOF];
#pragma unroll
    for (int i = 0; i < UG_NDOF; i++)
#pragma unroll
      for (int j = 0; j <= i; j++) L[i][j] = s.M[i][j];
#pragma unroll
    for (int k = 0; k < UG_NDOF; k++) {
      float dk = L[k][k];
#pragma unroll
      for (int t = 0; t < k; t++) dk -= L[k][t] * L[k][t];
      float inv = rsqrtf(dk);
      UG_TR("chol k=%d dk=%g inv=%g\n", k, dk, inv);
      L[k][k] = inv;  // store the reciprocal of the diagonal
#pragma unroll
      for (int i = k + 1; i < UG_NDOF; i++) {
        float v = L[i][k];
#pragma unroll
        for (int t = 0; t < k; t++) v -= L[i][t] * L[k][t];
        L[i][k] = v * inv;
      }
    }
    float x[UG_NDOF];
#pragma unroll
    for (int i = 0; i < UG_NDOF; i++) {
      float v = (i == lane) ? 1.0f : 0.0f;
#pragma unroll
      for (int t = 0; t < i; t++) v -= L[i][t] * x[t];
      x[i] = v * L[i][i];
    }
#pragma unroll
    for (int i = UG_NDOF - 1; i >= 0; i--) {
      float v = x[i];
#pragma unroll
      for (int t = i + 1; t < UG_NDOF; t++) v -= L[t][i] * x[t];
      x[i] = v * L[i][i];
    }
    __syncwarp();
    if (isd) {
#pragma unroll
      for (int i = 0; i < UG_NDOF; i++) s.M[i][lane] = x[i];
    }
    s.g[lane] = isd ? -tau : 0.f;
  }
  __syncwarp();
  if (isd) {
    float qdd = 0.f;
#pragma unroll
    for (int j = 0; j < UG_NDOF; j++) qdd += s.M[d][j] * s.g[j];
    s.vstar[d] = fminf(fmaxf(s.qd[d] + UGP_DT * qdd, -UGP_MAX_COORD_VEL), UGP_MAX_COORD_VEL);
  } else if (lane < UG_NDOF + UG_MAXFREE) {
    int b = lane - UG_NDOF;
    bool act = (s.flags & (FL_BODY0 << b)) != 0 && b < M.nfree;
    f3 lin = ld3(s.flin[b]), ang = ld3(s.fang[b]);
    m3 R = ldm(s.fR[b]);
    f3 I = ld3(M.finertia[b]);
    if (act) {
      f3 acc = mk3(0, 0, M.gravity) - lin * (UGP_LIN_DAMPING + UGP_LIN_DAMPING * norm3(lin));
      f3 wl = tmulv(R, ang);
      f3 Iwl = mk3(I.x * wl.x, I.y * wl.y, I.z * wl.z);
      f3 tq = -(Iwl * (UGP_ANG_DAMPING + UGP_ANG_DAMPING * norm3(wl))) - cross(wl, Iwl);
      f3 alf = mk3(tq.x / I.x, tq.y / I.y, tq.z / I.z);
      lin = lin + acc * UGP_DT; ang = ang + mulv(R, alf) * UGP_DT;
      float inv[6] = {1.0f / I.x, 0.f, 0.f, 1.0f / I.y, 0.f, 1.0f / I.z};
      rot_sym(R, inv, s.fIinv[b]);
    } else {
      lin = mk3(0, 0, 0); ang = mk3(0, 0, 0);
      for (int k = 0; k < 6; k++) s.fIinv[b][k] = 0.f;
    }
    float* v = s.vstar + UG_NDOF + 6 * b;
    v[0] = lin.x; v[1] = lin.y; v[2] = lin.z; v[3] = ang.x; v[4] = ang.y; v[5] = ang.z;
  } else if (lane >= 30) {
    s.vstar[lane] = 0.f;
  }
  __syncwarp();
#ifdef UG_TRACE
  if (lane == 0 && blockIdx.x == 0) { printf("vstar:"); for (int i = 0; i < 12; i++) printf(" %g", s.vstar[i]); printf("\nMinvdiag:"); for (int i = 0; i < 12; i++) printf(" %g", s.M[i][i]); printf("\n"); }
#endif
}

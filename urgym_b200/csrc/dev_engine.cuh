// One lane group (G = 16 or 32 lanes of a warp) = one env. Everything an env needs for a physics tick lives in the
// group's shared-memory block (~12 KB, so ~16 envs are resident per SM); lanes take the role of bodies / DoFs /
// collision pairs / solver rows depending on the phase.
//
// tick() restates one pybullet stepSimulation at default parameters in a GPU-native
// formulation: joint-space mass matrix by direct composite sums + in-register Cholesky instead of the recursive
// articulated-body pass, and a projected Gauss-Seidel that runs in "response space": joint rows (limits, motors,
// gears) act on 17 extended joint columns, contact rows on a triangular-packed Delassus block. Algebraically this is
// the iteration the CPU oracle performs (same rows, same order, same clamps).
#pragma once
#include "dev_collide.cuh"
#include "model.h"
#include "params.h"

#define UG_NCOL 17   // extended joint columns: 12 unit DoFs + 5 gear combinations e6 + ratio * e_child
#define UG_JMAX 24   // joint rows per tick: <= 12 limits + 7 motors + 5 gears

// Manifold point. Side A is a robot link (ra >= 0, pa = world point) or a free body (fa >= 0, pa = point relative
// to the body origin); side B is a free body (fb >= 0, rb relative) or static. n points from B to A.
// Restitution and rolling friction are non-zero in the Golf scenes only (club 0.1 x ball 0.1; the ball's rollingFriction) and Box-v0
// borrows `rest` for its motor row: the other scenes' records do without the two words (160 B per env: one more resident warp per SM).
template <bool FULL> struct ConT;
template <> struct ConT<true> {
  int shapes;   // sa | sb << 8
  int bodies;   // (fa + 1) | (fb + 1) << 8 | (ra + 1) << 16
  real pa[3], rb[3], n[3];
  real dist, mu, rest, roll;
};
template <> struct ConT<false> {
  int shapes, bodies;
  real pa[3], rb[3], n[3];
  real dist, mu;
};
UG_DEV real con_rest(const ConT<true>& c) { return c.rest; }
UG_DEV real con_rest(const ConT<false>&) { return 0._r; }
UG_DEV real con_roll(const ConT<true>& c) { return c.roll; }
UG_DEV real con_roll(const ConT<false>&) { return 0._r; }
UG_DEV void con_set_rest_roll(ConT<true>& c, real rest, real roll) { c.rest = rest; c.roll = roll; }
UG_DEV void con_set_rest_roll(ConT<false>&, real, real) {}

template <int NCAP, int RC, bool ANCH = false>
struct SmemT {
  static constexpr int NC_CAP = NCAP;  // manifold points kept per tick
  // the scene has world-anchored 1-DoF bodies (Box-v0's door and button): every branch that serves them is compiled under this flag,
  // so the kernels of the other scenes carry none of it (same instructions, same bits as before Box-v0 existed)
  static constexpr bool ANCHORS = ANCH;
  static constexpr bool FULLCON = ANCH || RC > 3 * NCAP;   // rolling-friction rows exist (Golf scenes) or Box-v0
  typedef ConT<FULLCON> Con;
  static constexpr int TTRI = RC * (RC + 1) / 2;  // packed lower triangle of the symmetric contact block
  // ---- state that lives across the ticks of one gym step
  real q[UG_NDOF], qd[UG_NDOF], tgt[8];
  real fpos[UG_MAXFREE][3], fquat[UG_MAXFREE][4], flin[UG_MAXFREE][3], fang[UG_MAXFREE][3];
  int flags, ncon, ovf, npseudo;   // npseudo: pseudo contacts at the head of con (anchored bodies' joint rows); ovf: ticks of this launch whose manifold points / rolling manifolds exceeded the capacity
  unsigned hit[UG_HITWORDS];     // pairs of the scene's pair list with at least one manifold point this tick (set before the cap)
  // ---- per tick
  real vstar[32];
  real bR[UG_NDOF][9], bp[UG_NDOF][3], ax[UG_NDOF][3];
  real fR[UG_MAXFREE][9], fIinv[UG_MAXFREE][6];
  Con con[NCAP];
  real lam[UG_JMAX + RC];
  // row descriptors in 16 bits (240 B per env less than as ints; together with the shorter manifold records: 14 instead of 13 resident warps)
  unsigned short jdesc[UG_JMAX];  // col | type << 8 | index << 12
  unsigned short jci4[UG_JMAX + 2];  // byte offset of row j's column inside an E row, at [j + 1] (one slack entry on each side)
  unsigned short cdesc[RC];       // contact | type << 8 | dir << 12
  // rows 0..16: extended M^-1 (17 x 17); rows 17..: robot velocity response of each contact row in the same 17 columns
  real E[(UG_NCOL + RC) * UG_NCOL];
  union alignas(16) {
    // Contact-space Delassus block T = J M^-1 J^T, symmetric, packed lower triangle: T(r, c) at r (r + 1) / 2 + c for c <= r.
    // A step of row r reads T(r, c) for the lane's own rows c: c <= r is contiguous in c, c > r sits at c (c + 1) / 2 + r, and
    // consecutive triangular numbers are distinct modulo 32 (16 o + l (l + 1) / 2 for c = 32 o + l), so both are conflict-free.
    real T[TTRI];
    struct {
      real M[UG_NDOF][UG_NDOF];  // mass matrix, then M^-1 (16-byte aligned rows for the row build)
      real cm[UG_NDOF][3], Iw[UG_NDOF][6];
      real w[UG_NDOF][3], vp[UG_NDOF][3], al[UG_NDOF][3], wr[UG_NDOF][6];
      real tauv[16];
    } d;
    real st[UG_STATE];
  } u;
};

// flags (also slot 31 of the state record)
#define FL_ARM 1
#define FL_GRIP 2
#define FL_APPROACHED 4
#define FL_GRASPED 8
#define FL_BODY0 16

// Pseudo contacts: joint rows of the anchored bodies (Box-v0 door / button limits, button motor) travel as manifold entries at
// the head of s.con so that they are solved between the robot's joint rows and the contact normals, as the checker does.
// Con::shapes of such an entry: UG_PSEUDO | kind (1 limit, 2 motor) | 4 when the row is angular (hinge).
#define UG_PSEUDO 0x10000
// linear part of an anchored or free body's generalized inverse inertia applied to a force direction
UG_DEV f3 lin_inv(const DevModel& M, int b, f3 J) {
  const int a = M.fanchor[b];
  if (a == 0) return J * (1.0_r / M.fmass[b]);
  if (a == 1) return mk3(0, 0, 0);
  const f3 ax = ld3(M.faxis[b]);
  return ax * (dot(ax, J) / M.fmass[b]);
}
// coordinate of an anchored body's single degree of freedom (angle about / offset along faxis)
template <class S>
UG_DEV real anchor_q(const S& s, const DevModel& M, int b) {
  const f3 ax = ld3(M.faxis[b]);
  if (M.fanchor[b] == 1) return 2.0_r * ratan2(s.fquat[b][0] * ax.x + s.fquat[b][1] * ax.y + s.fquat[b][2] * ax.z, s.fquat[b][3]);
  return dot(ld3(s.fpos[b]) - ld3(M.forigin[b]), ax);
}

enum { RT_LIM_LO = 0, RT_LIM_HI = 1, RT_MOTOR = 2, RT_GEAR = 3, RT_NORMAL = 4, RT_ROLL = 5, RT_FRIC = 6, RT_PAD = 7 };

__device__ __constant__ int c_gear_dof[5] = {7, 8, 9, 10, 11};
__device__ __constant__ real c_gear_ratio[5] = {1.0_r, -1.0_r, -1.0_r, 1.0_r, -1.0_r};

// ---------------------------------------------------------------------------------------------------------------
// Forward kinematics of the 12 movable bodies, level by level (arm chain, then the four gripper branches).
template <int G, class S>
UG_DEV void fk_bodies(S& s, const DevModel& M, const Gp<G>& gp) {
  const int l = gp.l;
  int depth = l < UG_NDOF ? __popc(M.anc[l]) - 1 : -1;
#pragma unroll 1
  for (int lv = 0; lv < 8; lv++) {
    if (depth == lv) {
      int d = l, p = M.par[d];
      m3 Rp; f3 pp;
      if (p >= 0) { Rp = ldm(s.bR[p]); pp = ld3(s.bp[p]); }
      else { Rp.m[0] = 1; Rp.m[1] = 0; Rp.m[2] = 0; Rp.m[3] = 0; Rp.m[4] = 1; Rp.m[5] = 0; Rp.m[6] = 0; Rp.m[7] = 0; Rp.m[8] = 1; pp = mk3(0, 0, 0); }
      m3 R0 = mulm(Rp, ldm(M.jrot[d]));
      f3 pos = pp + mulv(Rp, ld3(M.jpos[d]));
      real sn, cs; rsincos_fast(s.q[d], &sn, &cs);  // |q| <= pi: the fast path is accurate to ~5e-7 there
      int a = M.axis[d], j = (a + 1) % 3, k = (a + 2) % 3;
      f3 cj = col(R0, j), ck = col(R0, k);
      f3 nj = cj * cs + ck * sn, nk = ck * cs - cj * sn;
      m3 R = R0;
      R.m[j] = nj.x; R.m[3 + j] = nj.y; R.m[6 + j] = nj.z; R.m[k] = nk.x; R.m[3 + k] = nk.y; R.m[6 + k] = nk.z;
      stm(s.bR[d], R); st3(s.bp[d], pos); st3(s.ax[d], col(R, a));
    }
    if (lv == 0 && l >= UG_NDOF && l < UG_NDOF + UG_MAXFREE) {
      int b = l - UG_NDOF;
      stm(s.fR[b], quat_to_mat(mkq(s.fquat[b][0], s.fquat[b][1], s.fquat[b][2], s.fquat[b][3])));
    }
    gp.sync();
  }
}

// Unconstrained joint accelerations -> s.vstar[0..11] = qd + dt*qdd; extended M^-1 -> s.E rows 0..16 (and the
// aligned copy s.u.d.M). Free bodies -> s.vstar[12..29].
template <int G, class S>
UG_DEV void dynamics(S& s, const DevModel& M, const Gp<G>& gp) {
  const int l = gp.l;
  const bool isd = l < UG_NDOF;
  const int d = isd ? l : 0;
  const uint32_t anc = M.anc[d];
  f3 pd = ld3(s.bp[d]);
  m3 Rd = ldm(s.bR[d]);
  // --- merged COM and world inertia of body d; angular velocity and linear velocity of its joint origin
  f3 cmd = pd + mulv(Rd, ld3(M.com[d]));
  if (isd) { st3(s.u.d.cm[d], cmd); rot_sym(Rd, M.I[d], s.u.d.Iw[d]); }
  f3 w = mk3(0, 0, 0), vp = mk3(0, 0, 0);
#pragma unroll 1
  for (int j = 0; j < UG_NDOF; j++) if (anc >> j & 1) {
    f3 aj = ld3(s.ax[j]) * s.qd[j];
    w = w + aj;
    if (j != d) vp = vp + cross(aj, pd - ld3(s.bp[j]));
  }
  if (isd) { st3(s.u.d.w[d], w); st3(s.u.d.vp[d], vp); }
  gp.sync();
  // --- velocity-product angular acceleration
  f3 al = mk3(0, 0, 0);
#pragma unroll 1
  for (int j = 0; j < UG_NDOF; j++) if (anc >> j & 1) {
    int pj = M.par[j];
    if (pj >= 0) al = al + cross(ld3(s.u.d.w[pj]), ld3(s.ax[j])) * s.qd[j];
  }
  if (isd) st3(s.u.d.al[d], al);
  gp.sync();
  // --- velocity-product linear acceleration of the joint origin: walk up the chain
  f3 ap = mk3(0, 0, 0);
  {
    int child = d, j = M.par[d];
#pragma unroll 1
    while (j >= 0) {
      f3 r = ld3(s.bp[child]) - ld3(s.bp[j]);
      f3 wj = ld3(s.u.d.w[j]);
      ap = ap + cross(ld3(s.u.d.al[j]), r) + cross(wj, cross(wj, r));
      child = j; j = M.par[j];
    }
  }
  // --- wrench needed on body d (about its merged COM) for zero joint acceleration: inertia, gyroscopic, gravity
  //     and Bullet's per-link velocity damping (row P2)
  f3 rc = cmd - pd;
  f3 ac = ap + cross(al, rc) + cross(w, cross(w, rc));
  f3 f = (ac - mk3(0, 0, M.gravity)) * M.mass[d];
  f3 n;
  {
    real Iwd[6]; rot_sym(Rd, M.I[d], Iwd);
    n = symv(Iwd, al) + cross(w, symv(Iwd, w));
    real wn = norm3(w);
    f3 wl = tmulv(Rd, w);
#pragma unroll 1
    for (int k = 0; k < M.nsub[d]; k++) {
      const DevSub& Sb = M.subl[d][k];
      f3 cs = pd + mulv(Rd, ld3(Sb.com));
      f3 vc = vp + cross(w, cs - pd);
      f3 fd = vc * (Sb.mass * (UGP_LIN_DAMPING + UGP_LIN_DAMPING * norm3(vc)));
      f3 td = mulv(Rd, symv(Sb.I, wl)) * (UGP_ANG_DAMPING + UGP_ANG_DAMPING * wn);
      f = f + fd; n = n + td + cross(cs - cmd, fd);
    }
  }
  if (isd) { real* o = s.u.d.wr[d]; o[0] = n.x; o[1] = n.y; o[2] = n.z; o[3] = f.x; o[4] = f.y; o[5] = f.z; }
  for (int i = l; i < UG_NDOF * UG_NDOF; i += G) (&s.u.d.M[0][0])[i] = 0._r;
  gp.sync();
  // --- bias torque and mass-matrix row of dof d from composite sums over its subtree, about its own joint origin
  real tau = 0._r;
  {
    const uint32_t sub = M.sub[d];
    f3 ad = ld3(s.ax[d]);
    f3 h = mk3(0, 0, 0); real IO[6] = {0, 0, 0, 0, 0, 0};
#pragma unroll 1
    for (int b = 0; b < UG_NDOF; b++) if (sub >> b & 1) {
      f3 r = ld3(s.u.d.cm[b]) - pd;
      const real* wb = s.u.d.wr[b]; const real* Ib = s.u.d.Iw[b];
      f3 nb = mk3(wb[0], wb[1], wb[2]), fb = mk3(wb[3], wb[4], wb[5]);
      tau += dot(ad, nb + cross(r, fb));
      real mb = M.mass[b], rr = dot(r, r);
      h = h + r * mb;
      IO[0] += Ib[0] + mb * (rr - r.x * r.x); IO[1] += Ib[1] - mb * r.x * r.y; IO[2] += Ib[2] - mb * r.x * r.z;
      IO[3] += Ib[3] + mb * (rr - r.y * r.y); IO[4] += Ib[4] - mb * r.y * r.z; IO[5] += Ib[5] + mb * (rr - r.z * r.z);
    }
    f3 nn = symv(IO, ad), ff = cross(ad, h);
    if (isd)
#pragma unroll 1
    for (int j = 0; j < UG_NDOF; j++) if (anc >> j & 1) {
      f3 aj = ld3(s.ax[j]);
      real v = dot(aj, nn) + dot(cross(aj, pd - ld3(s.bp[j])), ff);
      s.u.d.M[d][j] = v; s.u.d.M[j][d] = v;
    }
  }
  gp.sync();
  // --- Cholesky factor in shared memory (rolled: lane i produces row i of column k), then lane k solves for column k
  //     of M^-1 with the factor read back as broadcasts. Diagonal entries hold reciprocals.
  {
    real (*A)[UG_NDOF] = s.u.d.M;
#pragma unroll 1
    for (int k = 0; k < UG_NDOF; k++) {
      real dk = A[k][k];
#pragma unroll 1
      for (int t = 0; t < k; t++) dk -= A[k][t] * A[k][t];
      const real inv = rrsqrt(dk);
      if (l > k && l < UG_NDOF) {
        real v = A[l][k];
#pragma unroll 1
        for (int t = 0; t < k; t++) v -= A[l][t] * A[k][t];
        A[l][k] = v * inv;
      }
      gp.sync();
      if (l == k) A[k][k] = inv;
    }
    gp.sync();
    real x[UG_NDOF];
#pragma unroll
    for (int i = 0; i < UG_NDOF; i++) {
      real v = (i == l) ? 1.0_r : 0.0_r;
#pragma unroll
      for (int t = 0; t < i; t++) v -= A[i][t] * x[t];
      x[i] = v * A[i][i];
    }
#pragma unroll
    for (int i = UG_NDOF - 1; i >= 0; i--) {
      real v = x[i];
#pragma unroll
      for (int t = i + 1; t < UG_NDOF; t++) v -= A[t][i] * x[t];
      x[i] = v * A[i][i];
    }
    gp.sync();
    if (isd) {
      // M^-1 is symmetric: column l is row l. Unit columns, then the five gear columns of row l.
#pragma unroll
      for (int i = 0; i < UG_NDOF; i++) { s.u.d.M[l][i] = x[i]; s.E[l * UG_NCOL + i] = x[i]; }
#pragma unroll
      for (int gi = 0; gi < 5; gi++) s.E[l * UG_NCOL + 12 + gi] = x[6] + c_gear_ratio[gi] * x[7 + gi];
    }
    if (l < 16) s.u.d.tauv[l] = isd ? -tau : 0._r;
  }
  gp.sync();
  // gear rows 12..16 of the extended matrix: E[12+g][b] = E[6][b] + ratio_g * E[gear_dof_g][b]
  for (int b = l; b < UG_NCOL; b += G) {
#pragma unroll
    for (int gi = 0; gi < 5; gi++) s.E[(12 + gi) * UG_NCOL + b] = s.E[6 * UG_NCOL + b] + c_gear_ratio[gi] * s.E[(7 + gi) * UG_NCOL + b];
  }
  if (isd) {
    real qdd = 0._r;
#pragma unroll
    for (int j = 0; j < UG_NDOF; j++) qdd += s.u.d.M[d][j] * s.u.d.tauv[j];
    s.vstar[d] = rmin(rmax(s.qd[d] + UGP_DT * qdd, -UGP_MAX_COORD_VEL), UGP_MAX_COORD_VEL);
  } else if (l < UG_NDOF + UG_MAXFREE) {
    int b = l - UG_NDOF;
    bool act = (s.flags & (FL_BODY0 << b)) != 0 && b < M.nfree;
    f3 lin = ld3(s.flin[b]), ang = ld3(s.fang[b]);
    m3 R = ldm(s.fR[b]);
    f3 I = ld3(M.finertia[b]);
    const int anchor = S::ANCHORS ? M.fanchor[b] : 0;
    if (act && anchor == 1) {        // hinge (Box-v0 door): gravity torque of the merged COM about the axis, angular damping
      const f3 ax = ld3(M.faxis[b]);
      real w = dot(ang, ax);
      const real tq = dot(ax, cross(mulv(R, ld3(M.fcom[b])), mk3(0, 0, M.fmass[b] * M.gravity)));
      w += UGP_DT * (tq / M.finertia_axis[b] - w * (UGP_ANG_DAMPING + UGP_ANG_DAMPING * rabs(w)));
      ang = ax * w; lin = mk3(0, 0, 0);
      const real ii = 1.0_r / M.finertia_axis[b];
      real* o = s.fIinv[b];
      o[0] = ax.x * ax.x * ii; o[1] = ax.x * ax.y * ii; o[2] = ax.x * ax.z * ii; o[3] = ax.y * ax.y * ii; o[4] = ax.y * ax.z * ii; o[5] = ax.z * ax.z * ii;
    } else if (act && anchor == 2) { // slider (Box-v0 button): gravity and linear damping along the axis, no rotation
      const f3 ax = ld3(M.faxis[b]);
      real v = dot(lin, ax);
      v += UGP_DT * (ax.z * M.gravity - v * (UGP_LIN_DAMPING + UGP_LIN_DAMPING * rabs(v)));
      lin = ax * v; ang = mk3(0, 0, 0);
      for (int k = 0; k < 6; k++) s.fIinv[b][k] = 0._r;
    } else if (act) {
      f3 acc = mk3(0, 0, M.gravity) - lin * (UGP_LIN_DAMPING + UGP_LIN_DAMPING * norm3(lin));
      f3 wl = tmulv(R, ang);
      f3 Iwl = mk3(I.x * wl.x, I.y * wl.y, I.z * wl.z);
      f3 tq = -(Iwl * (UGP_ANG_DAMPING + UGP_ANG_DAMPING * norm3(wl))) - cross(wl, Iwl);
      f3 alf = mk3(tq.x / I.x, tq.y / I.y, tq.z / I.z);
      lin = lin + acc * UGP_DT; ang = ang + mulv(R, alf) * UGP_DT;
      real inv[6] = {1.0_r / I.x, 0._r, 0._r, 1.0_r / I.y, 0._r, 1.0_r / I.z};
      rot_sym(R, inv, s.fIinv[b]);
    } else {
      lin = mk3(0, 0, 0); ang = mk3(0, 0, 0);
      for (int k = 0; k < 6; k++) s.fIinv[b][k] = 0._r;
    }
    real* v = s.vstar + UG_NDOF + 6 * b;
    v[0] = lin.x; v[1] = lin.y; v[2] = lin.z; v[3] = ang.x; v[4] = ang.y; v[5] = ang.z;
  }
  gp.sync();
}

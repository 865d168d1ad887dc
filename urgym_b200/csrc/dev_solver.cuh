// Collision -> solver rows -> contact-space projected Gauss-Seidel -> integration, for one env per warp.
#pragma once
#include "dev_engine.cuh"

UG_DEV void shape_box(const Smem& s, const DevShape& S, BoxW& b) {
  m3 lR = ldm(S.lR); f3 lc = ld3(S.lc);
  if (S.body == 0) { m3 R = ldm(s.bR[S.rbody]); b.R = mulm(R, lR); b.c = ld3(s.bp[S.rbody]) + mulv(R, lc); }
  else if (S.body > 0) { m3 R = ldm(s.fR[S.body - 1]); b.R = mulm(R, lR); b.c = ld3(s.fpos[S.body - 1]) + mulv(R, lc); }
  else { b.R = lR; b.c = lc; }
  b.h = ld3(S.half);
}

// Lane per candidate pair; manifold points are appended in pair order and capped at UG_MAXC.
__device__ __noinline__ void collide(Smem& s, const DevModel& M, int lane) {
  int base = 0;
  for (int p0 = 0; p0 < M.npairs && base < UG_MAXC; p0 += 32) {
    int p = p0 + lane, n = 0, sa = 0, sb = 0;
    CPt pts[4];
    if (p < M.npairs) {
      sa = M.pairs[p][0]; sb = M.pairs[p][1];
      const DevShape& A = M.shapes[sa]; const DevShape& B = M.shapes[sb];
      bool actA = A.body <= 0 || (s.flags & (FL_BODY0 << (A.body - 1)));
      bool actB = B.body <= 0 || (s.flags & (FL_BODY0 << (B.body - 1)));
      if (actA && actB) {
        float margin = fminf(A.cbt, B.cbt);
        BoxW a, b; shape_box(s, A, a); shape_box(s, B, b);
        f3 dc = a.c - b.c;
        bool near;
        if (B.body < 0) {
          f3 e = mk3(fmaxf(fabsf(dc.x) - b.h.x, 0.f), fmaxf(fabsf(dc.y) - b.h.y, 0.f), fmaxf(fabsf(dc.z) - b.h.z, 0.f));
          near = norm3(e) <= A.brad + margin;
        } else near = norm3(dc) <= A.brad + B.brad + margin;
        if (near) {
          if (A.kind == 0 && B.kind == 0) n = box_box(a, b, margin, pts);
          else if (A.kind == 1 && B.kind == 0) {
            f3 ob, nb; float d;
            if (sphere_box(a.c, A.radius, b, margin, ob, nb, d)) { pts[0].n = nb; pts[0].dist = d; pts[0].posB = ob; pts[0].posA = ob + nb * d; n = 1; }
          } else if (A.kind == 0 && B.kind == 1) {
            f3 oa, na; float d;
            if (sphere_box(b.c, B.radius, a, margin, oa, na, d)) { pts[0].n = -na; pts[0].dist = d; pts[0].posA = oa; pts[0].posB = oa + na * d; n = 1; }
          } else {
            float len = norm3(dc), d = len - A.radius - B.radius;
            if (d < margin) {
              f3 nn = len > 1e-12f ? dc * (1.0f / len) : mk3(1, 0, 0);
              pts[0].n = nn; pts[0].dist = d; pts[0].posB = b.c + nn * B.radius; pts[0].posA = pts[0].posB + nn * d; n = 1;
            }
          }
        }
      }
    }
    int incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int t = __shfl_up_sync(UG_FULL, incl, o); if (lane >= o) incl += t; }
    int excl = incl - n, total = __shfl_sync(UG_FULL, incl, 31);
    if (n > 0) {
      const DevShape& A = M.shapes[sa]; const DevShape& B = M.shapes[sb];
      float mu = fminf(fmaxf(A.mu * B.mu, -UGP_MAX_FRICTION), UGP_MAX_FRICTION);
      float roll = fminf(fmaxf(A.roll * B.mu + B.roll * A.mu, -UGP_MAX_FRICTION), UGP_MAX_FRICTION);
      for (int i = 0; i < n; i++) {
        int idx = base + excl + i;
        if (idx < UG_MAXC) {
          Con& c = s.con[idx];
          c.sa = sa; c.sb = sb; st3(c.posA, pts[i].posA); st3(c.posB, pts[i].posB); st3(c.n, pts[i].n);
          c.dist = pts[i].dist; c.mu = mu; c.rest = A.rest * B.rest; c.roll = i == 0 ? roll : 0.f;
        }
      }
    }
    base = min(base + total, UG_MAXC);
  }
  if (lane == 0) s.ncon = base;
  __syncwarp();
}

// Jacobian contribution of one side of a contact-type row. J is a per-lane register array (static indexing only).
UG_DEV void add_side(const Smem& s, const DevModel& M, const DevShape& S, f3 P, f3 dir, bool lin, float sgn, float* J) {
  if (S.body == 0) {
    uint32_t anc = M.anc[S.rbody];
#pragma unroll
    for (int d = 0; d < UG_NDOF; d++) if (anc >> d & 1) {
      f3 a = ld3(s.ax[d]);
      J[d] += sgn * (lin ? dot(dir, cross(a, P - ld3(s.bp[d]))) : dot(dir, a));
    }
  } else if (S.body > 0) {
#pragma unroll
    for (int k = 0; k < UG_MAXFREE; k++) if (S.body - 1 == k) {
      f3 jl = lin ? dir * sgn : mk3(0, 0, 0);
      f3 ja = lin ? cross(P - ld3(s.fpos[k]), dir) * sgn : dir * sgn;
      J[12 + 6 * k] += jl.x; J[13 + 6 * k] += jl.y; J[14 + 6 * k] += jl.z; J[15 + 6 * k] += ja.x; J[16 + 6 * k] += ja.y; J[17 + 6 * k] += ja.z;
    }
  }
}

// B = (M^-1 J^T) for a dense 30-wide row held in registers.
UG_DEV void minv_row(const Smem& s, const DevModel& M, const float* J, float* B) {
#pragma unroll
  for (int a = 0; a < UG_NDOF; a++) {
    float v = 0.f;
#pragma unroll
    for (int b = 0; b < UG_NDOF; b++) v += s.M[a][b] * J[b];
    B[a] = v;
  }
#pragma unroll
  for (int k = 0; k < UG_MAXFREE; k++) {
    float im = k < M.nfree && (s.flags & (FL_BODY0 << k)) ? 1.0f / M.fmass[k] : 0.f;
    B[12 + 6 * k] = J[12 + 6 * k] * im; B[13 + 6 * k] = J[13 + 6 * k] * im; B[14 + 6 * k] = J[14 + 6 * k] * im;
    f3 ba = symv(s.fIinv[k], mk3(J[15 + 6 * k], J[16 + 6 * k], J[17 + 6 * k]));
    B[15 + 6 * k] = ba.x; B[16 + 6 * k] = ba.y; B[17 + 6 * k] = ba.z;
  }
}

struct RowRegs { float rhs, lo, hi, den, mu; int nidx; };

// Builds row `row` (descriptor in s.desc): writes J to shared memory and returns the scalar parameters.
__device__ __noinline__ void build_row(Smem& s, const DevModel& M, int row, int n_nc, RowRegs& rr) {
  float J[30];
#pragma unroll
  for (int t = 0; t < 30; t++) J[t] = 0.f;
  int desc = s.desc[row], type = desc & 15, idx = (desc >> 4) & 255, dir = desc >> 12;
  f3 dirv = mk3(0, 0, 0);
  if (type == RT_LIM_LO || type == RT_LIM_HI || type == RT_MOTOR) {
#pragma unroll
    for (int d = 0; d < UG_NDOF; d++) if (d == idx) J[d] = type == RT_LIM_HI ? -1.0f : 1.0f;
  } else if (type == RT_GEAR) {
    int gd = c_gear_dof[idx]; float gr = c_gear_ratio[idx];
    J[6] = 1.0f;
#pragma unroll
    for (int d = 7; d < UG_NDOF; d++) if (d == gd) J[d] = gr;
  } else if (type != RT_PAD) {
    const Con& c = s.con[idx];
    f3 n = ld3(c.n), t1, t2; plane_space(n, t1, t2);
    dirv = type == RT_NORMAL ? n : (dir == 0 ? t1 : t2);
    bool lin = type != RT_ROLL;
    add_side(s, M, M.shapes[c.sa], ld3(c.posA), dirv, lin, 1.0f, J);
    add_side(s, M, M.shapes[c.sb], ld3(c.posB), dirv, lin, -1.0f, J);
  }
  float B[30]; minv_row(s, M, J, B);
  float den = 0.f, rel = 0.f;
#pragma unroll
  for (int t = 0; t < 30; t++) { den += J[t] * B[t]; rel += J[t] * s.vstar[t]; }
#pragma unroll
  for (int t = 0; t < 30; t++) s.J[row][t] = J[t];
  s.J[row][30] = 0.f; s.J[row][31] = 0.f;
  rr.den = den; rr.mu = 0.f; rr.nidx = 0; rr.lo = 0.f; rr.hi = 0.f; rr.rhs = 0.f;
  float inv = type == RT_PAD ? 0.f : 1.0f / den;
  if (type == RT_PAD) rr.den = 1.0f;
  if (type == RT_LIM_LO || type == RT_LIM_HI) {
    float pen = type == RT_LIM_LO ? s.q[idx] - M.lower[idx] : M.upper[idx] - s.q[idx];
    rr.rhs = (-pen * UGP_ERP * UGP_INV_DT - rel) * inv; rr.lo = 0.f; rr.hi = UGP_LIMIT_MAX_IMPULSE;
  } else if (type == RT_MOTOR) {
    float want = UGP_MOTOR_KP * (s.tgt[idx] - s.q[idx]) * UGP_INV_DT + (1.0f - UGP_MOTOR_KD) * s.vstar[idx];
    want = fminf(fmaxf(want, -M.velocity[idx]), M.velocity[idx]);
    rr.rhs = (want - s.vstar[idx]) * inv; rr.lo = -M.effort[idx] * UGP_DT; rr.hi = M.effort[idx] * UGP_DT;
  } else if (type == RT_GEAR) {
    float pe = UGP_GEAR_ERP * (s.q[6] + c_gear_ratio[idx] * s.q[c_gear_dof[idx]]);
    rr.rhs = (-pe * UGP_ERP * UGP_INV_DT - rel) * inv; rr.lo = -UGP_GEAR_MAX_FORCE * UGP_DT; rr.hi = UGP_GEAR_MAX_FORCE * UGP_DT;
  } else if (type == RT_NORMAL) {
    const Con& c = s.con[idx];
    float pen = c.dist + UGP_LINEAR_SLOP, rest = 0.f;
    if (fabsf(rel) >= UGP_RESTITUTION_VEL_THRESHOLD) rest = fmaxf(c.rest * -rel, 0.f);
    float pos_e = 0.f, vel_e = rest - rel;
    if (pen > 0.f) vel_e -= pen * UGP_INV_DT; else pos_e = -pen * UGP_ERP2 * UGP_INV_DT;
    rr.rhs = (pos_e + vel_e) * inv; rr.lo = 0.f; rr.hi = 1e10f;
  } else if (type == RT_ROLL || type == RT_FRIC) {
    const Con& c = s.con[idx];
    rr.rhs = -rel * inv; rr.mu = type == RT_ROLL ? c.roll : c.mu; rr.nidx = n_nc + idx;
  }
}

// Returns dt * sum |qd| over the 12 joints after the tick (the quantity wait_until_stable thresholds).
__device__ __noinline__ float solve_and_integrate(Smem& s, const DevModel& M, int lane, int* iters_out) {
  // ---- row table
  int n_nc, NC = s.ncon, NT, fr0, R;
  {
    bool lim = false;
    if (lane < 24) { int d = lane >> 1; float pen = (lane & 1) ? M.upper[d] - s.q[d] : s.q[d] - M.lower[d]; lim = !(pen > 0.f); }
    unsigned lm = __ballot_sync(UG_FULL, lim);
    int nl = __popc(lm);
    if (lim) s.desc[__popc(lm & ((1u << lane) - 1))] = ((lane & 1) ? RT_LIM_HI : RT_LIM_LO) | ((lane >> 1) << 4);
    int nm = 0;
    bool arm = s.flags & FL_ARM, grip = s.flags & FL_GRIP;
    if (arm) { if (lane < 6) s.desc[nl + lane] = RT_MOTOR | (lane << 4); nm = 6; }
    if (grip) { if (lane == 0) s.desc[nl + nm] = RT_MOTOR | (6 << 4); nm++; }
    if (lane < 5) s.desc[nl + nm + lane] = RT_GEAR | (lane << 4);
    n_nc = nl + nm + 5;
    if (lane < NC) s.desc[n_nc + lane] = RT_NORMAL | (lane << 4);
    bool rl = lane < NC && s.con[lane].roll > 0.f;
    unsigned rm = __ballot_sync(UG_FULL, rl);
    NT = 2 * __popc(rm);
    if (rl) { int r = __popc(rm & ((1u << lane) - 1)); s.desc[n_nc + NC + 2 * r] = RT_ROLL | (lane << 4); s.desc[n_nc + NC + 2 * r + 1] = RT_ROLL | (lane << 4) | (1 << 12); }
    fr0 = n_nc + NC + NT;
    if (fr0 & 1) { if (lane == 0) s.desc[fr0] = RT_PAD; fr0++; }
    if (lane < NC) { s.desc[fr0 + 2 * lane] = RT_FRIC | (lane << 4); s.desc[fr0 + 2 * lane + 1] = RT_FRIC | (lane << 4) | (1 << 12); }
    R = fr0 + 2 * NC;
  }
  __syncwarp();
  // ---- build rows (three slots per lane)
  RowRegs rr[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    int row = lane + 32 * k;
    rr[k].rhs = 0.f; rr[k].lo = 0.f; rr[k].hi = 0.f; rr[k].den = 1.0f; rr[k].mu = 0.f; rr[k].nidx = 0;
    if (row < R) build_row(s, M, row, n_nc, rr[k]);
  }
  __syncwarp();
  // ---- Delassus matrix A = J M^-1 J^T, column(s) of this lane's rows
#pragma unroll 1
  for (int k = 0; k < 3; k++) {
    int row = lane + 32 * k;
    if (row < R) {
      float J[30], B[30];
#pragma unroll
      for (int t = 0; t < 30; t++) J[t] = s.J[row][t];
      minv_row(s, M, J, B);
      for (int i = 0; i < R; i++) {
        const float4* Ji = reinterpret_cast<const float4*>(s.J[i]);
        float v = 0.f;
#pragma unroll
        for (int t4 = 0; t4 < 7; t4++) { float4 j4 = Ji[t4]; v += j4.x * B[4 * t4] + j4.y * B[4 * t4 + 1] + j4.z * B[4 * t4 + 2] + j4.w * B[4 * t4 + 3]; }
        { float4 j4 = Ji[7]; v += j4.x * B[28] + j4.y * B[29]; }
        s.A[i][row] = v;
      }
    }
  }
  if (lane == 0) s.nrows = R;
  for (int r = lane; r < UG_RMAX; r += 32) s.lam[r] = 0.f;
  __syncwarp();
  // ---- projected Gauss-Seidel in contact space. Lane owns rows lane, lane+32, lane+64: applied impulse lam[k] and
  //      w[k] = J_row . (accumulated delta velocity). Sweep order as Bullet's multibody solver (row P6).
  float lam[3] = {0.f, 0.f, 0.f}, w[3] = {0.f, 0.f, 0.f}, invd[3];
#pragma unroll
  for (int k = 0; k < 3; k++) invd[k] = 1.0f / rr[k].den;
  int it = 0;
  const int nrm1 = n_nc + NC, tr1 = n_nc + NC + NT;
#define UG_ROW_STEP(ROW, K, COND_EXTRA)                                                          \
  {                                                                                              \
    const int l_ = (ROW) & 31;                                                                   \
    float d_ = rr[K].rhs - w[K] * invd[K];                                                       \
    float s_ = fminf(fmaxf(lam[K] + d_, lo_), hi_);                                              \
    d_ = s_ - lam[K];                                                                            \
    if (!(COND_EXTRA)) d_ = 0.f;                                                                 \
    if (lane == l_) { lam[K] += d_; float r_ = d_ * rr[K].den; res = fmaxf(res, r_ * r_); }      \
    float delta = __shfl_sync(UG_FULL, d_, l_);                                                  \
    const float* Ar = s.A[ROW];                                                                  \
    w[0] += Ar[lane] * delta;                                                                    \
    if (R > 32) w[1] += Ar[lane + 32] * delta;                                                   \
    if (R > 64 && lane < UG_RMAX - 64) w[2] += Ar[lane + 64] * delta;                            \
  }
  for (; it < UGP_SOLVER_ITERS;) {
    float res = 0.f;
    // joint rows (all live in slot 0), order alternates with the iteration parity
    for (int j = 0; j < n_nc; j++) {
      int row = (it & 1) ? j : n_nc - 1 - j;
      float lo_ = rr[0].lo, hi_ = rr[0].hi;
      UG_ROW_STEP(row, 0, true)
    }
    // normal rows
#pragma unroll
    for (int k = 0; k < 3; k++) {
      int a = max(n_nc, 32 * k), b = min(nrm1, 32 * k + 32);
      for (int row = a; row < b; row++) {
        float lo_ = 0.f, hi_ = 1e10f;
        UG_ROW_STEP(row, k, true)
        if (lane == (row & 31)) s.lam[row] = lam[k];
      }
    }
    __syncwarp();
    // rolling-friction rows: limits follow the normal impulse; skipped while it is zero
#pragma unroll
    for (int k = 0; k < 3; k++) {
      int a = max(nrm1, 32 * k), b = min(tr1, 32 * k + 32);
      for (int row = a; row < b; row++) {
        float tot = s.lam[rr[k].nidx];
        float hi_ = rr[k].mu * tot, lo_ = -hi_;
        bool on = __shfl_sync(UG_FULL, tot, row & 31) > 0.f;
        UG_ROW_STEP(row, k, on)
      }
    }
    // friction pairs with the implicit cone
#pragma unroll
    for (int k = 0; k < 3; k++) {
      int a = max(fr0, 32 * k), b = min(R, 32 * k + 32);
      for (int row = a; row < b; row += 2) {
        const int l_ = row & 31;
        float lim = rr[k].mu * s.lam[rr[k].nidx];
        float d_ = rr[k].rhs - w[k] * invd[k];
        float sm = lam[k] + d_;
        float sa = __shfl_sync(UG_FULL, sm, l_), sb = __shfl_sync(UG_FULL, sm, l_ + 1);
        float lm = __shfl_sync(UG_FULL, lim, l_);
        float ca = lm, cb = lm;
        if (sa < -lm || sa > lm || sb < -lm || sb > lm) {
          float ih = rsqrtf(fmaxf(sa * sa + sb * sb, 1e-30f));
          ca = fabsf(lm * sa * ih); cb = fabsf(lm * sb * ih);
        }
        float mine = lane == l_ ? ca : cb;
        float sn = fminf(fmaxf(sm, -mine), mine);
        d_ = sn - lam[k];
        if (lane == l_ || lane == l_ + 1) { lam[k] = sn; float r_ = d_ * rr[k].den; res = fmaxf(res, r_ * r_); }
        float da = __shfl_sync(UG_FULL, d_, l_), db = __shfl_sync(UG_FULL, d_, l_ + 1);
        const float* Aa = s.A[row]; const float* Ab = s.A[row + 1];
        w[0] += Aa[lane] * da + Ab[lane] * db;
        if (R > 32) w[1] += Aa[lane + 32] * da + Ab[lane + 32] * db;
        if (R > 64 && lane < UG_RMAX - 64) w[2] += Aa[lane + 64] * da + Ab[lane + 64] * db;
      }
    }
    it++;
    res = warp_max(res);
    if (res <= UGP_RESIDUAL_THRESHOLD) break;
  }
#undef UG_ROW_STEP
  *iters_out = it;
  // ---- generalized impulse g = J^T lambda, delta velocity, integration
#pragma unroll
  for (int k = 0; k < 3; k++) { int row = lane + 32 * k; if (row < R) s.lam[row] = lam[k]; }
  __syncwarp();
  {
    float gt = 0.f;
    for (int r = 0; r < R; r++) gt += s.J[r][lane] * s.lam[r];
    s.g[lane] = gt;
  }
  __syncwarp();
  float absq = 0.f;
  if (lane < UG_NDOF) {
    float dv = 0.f;
#pragma unroll
    for (int b = 0; b < UG_NDOF; b++) dv += s.M[lane][b] * s.g[b];
    float v = fminf(fmaxf(s.vstar[lane] + dv, -UGP_MAX_COORD_VEL), UGP_MAX_COORD_VEL);
    s.qd[lane] = v; s.q[lane] += UGP_DT * v;
    absq = fabsf(UGP_DT * v);
  } else if (lane < UG_NDOF + UG_MAXFREE) {
    int b = lane - UG_NDOF;
    if (b < M.nfree && (s.flags & (FL_BODY0 << b))) {
      const float* gb = s.g + UG_NDOF + 6 * b; const float* vb = s.vstar + UG_NDOF + 6 * b;
      float im = 1.0f / M.fmass[b];
      f3 lin = ld3(vb) + ld3(gb) * im;
      f3 ang = ld3(vb + 3) + symv(s.fIinv[b], ld3(gb + 3));
      st3(s.flin[b], lin); st3(s.fang[b], ang);
      st3(s.fpos[b], ld3(s.fpos[b]) + lin * UGP_DT);
      float wn = norm3(ang);
      if (wn * UGP_DT > UGP_ANGULAR_MOTION_THRESHOLD) wn = UGP_ANGULAR_MOTION_THRESHOLD * UGP_INV_DT;
      f3 axv;
      if (wn < 0.001f) axv = ang * (0.5f * UGP_DT - UGP_DT * UGP_DT * UGP_DT * 0.020833333333f * wn * wn);
      else axv = ang * (sinf(0.5f * wn * UGP_DT) / wn);
      q4 qn = qnormalize(qmul(mkq(axv.x, axv.y, axv.z, cosf(0.5f * wn * UGP_DT)), mkq(s.fquat[b][0], s.fquat[b][1], s.fquat[b][2], s.fquat[b][3])));
      s.fquat[b][0] = qn.x; s.fquat[b][1] = qn.y; s.fquat[b][2] = qn.z; s.fquat[b][3] = qn.w;
    }
  }
  absq = warp_sum(absq);
  __syncwarp();
  return absq;
}

// One stepSimulation tick. Returns sum over joints of |q(t) - q(t-1)|.
UG_DEV float tick(Smem& s, const DevModel& M, int lane, int* iters) {
  UG_TR("tick: fk\n");
  fk_bodies(s, M, lane);
  UG_TR("tick: collide\n");
  collide(s, M, lane);
  UG_TR("tick: dynamics ncon=%d\n", s.ncon);
  dynamics(s, M, lane);
  UG_TR("tick: solve\n");
  float r = solve_and_integrate(s, M, lane, iters);
  UG_TR("tick: done moved=%g iters=%d rows=%d\n", r, *iters, s.nrows);
  return r;
}

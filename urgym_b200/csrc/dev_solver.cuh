// Collision -> solver rows -> projected Gauss-Seidel -> integration, for one env per lane group.
//
// Solver layout. Rows keep the oracle's order: joint rows (active limits by DoF, arm motors, gripper
// motor, five gears), then per sweep normals, rolling-friction rows, friction pairs. Every joint row is a bound on
// the velocity along one of 17 "extended joint columns" (12 unit DoFs + 5 gear combinations), so its response to any
// impulse is one entry of E; contact rows respond to each other through the packed triangle T. Lane l owns joint row
// l (+G ...) and contact rows l, l+G, ...: per row it keeps w = J.dv and the applied impulse in registers, so one
// Gauss-Seidel step is: owner clamps, one shuffle broadcasts the impulse change, every lane does one FMA per row it
// owns with an operand read from shared memory.
#pragma once
#include "dev_engine.cuh"

// Optional per-phase cycle accounting (build with -DUG_PHASE_TIMING; tools/phase_timing.py). Phases: 0 fk, 1 collide,
// 2 dynamics, 3 row setup + pass 1, 4 T build, 5 Gauss-Seidel sweeps, 6 velocity update + integration.
#ifdef UG_PHASE_TIMING
__device__ unsigned long long g_phase_cycles[8];
#define UG_T0() long long t_ph_ = clock64()
#define UG_TP(i) { long long t2_ = clock64(); if ((threadIdx.x & 31) == 0) atomicAdd(&g_phase_cycles[i], (unsigned long long)(t2_ - t_ph_)); t_ph_ = t2_; }
#else
#define UG_T0()
#define UG_TP(i)
#endif

template <class S>
UG_DEV void shape_box(const S& s, const DevShape& Sh, BoxW& b) {
  m3 lR = ldm(Sh.lR); f3 lc = ld3(Sh.lc);
  if (Sh.body == 0) { m3 R = ldm(s.bR[Sh.rbody]); b.R = mulm(R, lR); b.c = ld3(s.bp[Sh.rbody]) + mulv(R, lc); }
  else if (Sh.body > 0) { m3 R = ldm(s.fR[Sh.body - 1]); b.R = mulm(R, lR); b.c = ld3(s.fpos[Sh.body - 1]) + mulv(R, lc); }
  else { b.R = lR; b.c = lc; }
  b.h = ld3(Sh.half);
}

template <class S>
UG_DEV f3 shape_center(const S& s, const DevShape& Sh) {
  const f3 lc = ld3(Sh.lc);
  if (Sh.body == 0) return ld3(s.bp[Sh.rbody]) + mulv(ldm(s.bR[Sh.rbody]), lc);
  if (Sh.body > 0) return ld3(s.fpos[Sh.body - 1]) + mulv(ldm(s.fR[Sh.body - 1]), lc);
  return lc;
}

// Two passes. Pass 1, lane per candidate pair: bounding-sphere cull against the pair's bound M.cullr (sphere against the
// static boxes themselves); survivors are compacted in pair order into a byte list in the (idle) solver scratch. The
// bound is conservative for the narrowphase below: a culled pair could not have produced a manifold point
// (model_build.cpp). Pass 2, lane per surviving pair: narrowphase. Manifold points are appended in pair order up to the
// scene's capacity S::NC_CAP (the checker keeps them all; a tick that had more is counted in s.ovf and reported through
// urgym_stats); the bit of every pair with at least one point is set in s.hit BEFORE the cap, so the contact predicates
// and the pair export never depend on the capacity. A tick with 130 candidate pairs (Golf) and 9 close ones costs one
// narrowphase round.
template <int G, class S>
__device__ __noinline__ void collide(S& s, const DevModel& M, const Gp<G> gp) {
  constexpr int NCAP = S::NC_CAP;
  const int l = gp.l;
  unsigned char* near_list = reinterpret_cast<unsigned char*>(s.u.T);  // UG_MAXPAIRS bytes; T is dead between ticks
  int nnear = 0;
  if (l < UG_HITWORDS) s.hit[l] = 0u;
#pragma unroll 1
  for (int p0 = 0; p0 < M.npairs; p0 += G) {
    const int p = p0 + l;
    bool near = false;
    if (p < M.npairs) {
      const DevShape& A = M.shapes[M.pairs[p][0]]; const DevShape& B = M.shapes[M.pairs[p][1]];
      const bool actA = A.body <= 0 || (s.flags & (FL_BODY0 << (A.body - 1)));
      const bool actB = B.body <= 0 || (s.flags & (FL_BODY0 << (B.body - 1)));
      if (actA && actB) {
        const f3 ca = shape_center(s, A), cb = shape_center(s, B);
        const f3 dc = ca - cb;
        if (B.body < 0) {  // static boxes are axis-aligned in the world (table, plane, pitch)
          const f3 h = ld3(B.half);
          const f3 e = mk3(rmax(rabs(dc.x) - h.x, 0._r), rmax(rabs(dc.y) - h.y, 0._r), rmax(rabs(dc.z) - h.z, 0._r));
          near = norm3(e) <= M.cullr[p];
        } else near = norm3(dc) <= M.cullr[p];
      }
    }
    const unsigned m = gp.ballot(near);
    if (near) near_list[nnear + __popc(m & ((1u << l) - 1u))] = (unsigned char)p;
    nnear += __popc(m);
  }
  gp.sync();
  int base = 0;
  // joint rows of the anchored bodies (Box-v0 only), in the checker's order: per body lower limit, upper limit (when violated), motor
  if constexpr (S::ANCHORS) {
    if (l == 0) {
#pragma unroll 1
      for (int b = 0; b < M.nfree; b++) {
        const int an = M.fanchor[b];
        if (an == 0 || !(s.flags & (FL_BODY0 << b))) continue;
        const real qa = anchor_q(s, M, b);
        const f3 ax = ld3(M.faxis[b]);
#pragma unroll 1
        for (int k = 0; k < 3; k++) {
          real pen = k == 0 ? qa - M.flower[b] : (k == 1 ? M.fupper[b] - qa : qa);
          if (k < 2 ? pen > 0._r : !(M.fmotor[b] > 0._r)) continue;
          typename S::Con& c = s.con[base++];
          c.shapes = UG_PSEUDO | (k < 2 ? 1 : 2) | (an == 1 ? 4 : 0); c.bodies = b + 1;
          st3(c.pa, mk3(0, 0, 0)); st3(c.rb, mk3(0, 0, 0)); st3(c.n, k == 1 ? -ax : ax);
          c.dist = pen; c.mu = 0._r; con_set_rest_roll(c, M.fmotor[b], 0._r);
        }
      }
    }
    base = gp.shfl(base, 0);
  }
  if (l == 0) s.npseudo = base;   // 0 unless S::ANCHORS
#pragma unroll 1
  for (int i0 = 0; i0 < nnear; i0 += G) {
    int n = 0, sa = 0, sb = 0, p = 0;
    CPt pts[4];
    if (i0 + l < nnear) {
      p = near_list[i0 + l];
      sa = M.pairs[p][0]; sb = M.pairs[p][1];
      const DevShape& A = M.shapes[sa]; const DevShape& B = M.shapes[sb];
      const real margin = rmin(A.cbt, B.cbt);
      BoxW a, b; shape_box(s, A, a); shape_box(s, B, b);
      const f3 dc = a.c - b.c;
      if (A.kind != 1 && B.kind == 0) n = box_box(a, b, margin, pts);   // a polytope's bounding box: the mid-phase
      else if (A.kind == 2) { f3 oa, na; real d; n = sphere_box(b.c, B.radius, a, margin, oa, na, d) ? 1 : 0; }
      else if (A.kind == 1 && B.kind == 0) {
        f3 ob, nb; real d;
        if (sphere_box(a.c, A.radius, b, margin, ob, nb, d)) { pts[0].n = nb; pts[0].dist = d; pts[0].posB = ob; pts[0].posA = ob + nb * d; n = 1; }
      } else if (A.kind == 0 && B.kind == 1) {
        f3 oa, na; real d;
        if (sphere_box(b.c, B.radius, a, margin, oa, na, d)) { pts[0].n = -na; pts[0].dist = d; pts[0].posA = oa; pts[0].posB = oa + na * d; n = 1; }
      } else {
        real len = norm3(dc), d = len - A.radius - B.radius;
        if (d < margin) {
          f3 nn = len > 1e-12_r ? dc * (1.0_r / len) : mk3(1, 0, 0);
          pts[0].n = nn; pts[0].dist = d; pts[0].posB = b.c + nn * B.radius; pts[0].posA = pts[0].posB + nn * d; n = 1;
        }
      }
    }
    // robot links with mesh geometry whose bounding box is close: the polytope itself decides. Against a sphere that is
    // cheap and stays in the pair's lane; against a box the pairs are handed to groups of four lanes (poly_box_coop),
    // eight pairs at a time, through scratch in the idle solver block behind the near list.
    const bool is_poly = n > 0 && M.shapes[sa].kind == 2;
    if (is_poly && M.shapes[sb].kind != 0) {
      const DevShape& A = M.shapes[sa]; const DevShape& B = M.shapes[sb];
      const real margin = rmin(A.cbt, B.cbt);
      const int lb = M.link_body[A.link];
      const m3 Rb = ldm(s.bR[lb]);
      const f3 o = ld3(s.bp[lb]) + mulv(Rb, ld3(M.link_lpos[A.link]));
      const m3 Rl = mulm(Rb, ldm(M.link_lrot[A.link]));
      BoxW b; shape_box(s, B, b);
      f3 oa, na; real d; n = 0;
      if (sphere_poly(b.c, B.radius, M.polys[A.poly], o, Rl, margin, oa, na, d)) { pts[0].n = -na; pts[0].dist = d; pts[0].posA = oa; pts[0].posB = oa + na * d; n = 1; }
    }
    if constexpr (G == 32) {
      const bool job = is_poly && M.shapes[sb].kind == 0;
      const unsigned jm = gp.ballot(job);
      const int njobs = __popc(jm), jidx = __popc(jm & ((1u << l) - 1u));
      real* const scratch = s.u.T + 64;           // the near list holds UG_MAXPAIRS bytes at the start of T
      static_assert(64 + UG_PB_GROUPS * UG_PB_SCRATCH <= S::TTRI, "polytope scratch does not fit the idle solver block");
#pragma unroll 1
      for (int j0 = 0; j0 < njobs; j0 += UG_PB_GROUPS) {
        const int g = l >> 2, sub = l & 3, myjob = j0 + g;
        const bool active = myjob < njobs;
        const int owner = active ? (int)__fns(jm, 0, myjob + 1) : 0;
        const int pj = gp.shfl(p, owner);
        real* const sc = scratch + UG_PB_SCRATCH * g;
        int cnt = 0;
        if (active) {
          const DevShape& A = M.shapes[M.pairs[pj][0]]; const DevShape& B = M.shapes[M.pairs[pj][1]];
          const real margin = rmin(A.cbt, B.cbt);
          const int lb = M.link_body[A.link];
          const m3 Rb = ldm(s.bR[lb]);
          const f3 o = ld3(s.bp[lb]) + mulv(Rb, ld3(M.link_lpos[A.link]));
          const m3 Rl = mulm(Rb, ldm(M.link_lrot[A.link]));
          BoxW b; shape_box(s, B, b);
          cnt = poly_box_coop(M.polys[A.poly], o, Rl, b, margin, sc, sub, 0xFu << (l & ~3));
        }
        gp.sync();
        const bool mine = job && jidx >= j0 && jidx < j0 + UG_PB_GROUPS;
        const int got = gp.shfl(cnt, mine ? 4 * (jidx - j0) : 0);
        if (mine) {
          n = got;
          const real* r = scratch + UG_PB_SCRATCH * (jidx - j0);
          for (int i = 0; i < got; i++) { pts[i].posA = ld3(r + 10 * i); pts[i].posB = ld3(r + 10 * i + 3); pts[i].n = ld3(r + 10 * i + 6); pts[i].dist = r[10 * i + 9]; }
        }
        gp.sync();
      }
    } else if (is_poly && M.shapes[sb].kind == 0) {
      const DevShape& A = M.shapes[sa]; const DevShape& B = M.shapes[sb];
      const real margin = rmin(A.cbt, B.cbt);
      const int lb = M.link_body[A.link];
      const m3 Rb = ldm(s.bR[lb]);
      const f3 o = ld3(s.bp[lb]) + mulv(Rb, ld3(M.link_lpos[A.link]));
      const m3 Rl = mulm(Rb, ldm(M.link_lrot[A.link]));
      BoxW b; shape_box(s, B, b);
      n = poly_box(M.polys[A.poly], o, Rl, b, margin, pts);
    }
    int incl = n;
#pragma unroll
    for (int o = 1; o < G; o <<= 1) { int t = __shfl_up_sync(gp.mask, incl, o, G); if (l >= o) incl += t; }
    int excl = incl - n, total = gp.shfl(incl, G - 1);
    if (n > 0) {
      atomicOr(&s.hit[p >> 5], 1u << (p & 31));
      const DevShape& A = M.shapes[sa]; const DevShape& B = M.shapes[sb];
      real mu = rmin(rmax(A.mu * B.mu, -UGP_MAX_FRICTION), UGP_MAX_FRICTION);
      real roll = rmin(rmax(A.roll * B.mu + B.roll * A.mu, -UGP_MAX_FRICTION), UGP_MAX_FRICTION);
      const int fa = A.body > 0 ? A.body - 1 : -1, fb = B.body > 0 ? B.body - 1 : -1, ra = A.body == 0 ? A.rbody : -1;
#pragma unroll 1
      for (int i = 0; i < n; i++) {
        int idx = base + excl + i;
        if (idx < NCAP) {
          typename S::Con& c = s.con[idx];
          c.shapes = sa | (sb << 8); c.bodies = (fa + 1) | ((fb + 1) << 8) | ((ra + 1) << 16);
          st3(c.pa, fa >= 0 ? pts[i].posA - ld3(s.fpos[fa]) : pts[i].posA);
          st3(c.rb, fb >= 0 ? pts[i].posB - ld3(s.fpos[fb]) : pts[i].posB);
          st3(c.n, pts[i].n);
          c.dist = pts[i].dist; c.mu = mu; con_set_rest_roll(c, A.rest * B.rest, i == 0 ? roll : 0._r);
        }
      }
    }
    base += total;
  }
  if (l == 0) { s.ncon = min(base, NCAP); if (base > NCAP) s.ovf++; }
  gp.sync();
}

// Jacobian of one contact-type row: 12 robot entries plus the linear / angular blocks of the (at most two) free
// bodies it touches. Register-resident (static indexing only).
struct CRow {
  real Jr[UG_NDOF];
  f3 JlA, JaA, JlB, JaB;
  int fa, fb;
};

template <class S>
UG_DEV void build_crow(const S& s, const DevModel& M, int desc, CRow& r, bool robot_part) {
  const int k = desc & 255, type = (desc >> 8) & 15, dir = desc >> 12;
  const typename S::Con& c = s.con[k];
  f3 n = ld3(c.n), t1, t2; plane_space(n, t1, t2);
  f3 dv = type == RT_NORMAL ? n : (dir == 0 ? t1 : t2);
  const bool lin = type != RT_ROLL && (!S::ANCHORS || (c.shapes & (UG_PSEUDO | 4)) != (UG_PSEUDO | 4));
  const int bod = c.bodies;
  r.fa = (bod & 255) - 1; r.fb = ((bod >> 8) & 255) - 1;
  const int ra = ((bod >> 16) & 255) - 1;
  f3 pa = ld3(c.pa), rb = ld3(c.rb);
  const f3 z = mk3(0, 0, 0);
  r.JlA = z; r.JaA = z; r.JlB = z; r.JaB = z;
  if (r.fa >= 0) { r.JlA = lin ? dv : z; r.JaA = lin ? cross(pa, dv) : dv; }
  if (r.fb >= 0) { r.JlB = lin ? -dv : z; r.JaB = lin ? -cross(rb, dv) : -dv; }
  if (robot_part) {
    const uint32_t anc = ra >= 0 ? M.anc[ra] : 0u;
#pragma unroll
    for (int d = 0; d < UG_NDOF; d++) {
      real v = 0._r;
      if (anc >> d & 1) {
        f3 a = ld3(s.ax[d]);
        v = lin ? dot(dv, cross(a, pa - ld3(s.bp[d]))) : dot(dv, a);
      }
      r.Jr[d] = v;
    }
  }
}

UG_DEV f3 pick3(const f3* a, int b) { return b == 0 ? a[0] : (b == 1 ? a[1] : a[2]); }

// All PGS sweeps of a contact-free tick with exactly N joint rows, warp per env: lane l owns row l and keeps the
// couplings of its row to every row of the sweep in registers, so a Gauss-Seidel step is FFMA, 2 x FMNMX, SHFL, FFMA
// plus the owner's impulse update: no operand loads, no address arithmetic, no loop counter, no guards. The operations
// and their order are those of the generic sweep (same bits). Returns the number of sweeps done; lam0 = the impulse.
template <int N, class S>
UG_DEV int joint_sweeps_reg(const S& s, const Gp<32>& gp, unsigned aE_own, real rhs0, real invd0, real lo0, real hi0, real den0, real& lam0) {
  const int l = gp.l;
  real e[N];
#pragma unroll
  for (int r = 0; r < N; r++) e[r] = lds(aE_own + (unsigned)s.jci4[r + 1]);
  real w0 = 0._r;
  int it = 0;
#define UG_JR_STEP(R)                                                                                    \
  {                                                                                                      \
    real d_ = rfma(-w0, invd0, rhs0);                                                                   \
    d_ = rmin(rmax(d_, blo), bhi);                                                                     \
    const real delta = gp.shfl(d_, (R));                                                                \
    if (l == (R)) lam0 += delta;                                                                         \
    w0 = rfma(e[(R)], delta, w0);                                                                        \
  }
#pragma unroll 1
  for (; it < UGP_SOLVER_ITERS;) {
    // a lane's impulse changes only at its own step, so its clamp bounds are constant over the sweep
    const real blo = lo0 - lam0, bhi = hi0 - lam0, lamp0 = lam0;
    if (it & 1) {
#pragma unroll
      for (int r = 0; r < N; r++) UG_JR_STEP(r)
    } else {
#pragma unroll
      for (int r = N - 1; r >= 0; r--) UG_JR_STEP(r)
    }
    it++;
    const real r_ = (lam0 - lamp0) * den0;
    if (!gp.any(r_ * r_ > UGP_RESIDUAL_THRESHOLD)) break;
  }
#undef UG_JR_STEP
  return it;
}

// Returns dt * sum |qd| over the 12 joints after the tick (the quantity wait_until_stable thresholds).
// The sweep loops are kept rolled and small on purpose: one Gauss-Seidel sweep has to stay resident in the ~6 KB L0
// instruction cache, because it runs 50 times per tick and is ~80 % of the kernel's time (profiles/).
template <int G, int RC, class S>
__device__ __noinline__ real solve_and_integrate(S& s, const DevModel& M, const Gp<G> gp, int* iters_out) {
  constexpr int NSJ = (UG_JMAX + G - 1) / G, NSC = (RC + G - 1) / G, NS = NSJ + NSC;
  const int l = gp.l;
  const int NC = s.ncon;
  UG_T0();
  // ---- row tables
  int n_nc, NT, fr0, Rc;
  {
    int lim = 0;
    if (l < UG_NDOF) {
      if (!(s.q[l] - M.lower[l] > 0._r)) lim = 1;
      else if (!(M.upper[l] - s.q[l] > 0._r)) lim = 2;
    }
    unsigned lm = gp.ballot(lim != 0);
    int nl = __popc(lm);
    if (lim) { const int p_ = __popc(lm & ((1u << l) - 1u)); s.jdesc[p_] = (unsigned short)(l | ((lim == 1 ? RT_LIM_LO : RT_LIM_HI) << 8) | (l << 12)); s.jci4[p_ + 1] = (unsigned short)(UG_RB * l); }
    int nm = 0;
    const int fl = s.flags;
    if (fl & FL_ARM) { if (l < 6) { s.jdesc[nl + l] = (unsigned short)(l | (RT_MOTOR << 8) | (l << 12)); s.jci4[nl + l + 1] = (unsigned short)(UG_RB * l); } nm = 6; }
    if (fl & FL_GRIP) { if (l == 0) { s.jdesc[nl + nm] = (unsigned short)(6 | (RT_MOTOR << 8) | (6 << 12)); s.jci4[nl + nm + 1] = (unsigned short)(UG_RB * 6); } nm++; }
    if (l < 5) { s.jdesc[nl + nm + l] = (unsigned short)((12 + l) | (RT_GEAR << 8) | (l << 12)); s.jci4[nl + nm + l + 1] = (unsigned short)(UG_RB * (12 + l)); }
    n_nc = nl + nm + 5;
    if (l == 0) { s.jci4[0] = 0; s.jci4[n_nc + 1] = 0; }  // slack entries: the sweeps prefetch one row ahead
    if (l < NC) s.cdesc[l] = l | (RT_NORMAL << 8);
    bool rl = l < NC && con_roll(s.con[l]) > 0._r;
    unsigned rm = gp.ballot(rl);
    constexpr int NROLL = (RC - 3 * S::NC_CAP) / 2;  // rolling-friction manifolds a tick can carry (Golf scenes: 4, else 0)
    NT = 2 * min(__popc(rm), NROLL);
    if (l == 0 && __popc(rm) > NROLL) s.ovf++;
    if (rl) {
      int r = __popc(rm & ((1u << l) - 1u));
      if (r < NROLL) { s.cdesc[NC + 2 * r] = l | (RT_ROLL << 8); s.cdesc[NC + 2 * r + 1] = l | (RT_ROLL << 8) | (1 << 12); }
    }
    fr0 = NC + NT;
    if (fr0 & 1) { if (l == 0) s.cdesc[fr0] = RT_PAD << 8; fr0++; }
    if (l < NC) { s.cdesc[fr0 + 2 * l] = l | (RT_FRIC << 8); s.cdesc[fr0 + 2 * l + 1] = l | (RT_FRIC << 8) | (1 << 12); }
    Rc = NC > 0 ? fr0 + 2 * NC : 0;
    // Rc <= NC_CAP + 2 * NROLL + 2 * NC_CAP = RC
  }
  gp.sync();
  // ---- per-lane row registers: slots [0, NSJ) joint rows, [NSJ, NS) contact rows
  real w[NS], lam[NS], lamp[NS], rhs[NS], invd[NS], den[NS], lo[NS], hi[NS];
  int colown[NSJ], nidx[NSC];
#pragma unroll
  for (int k = 0; k < NS; k++) { w[k] = 0._r; lam[k] = 0._r; lamp[k] = 0._r; rhs[k] = 0._r; invd[k] = 0._r; den[k] = 0._r; lo[k] = 0._r; hi[k] = 0._r; }
  // joint rows (signed form: a hi-limit row is the unit row with bounds [-max, 0])
#pragma unroll
  for (int sj = 0; sj < NSJ; sj++) {
    const int j = l + G * sj;
    colown[sj] = 0;
    if (j < n_nc) {
      const int dsc = s.jdesc[j], col = dsc & 255, type = (dsc >> 8) & 15, idx = dsc >> 12;
      colown[sj] = col;
      const real dn = s.E[col * UG_NCOL + col], inv = 1.0_r / dn;
      den[sj] = dn; invd[sj] = inv;
      const real vs = s.vstar[type == RT_GEAR ? 6 : idx], qi = s.q[type == RT_GEAR ? 6 : idx];
      if (type == RT_LIM_LO) {
        rhs[sj] = (-(qi - M.lower[idx]) * UGP_ERP * UGP_INV_DT - vs) * inv; lo[sj] = 0._r; hi[sj] = UGP_LIMIT_MAX_IMPULSE;
      } else if (type == RT_LIM_HI) {
        rhs[sj] = ((M.upper[idx] - qi) * UGP_ERP * UGP_INV_DT - vs) * inv; lo[sj] = -UGP_LIMIT_MAX_IMPULSE; hi[sj] = 0._r;
      } else if (type == RT_MOTOR) {
        real want = UGP_MOTOR_KP * (s.tgt[idx] - qi) * UGP_INV_DT + (1.0_r - UGP_MOTOR_KD) * vs;
        want = rmin(rmax(want, -M.velocity[idx]), M.velocity[idx]);
        rhs[sj] = (want - vs) * inv; lo[sj] = -M.effort[idx] * UGP_DT; hi[sj] = M.effort[idx] * UGP_DT;
      } else {
        const int gd = c_gear_dof[idx]; const real gr = c_gear_ratio[idx];
        real rel = vs + gr * s.vstar[gd];
        real pe = UGP_GEAR_ERP * (qi + gr * s.q[gd]);
        rhs[sj] = (-pe * UGP_ERP * UGP_INV_DT - rel) * inv; lo[sj] = -UGP_GEAR_MAX_FORCE * UGP_DT; hi[sj] = UGP_GEAR_MAX_FORCE * UGP_DT;
      }
    }
  }
  // contact rows, pass 1 (one code copy, slot loop rolled): Jacobian, robot response (-> E), diagonal, right-hand
  // side. Row parameters go through a scratch area behind the M^-1 copy (dead dynamics temporaries).
  real* const prm = &s.u.d.cm[0][0];
#pragma unroll 1
  for (int so = 0; so * G < Rc; so++) {
    const int c = l + G * so;
    if (c < Rc) {
      const int dsc = s.cdesc[c];
      const int type = (dsc >> 8) & 15, k = dsc & 255;
      real4 p0 = make_real4(0._r, 0._r, 0._r, 0._r), p1 = make_real4(0._r, 0._r, 0._r, 0._r);
      real* Er = s.E + (UG_NCOL + c) * UG_NCOL;
      if (type != RT_PAD) {
        CRow r; build_crow(s, M, dsc, r, true);
        real Br[UG_NDOF];
#pragma unroll
        for (int a = 0; a < UG_NDOF; a++) {
          const real4* Mr = reinterpret_cast<const real4*>(s.u.d.M[a]);
          real4 m0 = Mr[0], m1 = Mr[1], m2 = Mr[2];
          Br[a] = m0.x * r.Jr[0] + m0.y * r.Jr[1] + m0.z * r.Jr[2] + m0.w * r.Jr[3] + m1.x * r.Jr[4] + m1.y * r.Jr[5] + m1.z * r.Jr[6] + m1.w * r.Jr[7] +
                  m2.x * r.Jr[8] + m2.y * r.Jr[9] + m2.z * r.Jr[10] + m2.w * r.Jr[11];
        }
        real dn = 0._r, rel = 0._r;
#pragma unroll
        for (int a = 0; a < UG_NDOF; a++) { Er[a] = Br[a]; dn += r.Jr[a] * Br[a]; rel += r.Jr[a] * s.vstar[a]; }
#pragma unroll
        for (int gi = 0; gi < 5; gi++) Er[12 + gi] = Br[6] + c_gear_ratio[gi] * Br[7 + gi];
        if (r.fa >= 0) {
          const real* v = s.vstar + UG_NDOF + 6 * r.fa;
          if (!S::ANCHORS || M.fanchor[r.fa] == 0) dn += dot(r.JlA, r.JlA) * (1.0_r / M.fmass[r.fa]) + dot(r.JaA, symv(s.fIinv[r.fa], r.JaA));
          else dn += dot(r.JlA, lin_inv(M, r.fa, r.JlA)) + dot(r.JaA, symv(s.fIinv[r.fa], r.JaA));
          rel += dot(r.JlA, ld3(v)) + dot(r.JaA, ld3(v + 3));
        }
        if (r.fb >= 0) {
          const real* v = s.vstar + UG_NDOF + 6 * r.fb;
          if (!S::ANCHORS || M.fanchor[r.fb] == 0) dn += dot(r.JlB, r.JlB) * (1.0_r / M.fmass[r.fb]) + dot(r.JaB, symv(s.fIinv[r.fb], r.JaB));
          else dn += dot(r.JlB, lin_inv(M, r.fb, r.JlB)) + dot(r.JaB, symv(s.fIinv[r.fb], r.JaB));
          rel += dot(r.JlB, ld3(v)) + dot(r.JaB, ld3(v + 3));
        }
        real inv = 1.0_r / dn;
        if (S::ANCHORS && !(dn > 1e-20_r)) inv = 0._r;   // a row no body can move along (a tangent across a hinge axis) stays inert
        p0.y = inv; p0.z = dn;
        const typename S::Con& cn = s.con[k];
        if (S::ANCHORS && type == RT_NORMAL && (cn.shapes & UG_PSEUDO)) {
          if (cn.shapes & 1) {   // joint limit of an anchored body: the robot's limit row formula
            p0.x = (-cn.dist * UGP_ERP * UGP_INV_DT - rel) * inv; p0.w = 0._r; p1.x = UGP_LIMIT_MAX_IMPULSE;
          } else {               // position motor to coordinate 0 (cn.dist = coordinate, cn.rest = force limit)
            const real want = UGP_MOTOR_KP * (0._r - cn.dist) * UGP_INV_DT + (1.0_r - UGP_MOTOR_KD) * rel;
            p0.x = (want - rel) * inv; p0.w = -con_rest(cn) * UGP_DT; p1.x = con_rest(cn) * UGP_DT;
          }
        } else if (type == RT_NORMAL) {
          real pen = cn.dist + UGP_LINEAR_SLOP, rest = 0._r;
          if (rabs(rel) >= UGP_RESTITUTION_VEL_THRESHOLD) rest = rmax(con_rest(cn) * -rel, 0._r);
          real pos_e = 0._r, vel_e = rest - rel;
          if (pen > 0._r) vel_e -= pen * UGP_INV_DT; else pos_e = -pen * UGP_ERP2 * UGP_INV_DT;
          p0.x = (pos_e + vel_e) * inv; p0.w = 0._r; p1.x = 1e10_r;
        } else {
          p0.x = -rel * inv; p1.x = type == RT_ROLL ? con_roll(cn) : cn.mu; p1.y = (real)k;
        }
      } else {
#pragma unroll
        for (int a = 0; a < UG_NCOL; a++) Er[a] = 0._r;
      }
      reinterpret_cast<real4*>(prm)[2 * c] = p0; reinterpret_cast<real4*>(prm)[2 * c + 1] = p1;
    }
  }
#pragma unroll
  for (int so = 0; so < NSC; so++) {
    const int c = l + G * so;
    nidx[so] = 0;
    if (c < Rc) {
      real4 p0 = reinterpret_cast<const real4*>(prm)[2 * c], p1 = reinterpret_cast<const real4*>(prm)[2 * c + 1];
      rhs[NSJ + so] = p0.x; invd[NSJ + so] = p0.y; den[NSJ + so] = p0.z; lo[NSJ + so] = p0.w; hi[NSJ + so] = p1.x; nidx[so] = (int)p1.y;
    }
  }
  gp.sync();
  UG_TP(3)
  // contact rows, pass 2: T(c, c') = J_c' . (M^-1 J_c^T) for c' <= c. The owner of row c keeps its Jacobian and its
  // free-body responses in registers; the loop over c' is group-uniform.
#pragma unroll 1
  for (int so = 0; so * G < Rc; so++) {
    const int c = l + G * so;
    const bool has = c < Rc;
    const int dsc = has ? s.cdesc[c] : (RT_PAD << 8);
    CRow r;
    f3 dl[UG_MAXFREE], da[UG_MAXFREE];
#pragma unroll
    for (int b = 0; b < UG_MAXFREE; b++) { dl[b] = mk3(0, 0, 0); da[b] = mk3(0, 0, 0); }
#pragma unroll
    for (int d = 0; d < UG_NDOF; d++) r.Jr[d] = 0._r;
    if (((dsc >> 8) & 15) != RT_PAD) {
      build_crow(s, M, dsc, r, true);
#pragma unroll
      for (int b = 0; b < UG_MAXFREE; b++) {
        if (r.fa == b) { dl[b] = (!S::ANCHORS || M.fanchor[b] == 0) ? r.JlA * (1.0_r / M.fmass[b]) : lin_inv(M, b, r.JlA); da[b] = symv(s.fIinv[b], r.JaA); }
        if (r.fb == b) { dl[b] = (!S::ANCHORS || M.fanchor[b] == 0) ? r.JlB * (1.0_r / M.fmass[b]) : lin_inv(M, b, r.JlB); da[b] = symv(s.fIinv[b], r.JaB); }
      }
    }
    const int cmax = min(Rc, so * G + G);
    real* Trow = s.u.T + (c * (c + 1)) / 2;   // packed lower triangle: this lane writes T(c, cp) for cp <= c
#pragma unroll 1
    for (int cp = 0; cp < cmax; cp++) {
      const int dp = s.cdesc[cp], tp = (dp >> 8) & 15;
      real acc = 0._r;
      if (tp != RT_PAD) {
        const typename S::Con& cn = s.con[dp & 255];
        f3 n = ld3(cn.n), t1, t2; plane_space(n, t1, t2);
        f3 dv = tp == RT_NORMAL ? n : ((dp >> 12) == 0 ? t1 : t2);
        const bool angrow = tp == RT_ROLL || (S::ANCHORS && (cn.shapes & (UG_PSEUDO | 4)) == (UG_PSEUDO | 4));
        const real* Ep = s.E + (UG_NCOL + cp) * UG_NCOL;
#pragma unroll
        for (int d = 0; d < UG_NDOF; d++) acc += r.Jr[d] * Ep[d];
        const int bod = cn.bodies, fa = (bod & 255) - 1, fb = ((bod >> 8) & 255) - 1;
        if (fa >= 0) {
          f3 a = pick3(da, fa);
          f3 vel = angrow ? a : pick3(dl, fa) + cross(a, ld3(cn.pa));
          acc += dot(dv, vel);
        }
        if (fb >= 0) {
          f3 a = pick3(da, fb);
          f3 vel = angrow ? a : pick3(dl, fb) + cross(a, ld3(cn.rb));
          acc -= dot(dv, vel);
        }
      }
      if (has && cp <= c) Trow[cp] = acc;   // T(c, cp) with cp > c is the owner of row cp's entry
    }
  }
  gp.sync();
  UG_TP(4)
  // ---- projected Gauss-Seidel, Bullet's sweep order (row P6). Operands are addressed with 32-bit shared addresses:
  //      a step of row r reads E / T at (uniform row offset) + (per-lane constant) + (per-slot immediate).
  const unsigned aE = smem_u32(s.E), aT = smem_u32(s.u.T);
  unsigned aEj[NSJ], aEq[NSJ];
#pragma unroll
  for (int k = 0; k < NSJ; k++) { aEj[k] = aE + UG_RB * (unsigned)(colown[k] * UG_NCOL); aEq[k] = aE + UG_RB * (unsigned)(UG_NCOL * UG_NCOL + colown[k]); }
  const unsigned aEc = aE + UG_RB * (unsigned)((UG_NCOL + l) * UG_NCOL);  // contact slot o: + o * G * UG_NCOL * 4
  // T(r, c) for the lane's row c = l + G o of contact slot o: c <= r -> aTrow[o] + tri(r), c > r -> aTcol[o] + r (in scalars).
  // Lanes beyond the last row address the last row (their operand is never used).
  unsigned aTrow[NSC], aTcol[NSC]; int cown[NSC];
#pragma unroll
  for (int o = 0; o < NSC; o++) { const int c_ = min(l + G * o, RC - 1); cown[o] = c_; aTrow[o] = aT + UG_RB * (unsigned)c_; aTcol[o] = aT + UG_RB * (unsigned)((c_ * (c_ + 1)) / 2); }
  real limv[NSC];
#pragma unroll
  for (int so = 0; so < NSC; so++) limv[so] = 0._r;
  // Operand loads are written BEFORE the clamp / shuffle chain of each step on purpose: every shuffle ends a basic
  // block (convergence check), ptxas does not move loads across it, and a load issued after the shuffle would put
  // its ~29-cycle latency on the step's critical path. Issued first, it overlaps the chain (profiles/ r01d -> r01e).
  //
  // One Gauss-Seidel step of a joint row: ROW is group-uniform, SJ its (unrolled) slot, NACT the number of contact
  // slots in use this tick (sweeps are instantiated per NACT so that the loop bodies carry no dead predicates).
  // CI is the row's column byte offset, loaded one iteration ahead; NEXT is the following row of the sweep.
#define UG_JOINT_STEP(ROW, NEXT, SJ, NACT)                                                               \
  {                                                                                                      \
    const unsigned ci_ = ci_cur;                                                                         \
    real ej_[NSJ], ec_[NACT > 0 ? NACT : 1];                                                            \
    _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) ej_[k_] = lds(aEj[k_] + ci_);                     \
    _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++) ec_[o_] = lds(aEc + ci_ + UG_RB * (unsigned)(o_ * G * UG_NCOL)); \
    ci_cur = (unsigned)s.jci4[(NEXT) + 1];                                                               \
    const int li_ = (ROW) - (SJ) * G;                                                                    \
    real d_ = rfma(-w[SJ], invd[SJ], rhs[SJ]);                                                          \
    d_ = rmin(rmax(d_, jlo[SJ]), jhi[SJ]);                                                             \
    const real delta = gp.shfl(d_, li_);                                                                \
    if (l == li_) lam[SJ] += delta;                                                                      \
    _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) w[k_] = rfma(ej_[k_], delta, w[k_]);              \
    _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++) w[NSJ + o_] = rfma(ec_[o_], delta, w[NSJ + o_]); \
  }
  // Operands of a contact-row step: E entries for the joint slots, T entries for the contact slots. Each operand has ONE
  // running shared-memory address per lane: the E entry of the lane's column advances by a row of E per stepping row; the T
  // entry of the lane's own row c against stepping row r sits at tri(r) + c for c <= r and at tri(c) + r for c > r, so going
  // from r to r + 1 it advances by one scalar while r < c and by r + 1 scalars from there on (the crossing r = c - 1 -> c is one
  // scalar too). Slots below the stepping slot SO hold only rows c < r and slots above it only rows c > r, so only the
  // stepping slot needs the compare. (Round 1-2a recomputed tri(r), two candidate addresses and a select per operand per
  // step: 13-18 integer instructions of the 27 / 47 of a normal / friction step; profiles/r02_ncu_full_*.)
  //   UG_TP_INIT: addresses for stepping row R0. UG_ROW_LOAD / UG_PAIR_LOAD: load the operands of row ROW (and ROW + 1) and
  //   advance to ROW + 1 (ROW + 2); RB1 = UG_RB * (ROW + 1).
#define UG_TP_INIT(SO, R0, NACT)                                                                         \
    {                                                                                                    \
      const unsigned tri0_ = UG_RB * (unsigned)(((R0) * ((R0) + 1)) >> 1), rb0_ = UG_RB * (unsigned)(R0); \
      _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) pE[k_] = aEq[k_] + (UG_RB * UG_NCOL) * (unsigned)(R0); \
      _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++)                                              \
        pT[o_] = (o_) < (SO) ? aTrow[o_] + tri0_ : ((o_) > (SO) ? aTcol[o_] + rb0_ : (cown[o_] <= (R0) ? aTrow[o_] + tri0_ : aTcol[o_] + rb0_)); \
    }
#define UG_ROW_LOAD(SO, ROW, RB1, NACT)                                                                  \
    real uj_[NSJ], uc_[NACT];                                                                           \
    _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) { uj_[k_] = lds(pE[k_]); pE[k_] += UG_RB * UG_NCOL; } \
    _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++) {                                              \
      uc_[o_] = lds(pT[o_]);                                                                             \
      pT[o_] += (o_) < (SO) ? (RB1) : ((o_) > (SO) ? UG_RB : ((ROW) < cown[o_] ? UG_RB : (RB1)));        \
    }
#define UG_PAIR_LOAD(SO, ROW, RB1, NACT)                                                                 \
    real uj_[NSJ], vj_[NSJ], uc_[NACT], vc_[NACT];                                                      \
    _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) { uj_[k_] = lds(pE[k_]); vj_[k_] = lds(pE[k_] + UG_RB * UG_NCOL); pE[k_] += 2u * UG_RB * UG_NCOL; } \
    _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++) {                                              \
      uc_[o_] = lds(pT[o_]);                                                                             \
      if ((o_) > (SO)) { vc_[o_] = lds(pT[o_] + UG_RB); pT[o_] += 2u * UG_RB; }                          \
      else {                                                                                             \
        const unsigned q_ = pT[o_] + ((o_) < (SO) ? (RB1) : ((ROW) < cown[o_] ? UG_RB : (RB1)));         \
        vc_[o_] = lds(q_);                                                                               \
        pT[o_] = q_ + ((o_) < (SO) ? (RB1) + UG_RB : ((ROW) + 1 < cown[o_] ? UG_RB : (RB1) + UG_RB));    \
      }                                                                                                  \
    }
#define UG_ROW_APPLY(DA, NACT)                                                                           \
    _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) w[k_] = rfma(uj_[k_], DA, w[k_]);                 \
    _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++) w[NSJ + o_] = rfma(uc_[o_], DA, w[NSJ + o_]);
#define UG_PAIR_APPLY(DA, DB, NACT)                                                                      \
    _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) { w[k_] = rfma(uj_[k_], DA, w[k_]); w[k_] = rfma(vj_[k_], DB, w[k_]); } \
    _Pragma("unroll") for (int o_ = 0; o_ < (NACT); o_++) { w[NSJ + o_] = rfma(uc_[o_], DA, w[NSJ + o_]); w[NSJ + o_] = rfma(vc_[o_], DB, w[NSJ + o_]); }
  // joint rows, order alternates with the iteration parity (backward on even sweeps)
#define UG_JOINT_SWEEP(NACT)                                                                             \
  /* a lane's impulse changes only at its own step, so its clamp bounds are constant over the sweep */  \
  real jlo[NSJ], jhi[NSJ];                                                                              \
  _Pragma("unroll") for (int k_ = 0; k_ < NSJ; k_++) { jlo[k_] = lo[k_] - lam[k_]; jhi[k_] = hi[k_] - lam[k_]; } \
  if (it & 1) {                                                                                          \
    unsigned ci_cur = (unsigned)s.jci4[1];                                                               \
    _Pragma("unroll") for (int sj = 0; sj < NSJ; sj++) {                                                 \
      const int b_ = min(n_nc, sj * G + G);                                                              \
      int row = sj * G;                                     /* two rows per trip: half the loop overhead */ \
      _Pragma("unroll 1") for (; row + 1 < b_; row += 2) { UG_JOINT_STEP(row, row + 1, sj, NACT) UG_JOINT_STEP(row + 1, row + 2, sj, NACT) } \
      if (row < b_) UG_JOINT_STEP(row, row + 1, sj, NACT)                                                \
    }                                                                                                    \
  } else {                                                                                               \
    unsigned ci_cur = (unsigned)s.jci4[n_nc];                                                            \
    _Pragma("unroll") for (int sj = NSJ - 1; sj >= 0; sj--) {                                            \
      int row = min(n_nc, sj * G + G) - 1;                                                               \
      _Pragma("unroll 1") for (; row - 1 >= sj * G; row -= 2) { UG_JOINT_STEP(row, row - 1, sj, NACT) UG_JOINT_STEP(row - 1, row - 2, sj, NACT) } \
      if (row >= sj * G) UG_JOINT_STEP(row, row - 1, sj, NACT)                                           \
    }                                                                                                    \
  }
  // normals (contacts 0..NC-1 live in contact slot 0: NC <= NC_CAP <= G), friction limits from this sweep's normal
  // impulses, rolling-friction rows (Golf only; skipped while the normal impulse is zero), friction pairs with the
  // implicit cone
#define UG_CONTACT_SWEEP(NACT)                                                                           \
  {                                                                                                      \
    unsigned pE[NSJ], pT[NACT];                                                                          \
    const real nlo = lo[NSJ] - lam[NSJ], nhi = hi[NSJ] - lam[NSJ];   /* constant over the sweep, as above */ \
    UG_TP_INIT(0, 0, NACT)                                                                               \
    {                                                                                                    \
      unsigned rb1 = UG_RB;                                                                              \
      _Pragma("unroll 1") for (int row = 0; row < NC; row++, rb1 += UG_RB) {                             \
        UG_ROW_LOAD(0, row, rb1, NACT)                                                                   \
        real d_ = rfma(-w[NSJ], invd[NSJ], rhs[NSJ]);                                                   \
        d_ = rmin(rmax(d_, nlo), nhi);                                                                 \
        const real delta = gp.shfl(d_, row);                                                            \
        if (l == row) lam[NSJ] += delta;                                                                 \
        UG_ROW_APPLY(delta, NACT)                                                        \
      }                                                                                                  \
    }                                                                                                    \
    _Pragma("unroll") for (int so = 0; so < (NACT); so++) limv[so] = hi[NSJ + so] * gp.shfl(lam[NSJ], nidx[so]); \
    if (NT > 0) {                                                                                        \
      _Pragma("unroll") for (int so = 0; so < (NACT); so++) {                                            \
        const int a_ = max(NC, so * G), b_ = min(NC + NT, so * G + G);                                   \
        if (a_ < b_) {                                                                                   \
          UG_TP_INIT(so, a_, NACT)                                                                       \
          unsigned rb1 = UG_RB * (unsigned)(a_ + 1);                                                     \
          _Pragma("unroll 1") for (int row = a_; row < b_; row++, rb1 += UG_RB) {                        \
            UG_ROW_LOAD(so, row, rb1, NACT)                                                              \
            const real bnd = limv[so];                                                                  \
            real d_ = rfma(-w[NSJ + so], invd[NSJ + so], rhs[NSJ + so]);                                \
            d_ = rmin(rmax(d_, -bnd - lam[NSJ + so]), bnd - lam[NSJ + so]);                            \
            if (!(bnd > 0._r)) d_ = 0._r;                                                                  \
            const real delta = gp.shfl(d_, row - so * G);                                               \
            if (l == row - so * G) lam[NSJ + so] += delta;                                               \
            UG_ROW_APPLY(delta, NACT)                                                    \
          }                                                                                              \
        }                                                                                                \
      }                                                                                                  \
    }                                                                                                    \
    _Pragma("unroll") for (int so = 0; so < (NACT); so++) {                                              \
      const int a_ = max(fr0, so * G), b_ = min(Rc, so * G + G);                                         \
      if (a_ < b_) {                                                                                     \
        UG_TP_INIT(so, a_, NACT)                                                                         \
        unsigned rb1 = UG_RB * (unsigned)(a_ + 1);                                                       \
        _Pragma("unroll 1") for (int row = a_; row < b_; row += 2, rb1 += 2u * UG_RB) {                  \
          UG_PAIR_LOAD(so, row, rb1, NACT)                                                               \
          const int li = row - so * G;                                                                   \
          real d_ = rfma(-w[NSJ + so], invd[NSJ + so], rhs[NSJ + so]);                                  \
          const real sm = lam[NSJ + so] + d_;                                                           \
          const real ot = __shfl_xor_sync(gp.mask, sm, 1, G); /* the pair's other tangent (pairs start on even rows) */ \
          const real lm = limv[so];                           /* both lanes of a pair carry the same limit */ \
          /* implicit cone, branch-free: inside the box |.| <= lm both tangents keep their value, else this one is */ \
          /* scaled onto the circle of radius lm (rsqrt.approx.ftz: the argument is >= lm^2 > 0 whenever it is used) */ \
          const real cs = rabs(lm * sm * rsqrt_ftz(rmax(rfma(sm, sm, ot * ot), 1e-30_r)));             \
          const real cl = rmax(rabs(sm), rabs(ot)) > lm ? cs : lm;                                   \
          const real sn = rmin(rmax(sm, -cl), cl);                                                    \
          d_ = sn - lam[NSJ + so];                                                                       \
          if ((l | 1) == (li | 1)) lam[NSJ + so] = sn;                                                   \
          const real dA = gp.shfl(d_, li), dB = gp.shfl(d_, li + 1);                                    \
          UG_PAIR_APPLY(dA, dB, NACT)                                                           \
        }                                                                                                \
      }                                                                                                  \
    }                                                                                                    \
  }
  const int nact = (Rc + G - 1) / G;
  int it = 0;
#if UG_JREG > 0
  // Contact-free tick (about half of the ticks of the ball scenes: the arm moving before the paddle exists) with a
  // warp per env: see joint_sweeps_reg. Instantiated for the two row counts that occur almost always (7 motors + 5
  // gears, plus one active limit); anything else takes the generic sweeps below.
  bool swept = false;
  if constexpr (G == 32 && NSJ == 1) {
    if (nact == 0 && (n_nc == UG_JREG || n_nc == UG_JREG + 1) && !M.generic_sweeps) {
      const unsigned a0 = aEj[0];
      if (n_nc == UG_JREG) it = joint_sweeps_reg<UG_JREG>(s, gp, a0, rhs[0], invd[0], lo[0], hi[0], den[0], lam[0]);
      else it = joint_sweeps_reg<UG_JREG + 1>(s, gp, a0, rhs[0], invd[0], lo[0], hi[0], den[0], lam[0]);
      swept = true;
    }
  }
  if (!swept)
#endif
#pragma unroll 1
  for (; it < UGP_SOLVER_ITERS;) {
    if (nact == 0) { UG_JOINT_SWEEP(0) }
    else if (nact == 1) { UG_JOINT_SWEEP(1) UG_CONTACT_SWEEP(1) }
    else if (NSC >= 2 && nact == 2) { UG_JOINT_SWEEP(2) UG_CONTACT_SWEEP(2) }
    else if (NSC >= 3 && nact == 3) { UG_JOINT_SWEEP(3) UG_CONTACT_SWEEP(3) }
    else if (NSC >= 4) { UG_JOINT_SWEEP(4) UG_CONTACT_SWEEP(4) }
    it++;
    // least-squares residual of the sweep: max over rows of the squared velocity change (Bullet's early exit)
    real res = 0._r;
#pragma unroll
    for (int k = 0; k < NS; k++) { real r_ = (lam[k] - lamp[k]) * den[k]; res = rmax(res, r_ * r_); lamp[k] = lam[k]; }
    if (!gp.any(res > UGP_RESIDUAL_THRESHOLD)) break;  // max over the rows <= threshold, as one vote
  }
#undef UG_JOINT_STEP
#undef UG_TP_INIT
#undef UG_ROW_LOAD
#undef UG_PAIR_LOAD
#undef UG_ROW_APPLY
#undef UG_PAIR_APPLY
#undef UG_JOINT_SWEEP
#undef UG_CONTACT_SWEEP
  *iters_out = it;
  UG_TP(5)
  // ---- velocity change: robot dv from the stored responses, free bodies from the summed generalized impulse
#pragma unroll
  for (int sj = 0; sj < NSJ; sj++) { const int j = l + G * sj; if (j < n_nc) s.lam[j] = lam[sj]; }
#pragma unroll
  for (int so = 0; so < NSC; so++) { const int c = l + G * so; if (c < Rc) s.lam[UG_JMAX + c] = lam[NSJ + so]; }
  gp.sync();
  f3 gl[UG_MAXFREE], ga[UG_MAXFREE];
#pragma unroll
  for (int b = 0; b < UG_MAXFREE; b++) { gl[b] = mk3(0, 0, 0); ga[b] = mk3(0, 0, 0); }
  if (Rc > 0) {
#pragma unroll 1
    for (int c = l; c < Rc; c += G) {
      const int dsc = s.cdesc[c];
      if (((dsc >> 8) & 15) != RT_PAD) {
        CRow r; build_crow(s, M, dsc, r, false);
        const real lm = s.lam[UG_JMAX + c];
#pragma unroll
        for (int b = 0; b < UG_MAXFREE; b++) {
          if (r.fa == b) { gl[b] = gl[b] + r.JlA * lm; ga[b] = ga[b] + r.JaA * lm; }
          if (r.fb == b) { gl[b] = gl[b] + r.JlB * lm; ga[b] = ga[b] + r.JaB * lm; }
        }
      }
    }
#pragma unroll
    for (int b = 0; b < UG_MAXFREE; b++) {
      gl[b] = mk3(gp.sum(gl[b].x), gp.sum(gl[b].y), gp.sum(gl[b].z));
      ga[b] = mk3(gp.sum(ga[b].x), gp.sum(ga[b].y), gp.sum(ga[b].z));
    }
  }
  real absq = 0._r;
  if (l < UG_NDOF) {
    real dv = 0._r;
#pragma unroll 1
    for (int j = 0; j < n_nc; j++) dv = rfma(s.lam[j], s.E[(s.jdesc[j] & 255) * UG_NCOL + l], dv);
#pragma unroll 1
    for (int c = 0; c < Rc; c++) dv = rfma(s.lam[UG_JMAX + c], s.E[(UG_NCOL + c) * UG_NCOL + l], dv);
    real v = rmin(rmax(s.vstar[l] + dv, -UGP_MAX_COORD_VEL), UGP_MAX_COORD_VEL);
    s.qd[l] = v; s.q[l] += UGP_DT * v;
    absq = rabs(UGP_DT * v);
  } else if (l < UG_NDOF + UG_MAXFREE) {
    int b = l - UG_NDOF;
    if (b < M.nfree && (s.flags & (FL_BODY0 << b))) {
      const real* vb = s.vstar + UG_NDOF + 6 * b;
      real im = 1.0_r / M.fmass[b];
      f3 lin;
      if (!S::ANCHORS || M.fanchor[b] == 0) lin = ld3(vb) + pick3(gl, b) * im;
      else lin = ld3(vb) + lin_inv(M, b, pick3(gl, b));
      f3 ang = ld3(vb + 3) + symv(s.fIinv[b], pick3(ga, b));
      st3(s.flin[b], lin); st3(s.fang[b], ang);
      st3(s.fpos[b], ld3(s.fpos[b]) + lin * UGP_DT);
      real wn = norm3(ang);
      if (wn * UGP_DT > UGP_ANGULAR_MOTION_THRESHOLD) wn = UGP_ANGULAR_MOTION_THRESHOLD * UGP_INV_DT;
      f3 axv;
      if (wn < 0.001_r) axv = ang * (0.5_r * UGP_DT - UGP_DT * UGP_DT * UGP_DT * 0.020833333333_r * wn * wn);
      else axv = ang * (rsin(0.5_r * wn * UGP_DT) / wn);
      q4 qn = qnormalize(qmul(mkq(axv.x, axv.y, axv.z, rcos(0.5_r * wn * UGP_DT)), mkq(s.fquat[b][0], s.fquat[b][1], s.fquat[b][2], s.fquat[b][3])));
      s.fquat[b][0] = qn.x; s.fquat[b][1] = qn.y; s.fquat[b][2] = qn.z; s.fquat[b][3] = qn.w;
    }
  }
  absq = gp.sum(absq);
  gp.sync();
  UG_TP(6)
  return absq;
}

// One stepSimulation tick. Returns sum over joints of |q(t) - q(t-1)|.
template <int G, int RC, class S>
UG_DEV real tick(S& s, const DevModel& M, const Gp<G>& gp, int* iters) {
  UG_T0();
  fk_bodies(s, M, gp);
  UG_TP(0)
  collide(s, M, gp);
  UG_TP(1)
  dynamics(s, M, gp);
  UG_TP(2)
  return solve_and_integrate<G, RC>(s, M, gp, iters);
}

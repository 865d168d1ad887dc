// Env layer on the device: action -> IK / motor targets, wait_until_stable, reward / termination / observation,
// and the reset sequences, for the five in-scope envs. One warp runs one env through a small phase machine so that
// the tick body has a single call site. Reference files: see each function.
#pragma once
#include "dev_solver.cuh"

struct KParams {
  const DevModel* model; real* state; int N;
  const real* action; const real* u; const real* draws; const uint8_t* mask;
  real* obs; real* reward; uint8_t* term; uint8_t* trunc; uint8_t* succ; real* final_obs; int32_t* nticks;
  int32_t* contacts; int32_t* ccount; uint32_t* hitmask; double* stats;
  const int32_t* inject_pairs; const int32_t* inject_count;  // urgym_step_logic only: contact set instead of physics
  unsigned long long seed; int gidx0; int autoreset; int mode; int raw_ticks; int sparse;
  // auto-reset in a launch of its own: the step launch flags finished envs in `defer` instead of
  // resetting them; a MODE_RESET launch over that mask (deferred_pass = 1) follows on the same stream
  uint8_t* defer; int deferred_pass;
};
enum { MODE_STEP = 0, MODE_RESET = 1, MODE_TICKS = 2 };
enum { PH_STEP = 0, PH_ARM10 = 1, PH_SETTLE1 = 2, PH_SETTLE2 = 3, PH_RAW = 4 };

__device__ __constant__ real c_rest_pose[6] = {-1.5690622952052096_r, -1.5446774605904932_r, 1.343946009733127_r,
                                                -1.3708613585093699_r, -1.5707970583733368_r, 0.0009377758247187636_r};  // robot.py:182-183

UG_DEV unsigned long long splitmix(unsigned long long x) {
  x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
UG_DEV real urand(unsigned long long seed, unsigned env, unsigned episode, unsigned draw) {
  unsigned long long h = splitmix(seed ^ splitmix(((unsigned long long)env << 32) | episode) ^ splitmix(0xD1B54A32D192ED03ull * (unsigned long long)(draw + 1)));
  return (real)(h >> 40) * (1.0_r / 16777216.0_r);
}
UG_DEV real gripper_angle(real len) { return 0.715_r - rasin((len - 0.010_r) / 0.1143_r); }  // robot.py:215
UG_DEV real geometric_distance_reward(real v, real ts, real tmax) {                       // utilities.py:192-207
  v = rmax(v, 1e-10_r);
  return -rtanh(((v - ts) / v) / rmax(tmax - v, 1e-10_r));
}

// World pose of ee_link (Bullet link 7) from the body that carries it (robot.py:142-145).
template <class S>
UG_DEV void ee_pose(const S& s, const DevModel& M, f3& p, m3& R) {
  int b = M.link_body[7];
  m3 Rb = ldm(s.bR[b]);
  p = ld3(s.bp[b]) + mulv(Rb, ld3(M.link_lpos[7]));
  R = mulm(Rb, ldm(M.link_lrot[7]));
}
// Quaternion of link `li` by chaining joint quaternions from the base, as the CPU oracle does (sign convention).
UG_DEV q4 link_quat_chain(const real* q, const DevModel& M, int li) {
  int chain[UG_NLINK], n = 0;
  for (int i = li; i >= 0; i = M.link_parent[i]) chain[n++] = i;
  q4 r = mkq(0, 0, 0, 1);
  for (int t = n - 1; t >= 0; t--) {
    int i = chain[t];
    r = qmul(r, mkq(M.link_jquat[i][0], M.link_jquat[i][1], M.link_jquat[i][2], M.link_jquat[i][3]));
    if (M.link_dof[i] >= 0) r = qmul(r, quat_axis(M.link_axis[i], q[M.link_dof[i]]));
    r = qnormalize(r);
  }
  return r;
}

// calculateInverseKinematics as robot.py:119-121 calls it: damped least squares with orientation on the 6 arm
// columns, seeded from the current joints, <= 20 iterations. Every lane computes the same.
__device__ __noinline__ void ik_solve(const real* q_in, const DevModel& M, f3 tpos, q4 tq, real* out6) {
  real qa[6];
#pragma unroll
  for (int i = 0; i < 6; i++) qa[i] = q_in[i];
  tq = qnormalize(tq);
  real diff = 1e30_r;
  for (int it = 0; it < UGP_IK_ITERS && diff > UGP_IK_RESIDUAL; it++) {
    m3 R; f3 p = mk3(0, 0, 0);
    R.m[0] = 1; R.m[1] = 0; R.m[2] = 0; R.m[3] = 0; R.m[4] = 1; R.m[5] = 0; R.m[6] = 0; R.m[7] = 0; R.m[8] = 1;
    // kept rolled: six inlined copies of this body (rsincos carries its range-reduction path) made the iteration a
    // 2 k-instruction straight line that missed the instruction caches on every pass (profiles/, CubesPush)
    f3 axw[6], org[6];
#pragma unroll 1
    for (int d = 0; d < 6; d++) {
      p = p + mulv(R, ld3(M.jpos[d]));
      m3 R0 = mulm(R, ldm(M.jrot[d]));
      real sn, cs; rsincos(qa[d], &sn, &cs);
      int a = M.axis[d], j = (a + 1) % 3, k = (a + 2) % 3;
      f3 cj = col(R0, j), ck = col(R0, k);
      f3 nj = cj * cs + ck * sn, nk = ck * cs - cj * sn;
      R = R0;
      R.m[j] = nj.x; R.m[3 + j] = nj.y; R.m[6 + j] = nj.z; R.m[k] = nk.x; R.m[3 + k] = nk.y; R.m[6 + k] = nk.z;
      axw[d] = col(R, a); org[d] = p;
    }
    f3 pe = p + mulv(R, ld3(M.link_lpos[7]));
    q4 eq = mat_to_quat(mulm(R, ldm(M.link_lrot[7])));
    f3 dp = tpos - pe;
    diff = norm3(dp);
    q4 dq = qmul(tq, qconj(eq));
    f3 v = mk3(dq.x, dq.y, dq.z);
#ifdef UG_F64
    // parity build: the CPU checker's expression, term by term (its ik(): the angle goes through a float, as
    // pybullet's IK helper holds it in one [recalled]; the axis falls back to x when the sine vanishes)
    real ang = (real)(float)(2.0 * acos(rmax(-1.0, rmin(1.0, dq.w))));
    const real s2 = 1.0 - dq.w * dq.w;
    f3 axq = s2 < 10.0 * 2.220446049250313e-16 ? mk3(1, 0, 0) : v * (1.0 / sqrt(s2));
    if (ang > UGP_PI) ang -= 2.0_r * UGP_PI;
    f3 dr = axq * (ang / norm3(axq));
#else
    // fp32: the same rotation vector through atan2 (acosf loses the small angles an IK iteration converges through)
    real vn = norm3(v);
    real ang = 2.0_r * ratan2(vn, dq.w);
    if (ang > UGP_PI) ang -= 2.0_r * UGP_PI;
    f3 dr = vn > 1e-12_r ? v * (ang / vn) : mk3(0, 0, 0);
#endif
    real e[6] = {dp.x, dp.y, dp.z, dr.x, dr.y, dr.z};
    real J[6][6];
#pragma unroll
    for (int d = 0; d < 6; d++) {
      f3 jl = cross(axw[d], pe - org[d]);
      J[0][d] = jl.x; J[1][d] = jl.y; J[2][d] = jl.z; J[3][d] = axw[d].x; J[4][d] = axw[d].y; J[5][d] = axw[d].z;
    }
    real A[6][6], b[6];
#pragma unroll
    for (int a = 0; a < 6; a++) {
#pragma unroll
      for (int c = 0; c <= a; c++) { real sum = 0._r;
#pragma unroll
        for (int k = 0; k < 6; k++) sum += J[k][a] * J[k][c];
        A[a][c] = sum + (a == c ? UGP_IK_DAMPING : 0._r); }
      real sum = 0._r;
#pragma unroll
      for (int k = 0; k < 6; k++) sum += J[k][a] * e[k];
      b[a] = sum;
    }
    // Cholesky solve (J^T J + 0.5 I is symmetric positive definite)
#pragma unroll
    for (int k = 0; k < 6; k++) {
      real dk = A[k][k];
#pragma unroll
      for (int t = 0; t < k; t++) dk -= A[k][t] * A[k][t];
      real inv = rrsqrt(dk); A[k][k] = inv;
#pragma unroll
      for (int i = k + 1; i < 6; i++) { real x = A[i][k];
#pragma unroll
        for (int t = 0; t < k; t++) x -= A[i][t] * A[k][t];
        A[i][k] = x * inv; }
    }
    real x[6];
#pragma unroll
    for (int i = 0; i < 6; i++) { real t0 = b[i];
#pragma unroll
      for (int t = 0; t < i; t++) t0 -= A[i][t] * x[t];
      x[i] = t0 * A[i][i]; }
#pragma unroll
    for (int i = 5; i >= 0; i--) { real t0 = x[i];
#pragma unroll
      for (int t = i + 1; t < 6; t++) t0 -= A[t][i] * x[t];
      x[i] = t0 * A[i][i]; }
    real mx = 0._r;
#pragma unroll
    for (int i = 0; i < 6; i++) mx = rmax(mx, rabs(x[i]));
    real sc = mx > UGP_IK_MAX_STEP ? UGP_IK_MAX_STEP / mx : 1.0_r;
#pragma unroll
    for (int i = 0; i < 6; i++) qa[i] += x[i] * sc;
  }
#pragma unroll
  for (int i = 0; i < 6; i++) out6[i] = qa[i];
}

// Midpoint of the union of the world AABBs of links 11 and 16 (env_ball_balance_v0.py:249-275). Needs fresh FK.
template <class S>
UG_DEV f3 gripper_center(const S& s, const DevModel& M) {
  f3 lo = mk3(1e30_r, 1e30_r, 1e30_r), hi = mk3(-1e30_r, -1e30_r, -1e30_r);
#pragma unroll
  for (int t = 0; t < 2; t++) {
    int b = t == 0 ? M.link_body[11] : M.link_body[16];
    m3 R = ldm(s.bR[b]);
    f3 c = ld3(s.bp[b]) + mulv(R, ld3(M.aabb_c[t]));
    f3 h = ld3(M.aabb_h[t]);
    f3 e = mk3(rabs(R.m[0]) * h.x + rabs(R.m[1]) * h.y + rabs(R.m[2]) * h.z, rabs(R.m[3]) * h.x + rabs(R.m[4]) * h.y + rabs(R.m[5]) * h.z,
               rabs(R.m[6]) * h.x + rabs(R.m[7]) * h.y + rabs(R.m[8]) * h.z);
    lo = mk3(rmin(lo.x, c.x - e.x), rmin(lo.y, c.y - e.y), rmin(lo.z, c.z - e.z));
    hi = mk3(rmax(hi.x, c.x + e.x), rmax(hi.y, c.y + e.y), rmax(hi.z, c.z + e.z));
  }
  return (lo + hi) * 0.5_r;
}
template <class S>
UG_DEV real gripper_opening(const S& s, const DevModel& M) {  // env_cubes_grasp_v0.py:340-377
  real v = norm3(ld3(s.bp[M.link_body[16]]) - ld3(s.bp[M.link_body[11]])) - 0.052_r;
  return rmin(rmax(v, 0._r), 0.085_r) / 0.085_r;
}

template <int G, class S>
UG_DEV void load_state(S& s, const real* rec, const Gp<G>& gp) {
  const int lane = gp.l;
  real* st = s.u.st;
  for (int i = lane; i < UG_STATE; i += G) st[i] = rec[i];
  gp.sync();
  if (lane < UG_NDOF) { s.q[lane] = st[lane]; s.qd[lane] = st[12 + lane]; }
  if (lane < 7) s.tgt[lane] = st[24 + lane];
  if (lane == 0) s.flags = (int)st[31];
  if (lane < UG_MAXFREE) {
    const real* p = st + 38 + 13 * lane;
    st3(s.fpos[lane], ld3(p)); s.fquat[lane][0] = p[3]; s.fquat[lane][1] = p[4]; s.fquat[lane][2] = p[5]; s.fquat[lane][3] = p[6];
    st3(s.flin[lane], ld3(p + 7)); st3(s.fang[lane], ld3(p + 10));
  }
  gp.sync();
}
// The record's unused slots (78..95) are rewritten as zeros; slot 77 carries the running episode return.
template <int G, class S>
UG_DEV void store_state(S& s, real* rec, const Gp<G>& gp, int step_count, int episode, q4 vert, real ep_ret) {
  const int lane = gp.l;
  real* st = s.u.st;
  gp.sync();
  for (int i = lane; i < UG_STATE; i += G) st[i] = 0._r;
  gp.sync();
  if (lane < UG_NDOF) { st[lane] = s.q[lane]; st[12 + lane] = s.qd[lane]; }
  if (lane < 7) st[24 + lane] = s.tgt[lane];
  if (lane == 0) { st[31] = (real)s.flags; st[32] = (real)step_count; st[33] = (real)episode; st[34] = vert.x; st[35] = vert.y; st[36] = vert.z; st[37] = vert.w; st[77] = ep_ret; }
  if (lane < UG_MAXFREE) {
    real* p = st + 38 + 13 * lane;
    st3(p, ld3(s.fpos[lane])); p[3] = s.fquat[lane][0]; p[4] = s.fquat[lane][1]; p[5] = s.fquat[lane][2]; p[6] = s.fquat[lane][3];
    st3(p + 7, ld3(s.flin[lane])); st3(p + 10, ld3(s.fang[lane]));
  }
  gp.sync();
  for (int i = lane; i < UG_STATE; i += G) rec[i] = st[i];
}

template <class S>
UG_DEV void spawn(S& s, int b, f3 pos, int lane) {
  if (lane == 0) {
    st3(s.fpos[b], pos); s.fquat[b][0] = 0._r; s.fquat[b][1] = 0._r; s.fquat[b][2] = 0._r; s.fquat[b][3] = 1._r;
    st3(s.flin[b], mk3(0, 0, 0)); st3(s.fang[b], mk3(0, 0, 0)); s.flags |= FL_BODY0 << b;
  }
}
template <class S>
UG_DEV void set_arm_targets(S& s, const real* t, int lane) {
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < 6; i++) s.tgt[i] = t[i];
    s.flags |= FL_ARM;
  }
}
template <int G, class S>
UG_DEV void move_ee(S& s, const DevModel& M, f3 pos, q4 quat, const Gp<G>& gp) {
  real t[6]; ik_solve(s.q, M, pos, quat, t);
  gp.sync();
  set_arm_targets(s, t, gp.l);
}
template <class S>
UG_DEV void move_gripper(S& s, real len, int lane) { if (lane == 0) { s.tgt[6] = gripper_angle(len); s.flags |= FL_GRIP; } }

// Contact predicates over the pairs that carried a manifold point in the last tick (s.hit, set before the capacity cut).
// touch_robot: any pair (robot link in [l0,l1]) x (B.body == bb, or any static when bb < 0; optional B.link == blink).
template <class S>
UG_DEV bool touch_robot(const S& s, const DevModel& M, int l0, int l1, int bb, int blink) {
  bool hit = false;
#pragma unroll 1
  for (int w = 0; w < UG_HITWORDS; w++) {
    unsigned m = s.hit[w];
    while (m) {
      const int p = 32 * w + __ffs(m) - 1; m &= m - 1;
      const DevShape& A = M.shapes[M.pairs[p][0]]; const DevShape& B = M.shapes[M.pairs[p][1]];
      if (A.body == 0 && A.link >= l0 && A.link <= l1 && (bb < 0 ? B.body < 0 : B.body == bb) && (blink < 0 || B.link == blink)) hit = true;
    }
  }
  return hit;
}
template <class S>
UG_DEV bool touch_free(const S& s, const DevModel& M, int ba, int bb, int blink) {
  bool hit = false;
#pragma unroll 1
  for (int w = 0; w < UG_HITWORDS; w++) {
    unsigned m = s.hit[w];
    while (m) {
      const int p = 32 * w + __ffs(m) - 1; m &= m - 1;
      const DevShape& A = M.shapes[M.pairs[p][0]]; const DevShape& B = M.shapes[M.pairs[p][1]];
      if (A.body == ba && (bb < 0 ? B.body < 0 : B.body == bb) && (blink < 0 || B.link == blink)) hit = true;
    }
  }
  return hit;
}
template <class S>
UG_DEV f3 club_foot(const S& s) {  // env_golf_v0.py:368-373: COM of the club's link 0 (the transversal part)
  m3 R = quat_to_mat(mkq(s.fquat[0][0], s.fquat[0][1], s.fquat[0][2], s.fquat[0][3]));
  return ld3(s.fpos[0]) + mulv(R, mk3(0._r, 0._r, (real)(-0.09 - UGP_CLUB_ZC)));
}

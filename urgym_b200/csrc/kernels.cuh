// The fused env kernel: one launch advances every env by one gym step (action -> IK -> 1..480 ticks -> epilogue ->
// optional in-kernel auto-reset). State crosses HBM once in and once out per gym step.
#pragma once
#include "dev_env.cuh"

// Registered env -> the world it lives in. CubesPushDiff (5) and GolfJoints (6) differ from CubesPush (2) and Golf (4)
// in the action mapping and the observation only (env_cubes_push_diff_v0.py:209-211,
// env_golf_joints_v0.py:60-63,85,123-124,229-234).
template <int ENVF> struct Scene { static constexpr int value = ENVF == 5 ? 2 : (ENVF == 6 ? 4 : ENVF); };
template <int ENVF> struct ObsDim { static constexpr int value = ENVF == 0 ? 6 : (ENVF == 3 ? 17 : (ENVF == 6 ? 15 : (ENVF == 7 ? 7 : 9))); };
template <int ENVF> struct ActDim { static constexpr int value = ENVF == 3 ? 9 : (ENVF == 6 ? 6 : (ENVF == 7 ? 8 : 3)); };

// One out-of-line copy per env type (three call sites; the ball scenes' Euler conversion alone is ~0.9 k instructions):
// once-per-step straight-line code is what the instruction caches miss on.
template <int ENVF, class S>
__device__ __noinline__ void observe(S& s, const DevModel& M, real* o, real* o2, int lane) {
  // caller guarantees fresh fk_bodies(); every lane computes, lane 0 writes (to o, and to o2 when given)
  constexpr int ENV = Scene<ENVF>::value;
  f3 ep; m3 eR; ee_pose(s, M, ep, eR);
  real v[17];
  if (ENV == 0 || ENV == 1) {
    f3 e = euler_from_quat(mkq(s.fquat[0][0], s.fquat[0][1], s.fquat[0][2], s.fquat[0][3]));
    v[0] = e.x; v[1] = e.y; v[2] = e.z;
#pragma unroll
    for (int b = 1; b <= (ENV == 0 ? 1 : 2); b++) { f3 d = ld3(s.fpos[0]) - ld3(s.fpos[b]); v[3 * b] = d.x; v[3 * b + 1] = d.y; v[3 * b + 2] = d.z; }
  } else if (ENV == 2) {
    v[0] = ep.x; v[1] = ep.y; v[2] = ep.z;
#pragma unroll
    for (int b = 0; b < 2; b++) { v[3 + 3 * b] = s.fpos[b][0]; v[4 + 3 * b] = s.fpos[b][1]; v[5 + 3 * b] = s.fpos[b][2]; }
    if (ENVF == 5) {
#pragma unroll
      for (int b = 0; b < 2; b++) { v[3 + 3 * b] -= ep.x; v[4 + 3 * b] -= ep.y; v[5 + 3 * b] -= ep.z; }
    }
  } else if (ENV == 3) {
    q4 eq = link_quat_chain(s.q, M, 7);
    v[0] = (s.flags & FL_APPROACHED) ? 1._r : 0._r; v[1] = (s.flags & FL_GRASPED) ? 1._r : 0._r;
    v[2] = ep.x; v[3] = ep.y; v[4] = ep.z; v[5] = eq.x; v[6] = eq.y; v[7] = eq.z; v[8] = eq.w;
    v[9] = gripper_opening(s, M);
    v[10] = s.fpos[0][0]; v[11] = s.fpos[0][1]; v[12] = s.fpos[0][2];
    v[13] = s.fquat[0][0]; v[14] = s.fquat[0][1]; v[15] = s.fquat[0][2]; v[16] = s.fquat[0][3];
  } else if (ENV == 7) {                // env_box_v0.py:161-170: end-effector position and quaternion
    q4 eq = link_quat_chain(s.q, M, 7);
    v[0] = ep.x; v[1] = ep.y; v[2] = ep.z; v[3] = eq.x; v[4] = eq.y; v[5] = eq.z; v[6] = eq.w;
  } else {
    f3 st = club_foot(s), b = ld3(s.fpos[1]), h = mk3(0._r, -0.6_r, 0.02_r);
    f3 d1 = st - b, d2 = h - b;
    constexpr int J0 = ENVF == 6 ? 6 : 0;
    if (ENVF == 6) {
#pragma unroll
      for (int i = 0; i < 6; i++) v[i] = s.q[i];
    }
    v[J0 + 0] = st.x; v[J0 + 1] = st.y; v[J0 + 2] = st.z; v[J0 + 3] = d1.x; v[J0 + 4] = d1.y; v[J0 + 5] = d1.z;
    v[J0 + 6] = d2.x; v[J0 + 7] = d2.y; v[J0 + 8] = d2.z;
  }
  constexpr int OD = ObsDim<ENVF>::value;
  if (lane == 0) {
#pragma unroll
    for (int i = 0; i < OD; i++) o[i] = v[i];
    if (o2) {
#pragma unroll
      for (int i = 0; i < OD; i++) o2[i] = v[i];
    }
  }
}

// Capacities per scene, from the distribution of manifold points per tick on the uncapped checker under the benchmarked
// random policies (tools/contact_hist.py, profiles/r02_contact_hist.txt; gripper links as convex polytopes): the ball scenes
// exceed 16 points on 11-14 % of the ticks (the paddle rests on pads, inner fingers and knuckles) and never 19; CubesGrasp
// peaks at 20; CubesPush exceeds 20 on 2-4e-4 of the ticks (two cubes on the table plus a finger on both) and never 23, and
// the fp64 build showed that those ticks alone changed trip counts against the uncapped checker, so it gets 24; the Golf scenes (club in the fingers and on the pitch) exceed 24 on 2.5-3.5 % and 32 on
// 0.07-1 %. Rolling friction exists in the Golf scenes only (<= 3 manifolds seen, 4 kept). Rows = 3 per point (normal +
// friction pair) + 2 per rolling manifold. Ticks that exceed a capacity are counted (urgym_stats out[6]).
template <int ENVF> struct EnvNC { static constexpr int value = Scene<ENVF>::value == 4 ? 32 : ((Scene<ENVF>::value == 2 || Scene<ENVF>::value == 7) ? 24 : 20); };  // Box-v0: 20 + its 5 joint rows
template <int ENVF> struct EnvRC { static constexpr int value = 3 * EnvNC<ENVF>::value + (Scene<ENVF>::value == 4 ? 8 : 0); };
template <int ENVF> using EnvSmem = SmemT<EnvNC<ENVF>::value, EnvRC<ENVF>::value, Scene<ENVF>::value == 7>;

// Resident CTAs per SM asked of ptxas = what the scene's shared memory allows: 14 at 15.5 KB per env (ball scenes, CubesGrasp), 11 at
// 19.8 KB (CubesPush, Box: they may use up to 186 registers instead of 128; +2.9 % CubesPush when first allowed 164). The Golf kernels keep the 128-register cap although only 6 CTAs fit their 34.5 KB: with 240
// registers and no spills they measured 2-5 % SLOWER (profiles/r02_ab_running_pointers.txt), so the cap stays.
#ifdef UG_F64
template <int ENVF> struct EnvMinCtas { static constexpr int value = UG_MIN_CTAS; };
#else
template <int ENVF> struct EnvMinCtas { static constexpr int value = EnvRC<ENVF>::value <= 60 ? 14 : (EnvRC<ENVF>::value <= 100 ? 11 : UG_MIN_CTAS); };
#endif
template <int ENVF, int G>
__global__ void __launch_bounds__(32, EnvMinCtas<ENVF>::value) env_kernel(KParams P) {
  constexpr int ENV = Scene<ENVF>::value;
  using S = EnvSmem<ENVF>;
  constexpr int RC = EnvRC<ENVF>::value;
  extern __shared__ __align__(16) unsigned char smraw[];
  const Gp<G> gp = make_gp<G>();
  S& s = reinterpret_cast<S*>(smraw)[(threadIdx.x & 31) / G];
  const DevModel& M = *P.model;
  const int env = blockIdx.x * (32 / G) + (threadIdx.x & 31) / G, lane = gp.l;
  constexpr int OD = ObsDim<ENVF>::value;
  constexpr int AD = ActDim<ENVF>::value;
  if (env >= P.N) return;
  if (P.mode == MODE_RESET && P.mask && !P.mask[env]) return;
  real* rec = P.state + (size_t)env * UG_STATE;
  load_state(s, rec, gp);
  if (gp.l == 0) s.ovf = 0;
  int step_count = (int)s.u.st[32], episode = (int)s.u.st[33];
  q4 vert = mkq(s.u.st[34], s.u.st[35], s.u.st[36], s.u.st[37]);
  real ep_ret = s.u.st[77];
  gp.sync();
  const unsigned gidx = (unsigned)(P.gidx0 + env);
  const real* dr = P.draws ? P.draws + 4 * env : nullptr;
  int phase, n = 0, limit = UGP_SETTLE_MAX_TICKS, ticks_step = 0, ticks_reset = 0;
  long long tick_total = 0;
  real rew = 0._r; bool term = false, succ = false, trunc = false;

  // ---- what starts the tick loop
  if (P.mode == MODE_TICKS) { phase = PH_RAW; limit = P.raw_ticks; }
  else if (P.mode == MODE_RESET) { phase = PH_ARM10; }
  else {
    phase = PH_STEP;
    const real* a = P.action + (size_t)AD * env;
    fk_bodies(s, M, gp);
    f3 ep; m3 eR; ee_pose(s, M, ep, eR);
    if (ENV == 0 || ENV == 1) {          // env_ball_balance_v0.py:112-118
      if (ENV == 1) rew -= rabs(a[0]) + rabs(a[1]) + rabs(a[2]);
      f3 e = euler_from_quat(mat_to_quat(eR));
      move_ee(s, M, ep, quat_from_euler(e.x + a[0], e.y + a[1], e.z + a[2]), gp);
    } else if (ENVF == 6) {              // env_golf_joints_v0.py:123-124: joint targets = current arm joints + action
      real t[6];
#pragma unroll
      for (int i = 0; i < 6; i++) t[i] = s.q[i] + a[i];
      gp.sync();
      set_arm_targets(s, t, lane);
    } else if (ENV == 7) {               // env_box_v0.py:124-126: pose (position + quaternion) + action[:7] component-wise, gripper length
      q4 eq = link_quat_chain(s.q, M, 7);
      move_ee(s, M, ep + mk3(a[0], a[1], a[2]), mkq(eq.x + a[3], eq.y + a[4], eq.z + a[5], eq.w + a[6]), gp);
      move_gripper(s, a[7], lane);
      limit = 120;                       // env_box_v0.py:127-128: a fixed 120 ticks instead of wait_until_stable
    } else if (ENV == 2 || ENV == 4) {   // env_cubes_push_v0.py:122-126, env_golf_v0.py:122-126
      move_ee(s, M, ep + mk3(a[0], a[1], a[2]), vert, gp);
    } else {                              // env_cubes_grasp_v0.py:121-144
      real a0 = a[0], a1 = a[1];
      real u = P.u ? P.u[env] : urand(P.seed, gidx, (unsigned)episode, 16u + (unsigned)step_count);
      int sel = (a0 + a1 == 0._r) ? (u < 0.5_r ? 0 : 1) : (u * (a0 + a1) < a0 ? 0 : 1);
      if (sel == 0) { real t[6];
#pragma unroll
        for (int i = 0; i < 6; i++) t[i] = a[2 + i];
        set_arm_targets(s, t, lane); }
      else move_gripper(s, a[8] < 0.5_r ? 0.085_r : 0.0_r, lane);
    }
  }
  // robot.py:85-101: arm to rest pose with zero velocity; episode counters restart. The previous episode's task objects
  // stay in the world for the ten arm ticks: every env's reset() calls robot.reset() before it removes them
  // (env_ball_balance_v0.py:147-152, env_cubes_push_v0.py:214-215 + 238-241, env_golf_v0.py:240-248)
  auto begin_reset = [&]() {
    episode++; step_count = 0; ep_ret = 0._r;
    gp.sync();
    if (lane == 0) s.flags &= ~(FL_APPROACHED | FL_GRASPED);
    if (lane < 6) { s.q[lane] = c_rest_pose[lane]; s.qd[lane] = 0._r; }
    phase = PH_ARM10; n = 0; limit = UGP_RESET_ARM_TICKS;
    gp.sync();
  };
  if (phase == PH_ARM10) begin_reset();
  gp.sync();

  // ---- the tick loop
  for (;;) {
    int iters;
    real moved = 0._r;
    if (P.inject_pairs) {  // urgym_step_logic: no physics, the caller names the contact set the predicates see
      gp.sync();
      const int nc = min(P.inject_count[env], UG_MAXC);
      if (lane < UG_HITWORDS) s.hit[lane] = 0u;
      if (lane == 0) s.ncon = 0;
      gp.sync();
      for (int k = 0; k < nc; k++) {
        const int sa = P.inject_pairs[((size_t)env * UG_MAXC + k) * 2], sb = P.inject_pairs[((size_t)env * UG_MAXC + k) * 2 + 1];
        for (int p = lane; p < M.npairs; p += G) if (M.pairs[p][0] == sa && M.pairs[p][1] == sb) atomicOr(&s.hit[p >> 5], 1u << (p & 31));
      }
      gp.sync();
    } else {
      moved = tick<G, RC>(s, M, gp, &iters);
    }
    n++; tick_total++;
    bool done = phase == PH_ARM10 || phase == PH_RAW || ENV == 7 ? n >= limit : (moved < UGP_SETTLE_THRESHOLD || n >= limit);
    if (!done) continue;
    if (phase == PH_RAW) break;
    fk_bodies(s, M, gp);
    f3 ep; m3 eR; ee_pose(s, M, ep, eR);
    if (phase == PH_STEP) {
      ticks_step = n;
      // ---------------- epilogue: reward / termination
      if (ENV == 0) {                      // env_ball_balance_v0.py:122-132
        rew += 1._r; term = s.fpos[1][2] < s.fpos[0][2];
      } else if (ENV == 1) {               // env_two_balls_balance_v0.py:128-146
        rew += 1._r;
#pragma unroll
        for (int b = 1; b <= 2; b++)
#pragma unroll
          for (int k = 0; k < 2; k++) { real x = rabs(s.fpos[b][k]); rew += x > 1._r ? 0._r : 1._r - x * x; }
        term = s.fpos[1][2] < s.fpos[0][2] || s.fpos[2][2] < s.fpos[0][2];
      } else if (ENV == 2) {               // env_cubes_push_v0.py:130-154
        bool table = touch_robot(s, M, 1, 18, -1, 0);   // the table (static id 0), not the ground plane under it
        bool fingers = touch_robot(s, M, 9, 18, 2, -1);
        if (table) rew -= 0.01_r;
        else if (fingers) {
          real d = norm3(ld3(s.fpos[0]) - ld3(s.fpos[1]));
          rew = 0._r;
          if (!P.sparse) rew += geometric_distance_reward(d, 0.1_r, 0.5_r) / 10._r;
          if (d < 0.05_r) { rew += 10._r; succ = true; }
          rew += 0.1_r; term = succ;
        }
        rew -= 0.1_r;
      } else if (ENV == 3) {               // env_cubes_grasp_v0.py:146-173
        f3 gc = gripper_center(s, M);
        real d = norm3(gc - ld3(s.fpos[0]));
        bool grasp_now = touch_robot(s, M, 11, 13, 1, -1) && touch_robot(s, M, 16, 18, 1, -1);
        int fl = s.flags;
        gp.sync();
        if (!(fl & FL_APPROACHED)) {
          rew += geometric_distance_reward(d, 0.5_r, 2._r) / 4._r;
          if (d < 0.05_r && gripper_opening(s, M) > 0.50_r) { fl |= FL_APPROACHED; rew += 5._r; }
        } else if (!(fl & FL_GRASPED)) {
          if (grasp_now) { fl |= FL_GRASPED; rew += 5._r; }
        } else if (grasp_now && s.fpos[0][2] > 0.2_r) { rew += 10._r; term = true; succ = true; }
        if (lane == 0) s.flags = fl;
        rew -= 0.1_r;
      } else if (ENV == 7) {               // env_box_v0.py:141-159 (button_touch_mode='any'), :129-137; sticky bits as in CubesGrasp
        const real door = anchor_q(s, M, 0), button = anchor_q(s, M, 1);
        int fl = s.flags;
        gp.sync();
        if (!(fl & FL_APPROACHED)) { if (door > 1.9_r) { fl |= FL_APPROACHED; rew = 1._r; } }      // box_opened
        else if (!(fl & FL_GRASPED)) { if (button < -0.02_r) { fl |= FL_GRASPED; rew = 1._r; } }   // button_pressed
        else if (door < 0.1_r) { rew = 1._r; term = true; succ = true; }                            // box_closed
        if (lane == 0) s.flags = fl;
        rew -= 0.1_r;
      } else {                             // env_golf_v0.py:130-183
        f3 hole = mk3(0._r, -0.6_r, 0.02_r), ball = ld3(s.fpos[1]);
        bool in_fingers = touch_robot(s, M, 9, 18, 1, -1);
        if (norm3(hole - ball) < 0.05_r) { rew += 10._r; succ = true; term = true; }
        else if (!in_fingers) { rew -= 5._r; term = true; }
        else if (ball.z < 0._r) { rew -= 10._r; term = true; }
        else {
          bool floor_hit = touch_robot(s, M, 0, 18, -1, 1) || touch_free(s, M, 1, -1, 1);
          if (floor_hit) rew -= 0.01_r;
          if (!P.sparse) rew += geometric_distance_reward(norm3(club_foot(s) - ball), 0.1_r, 0.5_r) / 10._r;
          if (touch_free(s, M, 1, 2, -1)) { rew += 0.1_r; if (!P.sparse) rew += geometric_distance_reward(norm3(hole - ball), 0.1_r, 0.5_r) / 10._r; }
        }
        rew -= 0.1_r;
      }
      gp.sync();
      step_count++; ep_ret += rew;
      trunc = step_count >= M.max_steps;
      if (P.contacts) {
        const int np0 = (!S::ANCHORS || P.inject_pairs) ? 0 : s.npseudo;   // joint rows of anchored bodies sit at the head of con and are not contacts
        int nc = P.inject_pairs ? 0 : s.ncon - np0;
        for (int k = lane; k < UG_MAXC; k += G) { P.contacts[((size_t)env * UG_MAXC + k) * 2] = k < nc ? (s.con[np0 + k].shapes & 255) : -1; P.contacts[((size_t)env * UG_MAXC + k) * 2 + 1] = k < nc ? (s.con[np0 + k].shapes >> 8) : -1; }
        if (lane == 0) P.ccount[env] = nc;
        if (lane < UG_HITWORDS) P.hitmask[(size_t)env * UG_HITWORDS + lane] = s.hit[lane];
      }
      observe<ENVF>(s, M, P.obs + (size_t)OD * env, P.final_obs ? P.final_obs + (size_t)OD * env : nullptr, lane);
      if (lane == 0) { P.reward[env] = rew; P.term[env] = term; P.trunc[env] = trunc; P.succ[env] = succ; }
      if ((term || trunc) && lane == 0 && P.stats) {
        atomicAdd(&P.stats[0], 1.0); atomicAdd(&P.stats[1], (double)ep_ret); atomicAdd(&P.stats[2], (double)step_count); atomicAdd(&P.stats[3], succ ? 1.0 : 0.0);
      }
      if (P.defer) {  // the reset runs in the launch that follows
        if (lane == 0) P.defer[env] = (uint8_t)(term || trunc);
        break;
      }
      if ((term || trunc) && P.autoreset) { begin_reset(); continue; }
      break;
    }
    ticks_reset += n;
    if (phase == PH_ARM10) {
      // ---------------- reset stage 1 (robot.py:100-101 opens the gripper; the env removes the old task objects; then the
      // env-specific first move)
      move_gripper(s, 0.085_r, lane);
      if (lane == 0) s.flags &= (FL_ARM | FL_GRIP);
      gp.sync();
      if (ENV == 0 || ENV == 1) {          // env_ball_balance_v0.py:154-158
        move_ee(s, M, mk3(0.0_r, -0.60_r, 0.60_r), mkq(0.50_r, -0.50_r, -0.50_r, 0.50_r), gp);
        move_gripper(s, 0.085_r * 0.2_r, lane);
        phase = PH_SETTLE1;
      } else if (ENV == 2) {               // env_cubes_push_v0.py:212-227,238-246
        real d0 = dr ? dr[0] : urand(P.seed, gidx, (unsigned)episode, 0u);
        f3 c1 = mk3(-0.1_r + 0.2_r * d0, -0.7_r, 0.1_r);
        spawn(s, 0, mk3(0.0_r, -0.5_r, 0.1_r), lane); spawn(s, 1, c1, lane);
        real sh, ch; rsincos(0.25_r * UGP_PI, &sh, &ch);
        vert = mkq(0._r, sh, 0._r, ch);
        move_ee(s, M, c1 + mk3(0._r, 0._r, 0.2_r), vert, gp);
        move_gripper(s, 0.0_r, lane);
        phase = PH_SETTLE2;
      } else if (ENV == 3) {               // env_cubes_grasp_v0.py:276-299,311-327
        real d0 = dr ? dr[0] : urand(P.seed, gidx, (unsigned)episode, 0u), d1 = dr ? dr[1] : urand(P.seed, gidx, (unsigned)episode, 1u);
        real r = 0.5_r + 0.3_r * d0, th = -UGP_PI + UGP_PI * d1, sn, cs; rsincos(th, &sn, &cs);
        spawn(s, 0, mk3(r * cs, r * sn, 0.1_r), lane);
        move_ee(s, M, ep, link_quat_chain(s.q, M, 7), gp);
        phase = PH_SETTLE2;
      } else if (ENV == 7) {               // env_box_v0.py:178-185: move_ee to the current pose; reset_box(): both joints to 0
        move_ee(s, M, ep, link_quat_chain(s.q, M, 7), gp);
        spawn(s, 0, ld3(M.forigin[0]), lane); spawn(s, 1, ld3(M.forigin[1]), lane);
        gp.sync();
        observe<ENVF>(s, M, P.obs + (size_t)OD * env, nullptr, lane);
        break;
      } else {                             // env_golf_v0.py:238-256
        vert = mkq(0.0_r, 0.0_r, -0.70_r, 0.70_r);
        move_ee(s, M, mk3(0.0_r, -0.60_r, 0.40_r), vert, gp);
        move_gripper(s, 0.085_r * 0.2_r, lane);
        phase = PH_SETTLE1;
      }
      n = 0; limit = UGP_SETTLE_MAX_TICKS;
      gp.sync();
      continue;
    }
    if (phase == PH_SETTLE1) {
      // ---------------- reset stage 2: spawn the held object and the ball(s), close the gripper
      f3 gc = gripper_center(s, M);
      if (ENV == 0 || ENV == 1) {          // env_ball_balance_v0.py:160-168,186-205; env_two_balls_balance_v0.py:200-233
        f3 pp = mk3(ep.x, ep.y - 0.2_r, gc.z);
        real d0 = dr ? dr[0] : urand(P.seed, gidx, (unsigned)episode, 0u), d1 = dr ? dr[1] : urand(P.seed, gidx, (unsigned)episode, 1u);
        spawn(s, 0, pp, lane);
        if (ENV == 0) spawn(s, 1, pp + mk3(-0.03_r + 0.06_r * d0, -0.08_r + 0.10_r * d1, 0.2_r + 0.02_r), lane);
        else {
          real d2 = dr ? dr[2] : urand(P.seed, gidx, (unsigned)episode, 2u), d3 = dr ? dr[3] : urand(P.seed, gidx, (unsigned)episode, 3u);
          spawn(s, 1, pp + mk3(-0.03_r + 0.06_r * d0, -0.08_r + 0.035_r * d1, 0.2_r + 0.003_r), lane);
          spawn(s, 2, pp + mk3(-0.03_r + 0.06_r * d2, -0.015_r + 0.035_r * d3, 0.2_r - 0.003_r), lane);
        }
      } else if (ENV == 4) {               // env_golf_v0.py:258-266,335-340
        real d0 = dr ? dr[0] : urand(P.seed, gidx, (unsigned)episode, 0u), d1 = dr ? dr[1] : urand(P.seed, gidx, (unsigned)episode, 1u);
        spawn(s, 0, gc + mk3(0._r, 0._r, -0.05_r + (real)UGP_CLUB_ZC), lane);
        spawn(s, 1, mk3(-0.2_r + 0.4_r * d0, -0.7_r + (-0.1_r + 0.2_r * d1), 0.1_r), lane);
      }
      move_gripper(s, 0.0_r, lane);
      phase = PH_SETTLE2; n = 0; limit = UGP_SETTLE_MAX_TICKS;
      gp.sync();
      continue;
    }
    // PH_SETTLE2: reset finished -> first observation of the new episode
    observe<ENVF>(s, M, P.obs + (size_t)OD * env, nullptr, lane);
    break;
  }
  if (P.nticks) {
    if (P.mode == MODE_STEP) { if (lane == 0) { P.nticks[2 * env] = ticks_step; P.nticks[2 * env + 1] = ticks_reset; } }
    else if (lane == 0) P.nticks[P.deferred_pass ? 2 * env + 1 : env] = P.mode == MODE_RESET ? ticks_reset : n;
  }
  if (lane == 0 && P.stats) {
    if (P.mode == MODE_STEP) atomicAdd(&P.stats[4], 1.0);
    atomicAdd(&P.stats[5], (double)tick_total);
    if (s.ovf) atomicAdd(&P.stats[6], (double)s.ovf);
  }
  store_state(s, rec, gp, step_count, episode, vert, ep_ret);
}

// ---- stand-alone kinematics kernels (one thread per query)
__global__ void fk_kernel(const DevModel* Mp, int n, const real* q, real* pose) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const DevModel& M = *Mp;
  const real* qi = q + (size_t)UG_NDOF * i;
  f3 P[UG_NLINK]; q4 Q[UG_NLINK];
  for (int l = 0; l < UG_NLINK; l++) {
    int p = M.link_parent[l];
    f3 pp = p >= 0 ? P[p] : mk3(0, 0, 0); q4 pq = p >= 0 ? Q[p] : mkq(0, 0, 0, 1);
    P[l] = pp + mulv(quat_to_mat(pq), ld3(M.link_jxyz[l]));
    q4 r = qmul(pq, mkq(M.link_jquat[l][0], M.link_jquat[l][1], M.link_jquat[l][2], M.link_jquat[l][3]));
    if (M.link_dof[l] >= 0) r = qmul(r, quat_axis(M.link_axis[l], qi[M.link_dof[l]]));
    Q[l] = qnormalize(r);
    real* o = pose + ((size_t)i * UG_NLINK + l) * 7;
    o[0] = P[l].x; o[1] = P[l].y; o[2] = P[l].z; o[3] = Q[l].x; o[4] = Q[l].y; o[5] = Q[l].z; o[6] = Q[l].w;
  }
}
__global__ void ik_kernel(const DevModel* Mp, int n, const real* q, const real* target, real* out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const real* t = target + 7 * (size_t)i;
  real o[6];
  ik_solve(q + (size_t)UG_NDOF * i, *Mp, mk3(t[0], t[1], t[2]), mkq(t[3], t[4], t[5], t[6]), o);
  for (int k = 0; k < 6; k++) out[6 * (size_t)i + k] = o[k];
}
// one warp per query (reuses the warp-level FK)
typedef SmemT<20, 60> SmemKin;
__global__ void __launch_bounds__(32) gripper_center_kernel(const DevModel* Mp, int n, const real* q, real* out) {
  extern __shared__ __align__(16) unsigned char smraw[];
  SmemKin& s = *reinterpret_cast<SmemKin*>(smraw);
  const Gp<32> gp = make_gp<32>();
  int i = blockIdx.x, lane = threadIdx.x;
  if (lane < UG_NDOF) s.q[lane] = q[(size_t)UG_NDOF * i + lane];
  if (lane < UG_MAXFREE) { s.fquat[lane][0] = 0; s.fquat[lane][1] = 0; s.fquat[lane][2] = 0; s.fquat[lane][3] = 1; }
  gp.sync();
  fk_bodies(s, *Mp, gp);
  f3 c = gripper_center(s, *Mp);
  if (lane == 0) { out[3 * (size_t)i] = c.x; out[3 * (size_t)i + 1] = c.y; out[3 * (size_t)i + 2] = c.z; }
}

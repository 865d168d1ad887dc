// TEST INFRASTRUCTURE — the CPU oracle's env layer and C ABI (see oracle/README.md). PARITY UNPINNED.
//
// Restates reset / step / wait_until_stable / reward / termination / observation of the five in-scope envs:
//   BallBalance      /root/reference/urgym/envs/env_ball_balance_v0.py:94-170
//   TwoBallsBalance  /root/reference/urgym/envs/env_two_balls_balance_v0.py:95-233
//   CubesPush        /root/reference/urgym/envs/env_cubes_push_v0.py:104-246
//   CubesGrasp       /root/reference/urgym/envs/env_cubes_grasp_v0.py:100-327
//   Golf             /root/reference/urgym/envs/env_golf_v0.py:104-358
//   Box              /root/reference/urgym/envs/env_box_v0.py:110-190 (fixed 120 ticks per step instead of wait_until_stable)
// and the robot helper /root/reference/urgym/base/robot.py:85-128,213-218. Episode limits from
// /root/reference/urgym/envs/__init__.py:10-44 (the TimeLimit wrapper's `truncated`).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
#include <algorithm>
#include <cstdio>
#include <initializer_list>
#include <string>

#include "urgo_engine.h"

namespace urgo {

static const double REST_POSE[6] = {-1.5690622952052096, -1.5446774605904932, 1.343946009733127,
                                    -1.3708613585093699, -1.5707970583733368, 0.0009377758247187636};  // robot.py:182-183
static const double PI = 3.14159265358979323846;

inline uint64_t splitmix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// Counter-based uniform in [0,1) with 24 bits, so fp32 and fp64 consumers see the same variate.
inline double urand(uint64_t seed, uint32_t env, uint32_t episode, uint32_t draw) {
  uint64_t h = splitmix(seed ^ splitmix(((uint64_t)env << 32) | episode) ^ splitmix(0xD1B54A32D192ED03ull * (uint64_t)(draw + 1)));
  return (double)(h >> 40) * (1.0 / 16777216.0);
}
inline double gripper_angle(double open_length) { return 0.715 - std::asin((open_length - 0.010) / 0.1143); }  // robot.py:215
// utilities.py:192-207
inline double geometric_distance_reward(double value, double ts, double tmax) {
  value = std::fmax(value, 1e-10);
  double f1 = (value - ts) / value, f2 = std::fmax(tmax - value, 1e-10);
  return -std::tanh(f1 / f2);
}

struct Env {
  World w;
  int env_type; uint64_t seed; uint32_t gidx;
  int step_count = 0; uint32_t episode = 0;
  bool approached = false, grasped = false;
  Q4 vertical;
  bool sparse = false;
  std::vector<std::pair<int, int>> last_contacts;
  long total_ticks = 0;
  // command trace for the env-logic parity tests: events of 9 doubles {code, v[8]} (tests/golden/make_golden.py)
  bool tracing = false; std::vector<double> trace;
  void ev(int code, std::initializer_list<double> v) {
    if (!tracing) return;
    size_t at = trace.size(); trace.resize(at + 9, 0.0); trace[at] = code;
    int k = 1; for (double x : v) if (k < 9) trace[at + k++] = x;
  }

  int variant;  // the registered env (ENV_PUSH_DIFF / ENV_GOLF_JOINTS differ from their scene in action / observation only)
  Env(int t, uint64_t s, uint32_t g) : w(t), env_type(scene_of(t)), seed(s), gidx(g), variant(t) {}

  static bool settled(double l1_motion) { return l1_motion < BP_SETTLE_THRESHOLD; }
  std::vector<double> settle_l1;  // summed joint motion of every tick of the last settle() (decision margins for the tests)
  int settle() {  // wait_until_stable, e.g. env_ball_balance_v0.py:94-103
    double prev[NDOF]; std::memcpy(prev, w.rob.q, sizeof(prev));
    int n = 0;
    settle_l1.clear();
    for (; n < BP_SETTLE_MAX_TICKS;) {
      w.tick(); n++;
      double s = 0; for (int d = 0; d < NDOF; d++) s += std::fabs(prev[d] - w.rob.q[d]);
      settle_l1.push_back(s);
      if (settled(s)) break;
      std::memcpy(prev, w.rob.q, sizeof(prev));
    }
    total_ticks += n;
    ev(EV_SETTLE, {(double)n});
    return n;
  }
  enum { EV_IK = 1, EV_GRIPPER = 2, EV_SPAWN = 3, EV_TICKS = 4, EV_SETTLE = 5, EV_JOINTS = 6, EV_RESET_ARM = 7 };
  void move_ee_end(V3 pos, Q4 q) {
    ev(EV_IK, {pos.x, pos.y, pos.z, q.x, q.y, q.z, q.w});
    double t[6]; w.ik(pos, q, t);
    for (int i = 0; i < 6; i++) w.target[i] = t[i];
    w.arm_armed = true;
  }
  void move_ee_joints(const double* a) {
    ev(EV_JOINTS, {a[0], a[1], a[2], a[3], a[4], a[5]});
    for (int i = 0; i < 6; i++) w.target[i] = a[i];
    w.arm_armed = true;
  }
  void move_gripper(double len) { w.target[6] = gripper_angle(len); w.grip_armed = true; ev(EV_GRIPPER, {w.target[6], len}); }
  void ee_pose(V3& p, Q4& q) { w.rob.fk(); p = w.rob.p[7]; q = w.rob.quat[7]; }
  void spawn(int b, V3 pos, Q4 q) {
    ev(EV_SPAWN, {(double)b, pos.x, pos.y, pos.z, q.x, q.y, q.z, q.w});
    FreeBody& B = w.fb[b]; B.pos = pos; B.quat = q; B.lin = V3(); B.ang = V3(); B.active = true;
  }
  double draw(const double* ov, int k) { return ov ? ov[k] : urand(seed, gidx, episode, (uint32_t)k); }

  bool touch(bool (*pred)(const ShapeDef&, const ShapeDef&, void*), void* ctx = nullptr) const {
    for (auto& c : w.contacts) if (pred(w.sc.shapes[c.sa], w.sc.shapes[c.sb], ctx)) return true;
    return false;
  }
  double opening() {  // env_cubes_grasp_v0.py:340-377
    w.rob.fk();
    double v = norm(w.rob.p[16] - w.rob.p[11]) - 0.052;
    v = std::fmax(0.0, std::fmin(0.085, v));
    return v / 0.085;
  }
  V3 club_foot() const { return w.fb[0].pos + quat_to_mat(w.fb[0].quat) * V3(0, 0, -0.09 - CLUB_ZC); }  // env_golf_v0.py:368-373

  int reset(const double* ov, double* obs) {
    episode++; step_count = 0; approached = grasped = false;
    int ticks = 0;
    // robot.reset() runs first, with the previous episode's objects still in the world (e.g. env_ball_balance_v0.py:147-152,
    // env_cubes_push_v0.py:214-215 + 238-241, env_golf_v0.py:240-248): they are removed after the ten arm ticks
    for (int i = 0; i < 6; i++) { w.rob.q[i] = REST_POSE[i]; w.rob.qd[i] = 0; }   // robot.py:93-94
    ev(EV_RESET_ARM, {REST_POSE[0], REST_POSE[1], REST_POSE[2], REST_POSE[3], REST_POSE[4], REST_POSE[5]});
    for (int i = 0; i < BP_RESET_ARM_TICKS; i++) { w.tick(); ticks++; total_ticks++; }  // robot.py:97-98
    ev(EV_TICKS, {(double)BP_RESET_ARM_TICKS});
    move_gripper(0.085);                                                            // robot.py:100-101
    for (int b = 0; b < MAXFREE; b++) w.fb[b].active = false;                       // removeBody of the old task objects
    V3 ep; Q4 eq;
    switch (env_type) {
      case ENV_BALL: case ENV_TWO: {
        move_ee_end(V3(0.0, -0.60, 0.60), Q4(0.50, -0.50, -0.50, 0.50));
        move_gripper(0.085 * 0.2);
        ticks += settle();
        ee_pose(ep, eq);
        V3 gc = w.gripper_center();
        V3 pp(ep.x, ep.y - 0.2, gc.z);
        spawn(0, pp, Q4());
        if (env_type == ENV_BALL) {
          spawn(1, pp + V3(-0.03 + 0.06 * draw(ov, 0), -0.08 + 0.10 * draw(ov, 1), 0.2 + 0.02), Q4());
        } else {
          double r = 0.003;
          spawn(1, pp + V3(-0.03 + 0.06 * draw(ov, 0), -0.08 + 0.035 * draw(ov, 1), 0.2 + r), Q4());
          spawn(2, pp + V3(-0.03 + 0.06 * draw(ov, 2), -0.015 + 0.035 * draw(ov, 3), 0.2 - r), Q4());
        }
        move_gripper(0.0);
        ticks += settle();
        break;
      }
      case ENV_PUSH: {
        spawn(0, V3(0.0, -0.5, 0.1), Q4());
        V3 c1(-0.1 + 0.2 * draw(ov, 0), -0.7, 0.1);
        spawn(1, c1, Q4());
        vertical = Q4() * axis_angle(V3(0, 1, 0), PI / 2);  // utilities.py:134-149
        move_ee_end(c1 + V3(0, 0, 0.2), vertical);
        move_gripper(0.0);
        ticks += settle();
        break;
      }
      case ENV_GRASP: {
        double r = 0.5 + 0.3 * draw(ov, 0), th = -PI + PI * draw(ov, 1);
        spawn(0, V3(r * std::cos(th), r * std::sin(th), 0.1), Q4());
        ee_pose(ep, eq);
        move_ee_end(ep, eq);
        ticks += settle();
        break;
      }
      case ENV_BOX: {  // env_box_v0.py:178-185: move_ee to the current pose, reset_box(): both joints to 0, button motor on
        ee_pose(ep, eq);
        move_ee_end(ep, eq);
        for (int b = 0; b < 2; b++) spawn(b, w.sc.anchor[b].origin, Q4());
        break;
      }
      case ENV_GOLF: {
        vertical = Q4(0.0, 0.0, -0.70, 0.70);
        move_ee_end(V3(0.0, -0.60, 0.40), vertical);
        move_gripper(0.085 * 0.2);
        ticks += settle();
        V3 gc = w.gripper_center();
        spawn(0, gc + V3(0, 0, -0.05 + CLUB_ZC), Q4());
        spawn(1, V3(0.0 + (-0.2 + 0.4 * draw(ov, 0)), -0.7 + (-0.1 + 0.2 * draw(ov, 1)), 0.1), Q4());
        move_gripper(0.0);
        ticks += settle();
        break;
      }
    }
    observe(obs);
    return ticks;
  }

  void observe(double* o) {
    V3 ep; Q4 eq;
    switch (env_type) {
      case ENV_BALL: case ENV_TWO: {
        V3 e = euler_from_quat(w.fb[0].quat);
        o[0] = e.x; o[1] = e.y; o[2] = e.z;
        for (int b = 1; b < w.sc.nfree; b++) { V3 d = w.fb[0].pos - w.fb[b].pos; o[3 * b] = d.x; o[3 * b + 1] = d.y; o[3 * b + 2] = d.z; }
        break;
      }
      case ENV_PUSH: {
        ee_pose(ep, eq);
        o[0] = ep.x; o[1] = ep.y; o[2] = ep.z;
        V3 ref = variant == ENV_PUSH_DIFF ? ep : V3();   // env_cubes_push_diff_v0.py:209-211
        for (int b = 0; b < 2; b++) { V3 d = w.fb[b].pos - ref; o[3 + 3 * b] = d.x; o[4 + 3 * b] = d.y; o[5 + 3 * b] = d.z; }
        break;
      }
      case ENV_GRASP:
        ee_pose(ep, eq);
        o[0] = approached; o[1] = grasped;
        o[2] = ep.x; o[3] = ep.y; o[4] = ep.z; o[5] = eq.x; o[6] = eq.y; o[7] = eq.z; o[8] = eq.w;
        o[9] = opening();
        o[10] = w.fb[0].pos.x; o[11] = w.fb[0].pos.y; o[12] = w.fb[0].pos.z;
        o[13] = w.fb[0].quat.x; o[14] = w.fb[0].quat.y; o[15] = w.fb[0].quat.z; o[16] = w.fb[0].quat.w;
        break;
      case ENV_BOX:  // env_box_v0.py:161-170: robot.get_joint_obs()['ee_pos'] = ee position + quaternion
        ee_pose(ep, eq);
        o[0] = ep.x; o[1] = ep.y; o[2] = ep.z; o[3] = eq.x; o[4] = eq.y; o[5] = eq.z; o[6] = eq.w;
        break;
      case ENV_GOLF: {
        V3 s = club_foot(), b = w.fb[1].pos, h(0, -0.6, 0.02);
        if (variant == ENV_GOLF_JOINTS) { for (int i = 0; i < 6; i++) o[i] = w.rob.q[i]; o += 6; }  // env_golf_joints_v0.py:229-234
        o[0] = s.x; o[1] = s.y; o[2] = s.z;
        V3 d1 = s - b, d2 = h - b;
        o[3] = d1.x; o[4] = d1.y; o[5] = d1.z; o[6] = d2.x; o[7] = d2.y; o[8] = d2.z;
        break;
      }
    }
  }

  // The part of step() before wait_until_stable: action -> motor targets. Returns the reward accrued so far.
  double prologue(const double* a, double u) {
    V3 ep; Q4 eq;
    double rew = 0;
    switch (env_type) {
      case ENV_BALL: case ENV_TWO: {
        if (env_type == ENV_TWO) for (int i = 0; i < 3; i++) rew -= std::fabs(a[i]);
        ee_pose(ep, eq);
        V3 e = euler_from_quat(eq);
        move_ee_end(ep, quat_from_euler(e.x + a[0], e.y + a[1], e.z + a[2]));
        break;
      }
      case ENV_PUSH: case ENV_GOLF:
        if (variant == ENV_GOLF_JOINTS) {  // env_golf_joints_v0.py:123-124
          double t[6]; for (int i = 0; i < 6; i++) t[i] = w.rob.q[i] + a[i];
          move_ee_joints(t);
          break;
        }
        ee_pose(ep, eq);
        move_ee_end(ep + V3(a[0], a[1], a[2]), vertical);
        break;
      case ENV_BOX: {  // env_box_v0.py:124-126: new_pose = ee_pose (position + quaternion, 7 numbers) + action[:-1], component-wise
        ee_pose(ep, eq);
        move_ee_end(ep + V3(a[0], a[1], a[2]), Q4(eq.x + a[3], eq.y + a[4], eq.z + a[5], eq.w + a[6]));
        move_gripper(a[7]);
        break;
      }
      case ENV_GRASP: {
        double a0 = a[0], a1 = a[1];
        int sel = (a0 + a1 == 0) ? (u < 0.5 ? 0 : 1) : (u * (a0 + a1) < a0 ? 0 : 1);
        if (sel == 0) move_ee_joints(a + 2);
        else move_gripper(a[8] < 0.5 ? 0.085 : 0.0);
        break;
      }
    }
    return rew;
  }

  // The part of step() after wait_until_stable: contact predicates, reward, termination, observation.
  void epilogue(double rew, double* obs, double* reward, uint8_t* terminated, uint8_t* truncated, uint8_t* success) {
    bool term = false, succ = false;
    switch (env_type) {
      case ENV_BALL: case ENV_TWO: {
        rew += 1;
        if (env_type == ENV_BALL) {
          term = w.fb[1].pos.z < w.fb[0].pos.z;
        } else {
          for (int b = 1; b <= 2; b++) for (int k = 0; k < 2; k++) { double x = std::fabs(w.fb[b].pos[k]); rew += x > 1 ? 0.0 : 1 - x * x; }
          term = w.fb[1].pos.z < w.fb[0].pos.z || w.fb[2].pos.z < w.fb[0].pos.z;
        }
        break;
      }
      case ENV_PUSH: {
        // getContactPoints(robot, table) (env_cubes_push_v0.py:176-183): the table only, not the ground plane under it
        bool table = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return A.body == 0 && B.body < 0 && B.link == 0 && A.link != 0; });
        bool fingers = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return A.body == 0 && B.body == 2 && A.link >= 9 && A.link <= 18; });
        if (table) { rew -= 0.01; }
        else if (fingers) {
          double d = norm(w.fb[0].pos - w.fb[1].pos);
          rew = 0;
          if (!sparse) rew += geometric_distance_reward(d, 0.1, 0.5) / 10;
          if (d < 0.05) { rew += 10; succ = true; }
          rew += 0.1; term = succ;
        }
        rew -= 0.1;
        break;
      }
      case ENV_GRASP: {
        V3 gc = w.gripper_center();
        double d = norm(gc - w.fb[0].pos);
        bool l = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return A.body == 0 && B.body == 1 && A.link >= 11 && A.link <= 13; });
        bool r = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return A.body == 0 && B.body == 1 && A.link >= 16 && A.link <= 18; });
        bool grasp_now = l && r;
        if (!approached) {
          rew += geometric_distance_reward(d, 0.5, 2) / 4;
          if (d < 0.05 && opening() > 0.50) { approached = true; rew += 5; }
        } else if (!grasped) {
          if (grasp_now) { grasped = true; rew += 5; }
        } else {
          if (grasp_now && w.fb[0].pos.z > 0.2) { rew += 10; term = true; succ = true; }
        }
        rew -= 0.1;
        break;
      }
      case ENV_BOX: {  // env_box_v0.py:141-159 (button_touch_mode='any', the registered default), :129-137
        const double door = w.anchor_q(0), button = w.anchor_q(1);
        if (!approached) { if (door > 1.9) { approached = true; rew = 1; } }          // box_opened
        else if (!grasped) { if (button < -0.02) { grasped = true; rew = 1; } }       // button_pressed
        else if (door < 0.1) { rew = 1; term = true; succ = true; }                   // box_closed
        rew -= 0.1;
        break;
      }
      case ENV_GOLF: {
        V3 hole(0, -0.6, 0.02), ball = w.fb[1].pos;
        bool in_fingers = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return A.body == 0 && B.body == 1 && A.link >= 9 && A.link <= 18; });
        if (norm(hole - ball) < 0.05) { rew += 10; succ = true; term = true; }
        else if (!in_fingers) { rew -= 5; term = true; }
        else if (ball.z < 0) { rew -= 10; term = true; }
        else {
          bool floor_hit = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return (A.body == 0 || A.body == 1) && B.body < 0 && B.link == 1; });
          if (floor_hit) rew -= 0.01;
          if (!sparse) rew += geometric_distance_reward(norm(club_foot() - ball), 0.1, 0.5) / 10;
          bool hit = touch([](const ShapeDef& A, const ShapeDef& B, void*) { return A.body == 1 && B.body == 2; });
          if (hit) { rew += 0.1; if (!sparse) rew += geometric_distance_reward(norm(hole - ball), 0.1, 0.5) / 10; }
        }
        rew -= 0.1;
        break;
      }
    }
    last_contacts.clear();
    for (auto& c : w.contacts) last_contacts.push_back({c.sa, c.sb});
    step_count++;
    *reward = rew; *terminated = term; *success = succ; *truncated = step_count >= w.sc.max_steps;
    observe(obs);
  }

  // One gym step without auto-reset. Returns settle ticks.
  int step(const double* a, double u, double* obs, double* reward, uint8_t* terminated, uint8_t* truncated, uint8_t* success) {
    double rew = prologue(a, u);
    int n;
    if (env_type == ENV_BOX) { for (n = 0; n < 120; n++) w.tick(); total_ticks += n; ev(EV_TICKS, {(double)n}); }  // env_box_v0.py:127-128
    else n = settle();
    epilogue(rew, obs, reward, terminated, truncated, success);
    return n;
  }

  void get_state(double* s) const {
    std::memset(s, 0, sizeof(double) * STATE_SLOTS);
    for (int d = 0; d < NDOF; d++) { s[d] = w.rob.q[d]; s[12 + d] = w.rob.qd[d]; }
    for (int m = 0; m < NMOTOR; m++) s[24 + m] = w.target[m];
    int flags = (w.arm_armed ? 1 : 0) | (w.grip_armed ? 2 : 0) | (approached ? 4 : 0) | (grasped ? 8 : 0);
    for (int b = 0; b < MAXFREE; b++) if (w.fb[b].active) flags |= 16 << b;
    s[31] = flags; s[32] = step_count; s[33] = episode;
    s[34] = vertical.x; s[35] = vertical.y; s[36] = vertical.z; s[37] = vertical.w;
    for (int b = 0; b < MAXFREE; b++) {
      const FreeBody& B = w.fb[b]; double* p = s + 38 + 13 * b;
      p[0] = B.pos.x; p[1] = B.pos.y; p[2] = B.pos.z; p[3] = B.quat.x; p[4] = B.quat.y; p[5] = B.quat.z; p[6] = B.quat.w;
      p[7] = B.lin.x; p[8] = B.lin.y; p[9] = B.lin.z; p[10] = B.ang.x; p[11] = B.ang.y; p[12] = B.ang.z;
    }
  }
  void set_state(const double* s) {
    for (int d = 0; d < NDOF; d++) { w.rob.q[d] = s[d]; w.rob.qd[d] = s[12 + d]; }
    for (int m = 0; m < NMOTOR; m++) w.target[m] = s[24 + m];
    int flags = (int)s[31];
    w.arm_armed = flags & 1; w.grip_armed = flags & 2; approached = flags & 4; grasped = flags & 8;
    step_count = (int)s[32]; episode = (uint32_t)s[33];
    vertical = Q4(s[34], s[35], s[36], s[37]);
    for (int b = 0; b < MAXFREE; b++) {
      FreeBody& B = w.fb[b]; const double* p = s + 38 + 13 * b;
      B.active = (flags >> (4 + b)) & 1;
      B.pos = V3(p[0], p[1], p[2]); B.quat = Q4(p[3], p[4], p[5], p[6]); B.lin = V3(p[7], p[8], p[9]); B.ang = V3(p[10], p[11], p[12]);
    }
  }
};

struct Batch { std::vector<Env> envs; int env_type; };

}  // namespace urgo

using namespace urgo;

extern "C" {

void* urgo_create(int env_type, int num_envs, uint64_t seed, int env_index_offset) {
  if (env_type < 0 || env_type >= ENV_COUNT || num_envs <= 0) return nullptr;
  Batch* b = new Batch; b->env_type = env_type;
  b->envs.reserve(num_envs);
  for (int i = 0; i < num_envs; i++) b->envs.emplace_back(env_type, seed, (uint32_t)(env_index_offset + i));
  return b;
}
void urgo_destroy(void* h) { delete (Batch*)h; }
int urgo_obs_dim(int t) { return make_scene(t).obs_dim; }
int urgo_act_dim(int t) { return make_scene(t).act_dim; }
int urgo_max_steps(int t) { return make_scene(t).max_steps; }
int urgo_num_shapes(int t) { return (int)make_scene(t).shapes.size(); }
// shape descriptor: body, link
void urgo_shape_info(int t, int idx, int* body, int* link) { Scene s = make_scene(t); *body = s.shapes[idx].body; *link = s.shapes[idx].link; }
void urgo_set_sparse(void* h, int sparse) { for (auto& e : ((Batch*)h)->envs) e.sparse = sparse != 0; }

// env range [lo, hi) so that callers can spread envs over host threads (envs are independent)
int urgo_reset_range(void* h, int lo, int hi, const uint8_t* mask, const double* draws, double* obs, int32_t* nticks) {
  Batch* b = (Batch*)h; int od = make_scene(b->env_type).obs_dim;
  for (int i = lo; i < hi; i++) {
    if (mask && !mask[i]) continue;
    int n = b->envs[i].reset(draws ? draws + 4 * i : nullptr, obs + (size_t)od * i);
    if (nticks) nticks[i] = n;
  }
  return 0;
}
int urgo_reset(void* h, const uint8_t* mask, const double* draws, double* obs, int32_t* nticks) {
  return urgo_reset_range(h, 0, (int)((Batch*)h)->envs.size(), mask, draws, obs, nticks);
}
int urgo_step_range(void* h, int lo, int hi, const double* action, const double* u, const double* draws, double* obs, double* reward,
                    uint8_t* terminated, uint8_t* truncated, uint8_t* success, double* final_obs, int32_t* nticks, int autoreset) {
  Batch* b = (Batch*)h; Scene sc = make_scene(b->env_type); int od = sc.obs_dim, ad = sc.act_dim;
  for (int i = lo; i < hi; i++) {
    Env& e = b->envs[i];
    double uu = u ? u[i] : urand(e.seed, e.gidx, e.episode, 16u + (uint32_t)e.step_count);
    int n = e.step(action + (size_t)ad * i, uu, obs + (size_t)od * i, reward + i, terminated + i, truncated + i, success + i);
    int nr = 0;
    if (final_obs) std::memcpy(final_obs + (size_t)od * i, obs + (size_t)od * i, sizeof(double) * od);
    if (autoreset && (terminated[i] || truncated[i])) nr = e.reset(draws ? draws + 4 * i : nullptr, obs + (size_t)od * i);
    if (nticks) { nticks[2 * i] = n; nticks[2 * i + 1] = nr; }
  }
  return 0;
}
int urgo_step(void* h, const double* action, const double* u, const double* draws, double* obs, double* reward, uint8_t* terminated,
              uint8_t* truncated, uint8_t* success, double* final_obs, int32_t* nticks, int autoreset) {
  return urgo_step_range(h, 0, (int)((Batch*)h)->envs.size(), action, u, draws, obs, reward, terminated, truncated, success, final_obs, nticks, autoreset);
}
int urgo_get_state(void* h, double* out) { Batch* b = (Batch*)h; for (size_t i = 0; i < b->envs.size(); i++) b->envs[i].get_state(out + STATE_SLOTS * i); return 0; }
int urgo_set_state(void* h, const double* in) { Batch* b = (Batch*)h; for (size_t i = 0; i < b->envs.size(); i++) b->envs[i].set_state(in + STATE_SLOTS * i); return 0; }
// pairs[N][EXPORT_MAXC][2] (-1 padded), count[N] = the true number of manifold points (may exceed EXPORT_MAXC)
int urgo_contacts(void* h, int32_t* pairs, int32_t* count) {
  Batch* b = (Batch*)h;
  for (size_t i = 0; i < b->envs.size(); i++) {
    const Env& e = b->envs[i]; const int n = (int)e.last_contacts.size(); count[i] = n;
    for (int k = 0; k < EXPORT_MAXC; k++) { pairs[(i * EXPORT_MAXC + k) * 2] = k < n ? e.last_contacts[k].first : -1; pairs[(i * EXPORT_MAXC + k) * 2 + 1] = k < n ? e.last_contacts[k].second : -1; }
  }
  return 0;
}
int urgo_export_maxc(void) { return EXPORT_MAXC; }
// ---- env-logic entry points (no physics): the two halves of step() around wait_until_stable, and the command trace
// out[16] = reward so far, motor targets[7], arm_armed, grip_armed
int urgo_prologue(void* h, int env, const double* action, double u, double* out) {
  Env& e = ((Batch*)h)->envs[env];
  out[0] = e.prologue(action, u);
  for (int m = 0; m < NMOTOR; m++) out[1 + m] = e.w.target[m];
  out[8] = e.w.arm_armed; out[9] = e.w.grip_armed;
  return 0;
}
// pairs[npairs][2] = shape indices in contact after the (skipped) settle
int urgo_epilogue(void* h, int env, double reward_so_far, int npairs, const int32_t* pairs, double* obs, double* reward, uint8_t* terminated,
                  uint8_t* truncated, uint8_t* success) {
  Env& e = ((Batch*)h)->envs[env];
  e.w.contacts.clear();
  for (int i = 0; i < npairs; i++) { Contact c; c.sa = pairs[2 * i]; c.sb = pairs[2 * i + 1]; e.w.contacts.push_back(c); }
  e.epilogue(reward_so_far, obs, reward, terminated, truncated, success);
  return 0;
}
void urgo_set_trace(void* h, int on) { for (auto& e : ((Batch*)h)->envs) { e.tracing = on != 0; e.trace.clear(); } }
int urgo_trace(void* h, int env, double* out, int max_events) {
  Env& e = ((Batch*)h)->envs[env]; int n = (int)(e.trace.size() / 9);
  for (int i = 0; i < n && i < max_events; i++) std::memcpy(out + 9 * i, e.trace.data() + 9 * i, sizeof(double) * 9);
  return n;
}
// world-frame kinematic quantities of the current state that the env logic reads: ee pos3 quat4, link 11 pos3, link 16 pos3,
// gripper centre3, club foot3 (Golf only)
void urgo_kinematics(void* h, int env, double* out19) {
  Env& e = ((Batch*)h)->envs[env]; V3 p; Q4 q; e.ee_pose(p, q);
  double* o = out19; o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = q.x; o[4] = q.y; o[5] = q.z; o[6] = q.w;
  V3 a = e.w.rob.p[11], b = e.w.rob.p[16], c = e.w.gripper_center();
  o[7] = a.x; o[8] = a.y; o[9] = a.z; o[10] = b.x; o[11] = b.y; o[12] = b.z; o[13] = c.x; o[14] = c.y; o[15] = c.z;
  V3 f = e.env_type == ENV_GOLF ? e.club_foot() : V3(); o[16] = f.x; o[17] = f.y; o[18] = f.z;
}
// the settle rule alone: ticks wait_until_stable takes when the summed |joint motion| of tick k is l1[min(k, n-1)]
int urgo_settle_count(const double* l1, int n) {
  int k = 0;
  for (; k < BP_SETTLE_MAX_TICKS;) { double s = l1[k < n ? k : n - 1]; k++; if (Env::settled(s)) break; }
  return k;
}
// static model facts the env-logic tests compare with the reference's robot set-up
// out: dof_link[12], effort[12], velocity[12], lower[12], upper[12], gear_child_link[5], gear_ratio[5], rest_pose[6],
//      reset_arm_ticks, settle_max_ticks, settle_threshold, club_zc, ik_iters  (= 75 doubles)
void urgo_model_info(double* o) {
  int k = 0, dof_link[NDOF], nd = 0;
  for (int i = 0; i < URG_NLINK; i++) if (URG_LINKS[i].jtype == 1) dof_link[nd++] = i;
  for (int d = 0; d < NDOF; d++) o[k++] = dof_link[d];
  for (int d = 0; d < NDOF; d++) o[k++] = URG_LINKS[dof_link[d]].effort;
  for (int d = 0; d < NDOF; d++) o[k++] = URG_LINKS[dof_link[d]].velocity;
  for (int d = 0; d < NDOF; d++) o[k++] = URG_LINKS[dof_link[d]].lower;
  for (int d = 0; d < NDOF; d++) o[k++] = URG_LINKS[dof_link[d]].upper;
  for (int g = 0; g < 5; g++) o[k++] = dof_link[GEAR_DOF[g]];
  for (int g = 0; g < 5; g++) o[k++] = GEAR_RATIO[g];
  for (int i = 0; i < 6; i++) o[k++] = REST_POSE[i];
  o[k++] = BP_RESET_ARM_TICKS; o[k++] = BP_SETTLE_MAX_TICKS; o[k++] = BP_SETTLE_THRESHOLD; o[k++] = CLUB_ZC; o[k++] = BP_IK_ITERS;
}
// scene facts: out[0]=gravity z, out[1..3]=free-body masses; per shape (urgo_shape_detail): kind (0 box, 1 sphere),
// half3 or radius, local centre3, mu, restitution, rolling friction
void urgo_scene_info(int t, double* out4) { Scene s = make_scene(t); out4[0] = s.gravity; for (int b = 0; b < MAXFREE; b++) out4[1 + b] = b < s.nfree ? s.fmass[b] : 0.0; }
void urgo_shape_detail(int t, int idx, double* o) {
  Scene s = make_scene(t); const ShapeDef& d = s.shapes[idx];
  o[0] = d.kind; o[1] = d.kind == 0 ? d.half.x : d.radius; o[2] = d.half.y; o[3] = d.half.z; o[4] = d.lc.x; o[5] = d.lc.y; o[6] = d.lc.z;
  o[7] = d.mu; o[8] = d.rest; o[9] = d.roll;
}
long urgo_total_ticks(void* h) { long s = 0; for (auto& e : ((Batch*)h)->envs) s += e.total_ticks; return s; }

// ---- unit-level entry points for parity tests
// n ticks of raw physics on every env (no env logic); detail outputs of the last tick are optional
int urgo_tick(void* h, int nticks) { for (auto& e : ((Batch*)h)->envs) for (int k = 0; k < nticks; k++) e.w.tick(); return 0; }
int urgo_last_contacts_detail(void* h, int env, double* out /*[EXPORT_MAXC][11]: sa, sb, posB3, n3, dist, mu, roll*/) {
  Env& e = ((Batch*)h)->envs[env]; int n = std::min((int)e.w.contacts.size(), (int)EXPORT_MAXC);
  for (int i = 0; i < n; i++) {
    const Contact& c = e.w.contacts[i]; double* o = out + 11 * i;
    o[0] = c.sa; o[1] = c.sb; o[2] = c.posB.x; o[3] = c.posB.y; o[4] = c.posB.z; o[5] = c.n.x; o[6] = c.n.y; o[7] = c.n.z; o[8] = c.dist; o[9] = c.mu; o[10] = c.roll;
  }
  return n;
}
// summed profiling counters over all envs: ticks, solver iterations, rows, contacts, rows*iterations
void urgo_tick_stats(void* h, long* out5) {
  for (int k = 0; k < 5; k++) out5[k] = 0;
  for (auto& e : ((Batch*)h)->envs) { out5[0] += e.w.st_ticks; out5[1] += e.w.st_iters; out5[2] += e.w.st_rows; out5[3] += e.w.st_contacts; out5[4] += e.w.st_rowiters; }
}
// histograms over ticks: solver rows [192], manifold points [64], joint rows [32], rolling-friction manifolds [16]
void urgo_tick_hist(void* h, long* rows192, long* nc64, long* nnc32, long* roll16) {
  for (int k = 0; k < 192; k++) rows192[k] = 0;
  for (int k = 0; k < 64; k++) nc64[k] = 0;
  for (int k = 0; k < 32; k++) nnc32[k] = 0;
  for (int k = 0; k < 16; k++) roll16[k] = 0;
  for (auto& e : ((Batch*)h)->envs) {
    for (int k = 0; k < 192; k++) rows192[k] += e.w.st_hist_rows[k];
    for (int k = 0; k < 64; k++) nc64[k] += e.w.st_hist_nc[k];
    for (int k = 0; k < 32; k++) nnc32[k] += e.w.st_hist_nnc[k];
    for (int k = 0; k < 16; k++) roll16[k] += e.w.st_hist_roll[k];
  }
}
// run-time copy of a recalled parameter (bullet_params.h: BulletParams); affects worlds created afterwards. Returns 0 or -1.
int urgo_set_param(const char* name, double v) {
  BulletParams& P = bp(); std::string n(name);
#define URGO_P(f) if (n == #f) { P.f = v; return 0; }
  URGO_P(erp) URGO_P(erp2) URGO_P(linear_slop) URGO_P(restitution_threshold) URGO_P(lin_damping) URGO_P(ang_damping) URGO_P(contact_breaking)
  URGO_P(motor_kp) URGO_P(motor_kd) URGO_P(ik_damping) URGO_P(max_friction) URGO_P(solver_iters) URGO_P(robot_mu) URGO_P(body_mu) URGO_P(cube_mu)
  URGO_P(plane_mu) URGO_P(cube_mass) URGO_P(friction_cone) URGO_P(roll_points) URGO_P(link_mask) URGO_P(box_shrink) URGO_P(hull_links) URGO_P(disc_half)
#undef URGO_P
  return -1;
}
void urgo_reset_params(void) { bp() = BulletParams(); }
// summed joint motion of each tick of the env's last wait_until_stable (out[max]); returns the tick count
int urgo_settle_trace(void* h, int env, double* out, int max) {
  Env& e = ((Batch*)h)->envs[env]; int n = (int)e.settle_l1.size();
  for (int i = 0; i < n && i < max; i++) out[i] = e.settle_l1[i];
  return n;
}
// how close the last tick's contact set came to being another one (see decision_margin in urgo_collide.h)
double urgo_contact_margin(void* h, int env) { return ((Batch*)h)->envs[env].w.contact_margin; }
int urgo_last_iters(void* h, int env) { return ((Batch*)h)->envs[env].w.last_iters; }
// link frames for joint vector q[12]: out[19][7] = pos xyz, quat xyzw (getLinkState(...)[4], [5])
void urgo_fk(const double* q, double* out) {
  Robot r; std::memcpy(r.q, q, sizeof(double) * NDOF); r.fk();
  for (int i = 0; i < URG_NLINK; i++) { double* o = out + 7 * i; o[0] = r.p[i].x; o[1] = r.p[i].y; o[2] = r.p[i].z; o[3] = r.quat[i].x; o[4] = r.quat[i].y; o[5] = r.quat[i].z; o[6] = r.quat[i].w; }
}
void urgo_ik(const double* q, const double* target7, double* out6) {
  World w(ENV_BALL); std::memcpy(w.rob.q, q, sizeof(double) * NDOF);
  w.ik(V3(target7[0], target7[1], target7[2]), Q4(target7[3], target7[4], target7[5], target7[6]), out6);
}
// the same with the robot's base link at `base3` (Box-v0: (0, 0.5, 0), env_box_v0.py:55)
void urgo_ik_base(const double* q, const double* target7, const double* base3, double* out6) {
  World w(ENV_BALL); std::memcpy(w.rob.q, q, sizeof(double) * NDOF); w.rob.base = V3(base3[0], base3[1], base3[2]);
  w.ik(V3(target7[0], target7[1], target7[2]), Q4(target7[3], target7[4], target7[5], target7[6]), out6);
}
void urgo_gripper_center(const double* q, double* out3) {
  World w(ENV_BALL); std::memcpy(w.rob.q, q, sizeof(double) * NDOF);
  V3 c = w.gripper_center(); out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}
// free (unconstrained) joint accelerations by the 19-link ABA and M^-1 tau via the cached articulated inertias
void urgo_forward_dynamics(const double* q, const double* qd, double gz, double* qdd) {
  Robot r; std::memcpy(r.q, q, sizeof(double) * NDOF); std::memcpy(r.qd, qd, sizeof(double) * NDOF); r.fk(); r.aba(gz, qdd);
}
void urgo_minv(const double* q, double* out144) {
  Robot r; std::memcpy(r.q, q, sizeof(double) * NDOF); r.fk(); double qdd[NDOF]; r.aba(0.0, qdd);
  for (int d = 0; d < NDOF; d++) { double tau[NDOF] = {0}; tau[d] = 1; double col[NDOF]; r.minv_mul(tau, col); for (int k = 0; k < NDOF; k++) out144[k * NDOF + d] = col[k]; }
}
void urgo_euler_from_quat(const double* q, double* e) { V3 v = euler_from_quat(Q4(q[0], q[1], q[2], q[3])); e[0] = v.x; e[1] = v.y; e[2] = v.z; }
void urgo_quat_from_euler(const double* e, double* q) { Q4 v = quat_from_euler(e[0], e[1], e[2]); q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = v.w; }
void urgo_quat_mul(const double* a, const double* b, double* o) { Q4 v = Q4(a[0], a[1], a[2], a[3]) * Q4(b[0], b[1], b[2], b[3]); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
double urgo_geometric_distance_reward(double v, double ts, double tm) { return geometric_distance_reward(v, ts, tm); }
double urgo_gripper_angle(double len) { return gripper_angle(len); }
double urgo_urand(uint64_t seed, uint32_t env, uint32_t episode, uint32_t draw) { return urand(seed, env, episode, draw); }
// box-box narrowphase: a,b = [c3, R9 row-major, h3]; out = [n][10] posA3 posB3 n3 dist
int urgo_box_box(const double* a, const double* b, double margin, double* out) {
  BoxW A, B; A.c = V3(a[0], a[1], a[2]); for (int i = 0; i < 9; i++) A.R.m[i] = a[3 + i]; A.h = V3(a[12], a[13], a[14]);
  B.c = V3(b[0], b[1], b[2]); for (int i = 0; i < 9; i++) B.R.m[i] = b[3 + i]; B.h = V3(b[12], b[13], b[14]);
  CPoint p[4]; int n = box_box(A, B, margin, p);
  for (int i = 0; i < n; i++) { double* o = out + 10 * i; o[0] = p[i].posA.x; o[1] = p[i].posA.y; o[2] = p[i].posA.z; o[3] = p[i].posB.x; o[4] = p[i].posB.y; o[5] = p[i].posB.z; o[6] = p[i].n.x; o[7] = p[i].n.y; o[8] = p[i].n.z; o[9] = p[i].dist; }
  return n;
}

}  // extern "C"

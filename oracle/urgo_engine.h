// TEST INFRASTRUCTURE — part of the CPU oracle (see oracle/README.md). PARITY UNPINNED (see bullet_params.h).
//
// fp64, single-thread restatement of what one pybullet world does for one URGym env instance: the UR5 + Robotiq-85
// multibody (19 links, fixed joints kept as zero-DoF links as pybullet does when URDF_MERGE_FIXED_LINKS is not
// passed, /root/reference/urgym/base/robot.py:186-187), free rigid bodies, primitive collision, Bullet-style
// velocity-space projected Gauss-Seidel, and damped-least-squares IK. SURVEY.md §8a rows P1-P6, A3, A5, A7.
#pragma once
#include <cstdint>
#include <cstring>
#include <vector>

#include "bullet_params.h"
#include "model_data.h"
#include "urgo_collide.h"
#include "urgo_math.h"

namespace urgo {

enum { ENV_BALL = 0, ENV_TWO = 1, ENV_PUSH = 2, ENV_GRASP = 3, ENV_GOLF = 4, ENV_PUSH_DIFF = 5, ENV_GOLF_JOINTS = 6, ENV_BOX = 7, ENV_COUNT = 8 };
// The two variants share their parent's world and differ in the action mapping / observation only:
// env_cubes_push_diff_v0.py:209-211 (cube positions relative to the end effector), env_golf_joints_v0.py:60-63,85,123-124,229-234
// (six joint displacements instead of an end-effector displacement; the six arm joints prepended to the observation).
inline int scene_of(int env_type) { return env_type == ENV_PUSH_DIFF ? ENV_PUSH : (env_type == ENV_GOLF_JOINTS ? ENV_GOLF : env_type); }
// The oracle keeps every manifold point and every rolling-friction row (Bullet has no global cap, only 4 points per
// manifold). EXPORT_MAXC is the width of the contact-pair export only (urgo_contacts).
enum { NDOF = 12, NMOTOR = 7, MAXFREE = 3, EXPORT_MAXC = 64, STATE_SLOTS = 96 };
// Movable joints in Bullet link numbering; env code agrees (robot.py:68-73, first six = arm).
static const int MOV[NDOF] = {1, 2, 3, 4, 5, 6, 9, 11, 13, 14, 16, 18};
// Gear rows: parent finger_joint (link 9, dof 6) with each mimic child; gearRatio = -multiplier (robot.py:190-211).
static const int GEAR_DOF[5] = {7, 8, 9, 10, 11};            // links 11, 13, 14, 16, 18
static const double GEAR_RATIO[5] = {1.0, -1.0, -1.0, 1.0, -1.0};
// Order in which robot links enter the pair list, and so the order of the contact rows inside a Gauss-Seidel sweep
// (Bullet's own order is that of its manifold array, i.e. of broadphase pair creation: arbitrary). Gripper first.
static const int LINK_ORDER[17] = {12, 17, 11, 16, 10, 15, 9, 14, 13, 18, 8, 6, 5, 4, 3, 2, 1};

struct ShapeDef {
  int kind;   // 0 box, 1 sphere, 2 convex polytope (poly = index into URG_POLYS, in the frame of robot link `link`)
  int poly = -1;
  int body;   // 0 robot, 1..3 free body index+1, -1 static
  int link;   // robot link index, or sub-shape id on a free body
  V3 lc; M3 lR; V3 half; double radius;
  double mu, rest, roll, cbt;
};

// A body slot is a free rigid body (anchor 0) or, for the second articulated body of Box-v0 (skew-box-button.urdf: both of its
// joints hang off the fixed base), a world-anchored 1-DoF body: a hinge (anchor 1: rotation about the world axis `axis` through
// the body origin, which stays at `origin`) or a slider (anchor 2: translation along `axis` from `origin`, no rotation). Their
// generalized inverse inertia is rank one: axis axis^T / inertia_axis (angular) or axis axis^T / mass (linear).
struct Anchor {
  int kind = 0; V3 axis, origin; double inertia_axis = 0, lower = 0, upper = 0;
  V3 com;                 // hinge: merged centre of mass in the body frame (gravity torque about the axis)
  double motor_force = 0; // slider: POSITION_CONTROL motor to coordinate 0 with this force limit (env_box_v0.py:176), 0 = none
};
struct Scene {
  int env_type = 0, nfree = 0;
  double gravity = -9.8;
  V3 base;                // robot base position (env_box_v0.py:55: (0, 0.5, 0); every other env: the origin)
  Anchor anchor[MAXFREE];
  double fmass[MAXFREE]; V3 finertia[MAXFREE];
  std::vector<ShapeDef> shapes;
  std::vector<std::pair<int, int>> pairs;
  int obs_dim = 0, act_dim = 0, max_steps = 0;
};

inline ShapeDef mk_box(int body, int link, V3 lc, V3 half, double mu, double rest = 0, double roll = 0) {
  ShapeDef s; s.kind = 0; s.body = body; s.link = link; s.lc = lc; s.lR = M3::identity(); s.half = half; s.radius = 0;
  s.mu = mu; s.rest = rest; s.roll = roll; s.cbt = bp().contact_breaking * norm(half); return s;
}
inline ShapeDef mk_sphere(int body, int link, double r, double mu, double rest = 0, double roll = 0) {
  ShapeDef s; s.kind = 1; s.body = body; s.link = link; s.lc = V3(); s.lR = M3::identity(); s.half = V3(r, r, r); s.radius = r;
  s.mu = mu; s.rest = rest; s.roll = roll; s.cbt = bp().contact_breaking * r; return s;
}
inline V3 box_inertia(double m, V3 h) {
  double lx = 2 * h.x, ly = 2 * h.y, lz = 2 * h.z;
  return V3(m / 12 * (ly * ly + lz * lz), m / 12 * (lx * lx + lz * lz), m / 12 * (lx * lx + ly * ly));
}

// Club geometry (env_golf_v0.py:289-332): three 1 kg parts welded; one rigid body about the merged COM.
static const double CLUB_ZC = (-0.09 + 0.08) / 3.0;  // merged COM z in the club's base frame

inline Scene make_scene(int env_type_full) {
  const int env_type = scene_of(env_type_full);
  Scene sc; sc.env_type = env_type;
  for (int k = 0; k < 17; k++) {  // robot link boxes (link 0 rests on the table and link 7's 1 cm box is interior)
    const UrgLinkModel& L = URG_LINKS[LINK_ORDER[k]];
    ShapeDef s = mk_box(0, LINK_ORDER[k], V3(L.box_center[0], L.box_center[1], L.box_center[2]),
                        V3(L.box_half[0], L.box_half[1], L.box_half[2]), bp().robot_mu);
    for (int i = 0; i < 9; i++) s.lR.m[i] = L.box_rot[i];
    if (bp().hull_links != 0.0 && URG_LINK_POLY[LINK_ORDER[k]] >= 0) { s.kind = 2; s.poly = URG_LINK_POLY[LINK_ORDER[k]]; }
    if (!(((long)bp().link_mask >> LINK_ORDER[k]) & 1)) { s.half = V3(1e-6, 1e-6, 1e-6); s.kind = 0; }  // ablation: link does not collide
    else if (bp().box_shrink > 0 && LINK_ORDER[k] != 12 && LINK_ORDER[k] != 17) s.half = V3(std::fmax(s.half.x - bp().box_shrink, 1e-4), std::fmax(s.half.y - bp().box_shrink, 1e-4), std::fmax(s.half.z - bp().box_shrink, 1e-4));
    sc.shapes.push_back(s);
  }
  const V3 table_c(0, 0, -0.25), table_h(0.5, 1.0, 0.25);  // env_ball_balance_v0.py:69-71
  const double bm = bp().body_mu, cm = bp().cube_mu, cmass = bp().cube_mass;
  // plane.urdf (30 x 30 x 10 box collider, 5 m below its origin, friction 1 [recalled]) at z = -0.5 under the table
  // (env_ball_balance_v0.py:70, env_cubes_push_v0.py:77, env_golf_v0.py:77); static id 2. Whatever leaves the table
  // comes to rest on it instead of falling for the rest of the episode.
  const ShapeDef ground = mk_box(-1, 2, V3(0, 0, -0.5 - 5.0), V3(15, 15, 5), bp().plane_mu);
  switch (env_type) {
    case ENV_BALL:  // env_ball_balance_v0.py:177-203
      sc.nfree = 2; sc.obs_dim = 6; sc.act_dim = 3; sc.max_steps = 500;
      sc.fmass[0] = 0.1; sc.finertia[0] = box_inertia(0.1, V3(0.05, 0.1, 0.005));
      sc.fmass[1] = 0.01; sc.finertia[1] = V3(1, 1, 1) * (0.4 * 0.01 * 0.02 * 0.02);
      sc.shapes.push_back(mk_box(1, -1, V3(), V3(0.05, 0.1, 0.005), bm));
      sc.shapes.push_back(mk_sphere(2, -1, 0.02, bm));
      sc.shapes.push_back(mk_box(-1, 0, table_c, table_h, bm));
      sc.shapes.push_back(ground);
      break;
    case ENV_TWO:  // env_two_balls_balance_v0.py:187,200-233
      sc.nfree = 3; sc.obs_dim = 9; sc.act_dim = 3; sc.max_steps = 500;
      sc.fmass[0] = 0.1; sc.finertia[0] = box_inertia(0.1, V3(0.05, 0.1, 0.005));
      for (int b = 1; b < 3; b++) { sc.fmass[b] = 0.01; sc.finertia[b] = V3(1, 1, 1) * (0.4 * 0.01 * 0.003 * 0.003); }
      sc.shapes.push_back(mk_box(1, -1, V3(), V3(0.05, 0.1, 0.005), bm));
      sc.shapes.push_back(mk_sphere(2, -1, 0.003, bm));
      sc.shapes.push_back(mk_sphere(3, -1, 0.003, bm));
      sc.shapes.push_back(mk_box(-1, 0, table_c, table_h, bm));
      sc.shapes.push_back(ground);
      break;
    case ENV_PUSH:  // env_cubes_push_v0.py:233-246; cube.urdf: unit box, mass 1, lateral friction 1 [recalled]
      sc.nfree = 2; sc.obs_dim = 9; sc.act_dim = 3; sc.max_steps = 50;
      for (int b = 0; b < 2; b++) {
        sc.fmass[b] = cmass; sc.finertia[b] = box_inertia(cmass, V3(0.02, 0.02, 0.02));
        sc.shapes.push_back(mk_box(b + 1, -1, V3(), V3(0.02, 0.02, 0.02), cm));
      }
      sc.shapes.push_back(mk_box(-1, 0, table_c, table_h, bm));
      sc.shapes.push_back(ground);
      break;
    case ENV_GRASP:  // env_cubes_grasp_v0.py:43,48,305; plane.urdf: 30x30x10 box 5 m below origin, friction 1 [recalled]
      sc.nfree = 1; sc.obs_dim = 17; sc.act_dim = 9; sc.max_steps = 20; sc.gravity = -10.0;
      sc.fmass[0] = cmass; sc.finertia[0] = box_inertia(cmass, V3(0.02, 0.02, 0.02));
      sc.shapes.push_back(mk_box(1, -1, V3(), V3(0.02, 0.02, 0.02), cm));
      sc.shapes.push_back(mk_box(-1, 0, V3(0, 0, -5.0), V3(15, 15, 5), bp().plane_mu));
      break;
    case ENV_GOLF: {  // env_golf_v0.py:87,289-358
      sc.nfree = 2; sc.obs_dim = 9; sc.act_dim = 3; sc.max_steps = 100;
      double zc = CLUB_ZC;
      V3 ib = box_inertia(1.0, V3(0.015, 0.015, 0.075)), it = box_inertia(1.0, V3(0.05, 0.015, 0.015));
      V3 id(1.0 / 12 * 0.01 * 0.01 + 0.25 * 0.04 * 0.04, 1.0 / 12 * 0.01 * 0.01 + 0.25 * 0.04 * 0.04, 0.5 * 0.04 * 0.04);
      double d0 = 0 - zc, d1 = -0.09 - zc, d2 = 0.08 - zc;
      sc.fmass[0] = 3.0;
      sc.finertia[0] = V3(ib.x + it.x + id.x + d0 * d0 + d1 * d1 + d2 * d2, ib.y + it.y + id.y + d0 * d0 + d1 * d1 + d2 * d2, ib.z + it.z + id.z);
      sc.fmass[1] = 0.01; sc.finertia[1] = V3(1, 1, 1) * (0.4 * 0.01 * 0.015 * 0.015);
      sc.shapes.push_back(mk_box(1, -1, V3(0, 0, d0), V3(0.015, 0.015, 0.075), 10.0, 0.1));
      sc.shapes.push_back(mk_box(1, 0, V3(0, 0, d1), V3(0.05, 0.015, 0.015), bm));
      sc.shapes.push_back(mk_box(1, 1, V3(0, 0, d2), V3(bp().disc_half, bp().disc_half, 0.005), bm));  // disc r=0.04 h=0.01 as a box
      sc.shapes.push_back(mk_sphere(2, -1, 0.015, bm, 0.1, 10.0));
      sc.shapes.push_back(mk_box(-1, 0, table_c, table_h, bm));
      sc.shapes.push_back(mk_box(-1, 1, V3(0, -0.6, 0.02), V3(URG_PITCH_HALF[0], URG_PITCH_HALF[1], URG_PITCH_HALF[2]), bm));
      sc.shapes.push_back(ground);
      break;
    }
    case ENV_BOX: {  // env_box_v0.py:55-85, urgym/urdf/skew-box-button.urdf (URDF_MERGE_FIXED_LINKS | URDF_USE_SELF_COLLISION, fixed base)
      sc.nfree = 2; sc.obs_dim = 7; sc.act_dim = 8; sc.max_steps = 200; sc.gravity = -10.0;
      sc.base = V3(0, 0.5, 0);
      // body 1: the door (link `door` + the welded handle bars h1, h2, h3), hinge about y at (0.075, 0, 0.1), limits [0, 2];
      // inertia about the axis: iyy = 0.5 of the door link (no inertial origin: its COM is the hinge) + the bars as point masses
      const double mh1 = 0.05, mh3 = 0.1;
      sc.anchor[0].kind = 1; sc.anchor[0].axis = V3(0, 1, 0); sc.anchor[0].origin = V3(0.075, 0, 0.1); sc.anchor[0].lower = 0; sc.anchor[0].upper = 2;
      sc.anchor[0].inertia_axis = 0.5 + 2 * mh1 * (0.13 * 0.13 + 0.025 * 0.025) + mh3 * (0.13 * 0.13 + 0.05 * 0.05);
      sc.fmass[0] = 0.1 + 2 * mh1 + mh3; sc.finertia[0] = V3(1, 1, 1);
      sc.anchor[0].com = V3(-0.13 * (2 * mh1 + mh3), 0, 2 * mh1 * 0.025 + mh3 * 0.05) * (1.0 / sc.fmass[0]);
      // body 2: the button, slider along z from (0, 0, 0.07), limits [-0.05, 0], position motor with force 1 (reset_box)
      sc.anchor[1].kind = 2; sc.anchor[1].axis = V3(0, 0, 1); sc.anchor[1].origin = V3(0, 0, 0.07); sc.anchor[1].lower = -0.05; sc.anchor[1].upper = 0;
      sc.anchor[1].motor_force = 1.0; sc.fmass[1] = 0.01; sc.finertia[1] = V3(1, 1, 1);
      // collision shapes. The URDF's four cylinders (handle posts r = 5 mm, bar r = 10 mm, button r = 20 mm) enter as their
      // circumscribed square prisms: the narrowphase has boxes, spheres and the gripper's polytopes
      sc.shapes.push_back(mk_box(1, 0, V3(-0.076, 0, 0), V3(0.08, 0.1, 0.005), bm));          // door plate
      sc.shapes.push_back(mk_box(1, 1, V3(-0.13, 0.05, 0.025), V3(0.005, 0.005, 0.025), bm));  // h1
      sc.shapes.push_back(mk_box(1, 2, V3(-0.13, -0.05, 0.025), V3(0.005, 0.005, 0.025), bm)); // h2
      sc.shapes.push_back(mk_box(1, 3, V3(-0.13, 0, 0.05), V3(0.01, 0.05, 0.01), 1.0));        // h3 (contact_coefficients mu = 1), axis along y
      sc.shapes.push_back(mk_box(2, 0, V3(), V3(0.02, 0.02, 0.025), bm));                      // button
      sc.shapes.push_back(mk_box(-1, 0, V3(0.075, 0, 0.045), V3(0.005, 0.1, 0.045), bm));      // box1..box4: the walls
      sc.shapes.push_back(mk_box(-1, 1, V3(-0.075, 0, 0.045), V3(0.005, 0.1, 0.045), bm));
      sc.shapes.push_back(mk_box(-1, 2, V3(0, 0.1, 0.045), V3(0.075, 0.005, 0.045), bm));
      sc.shapes.push_back(mk_box(-1, 3, V3(0, -0.1, 0.045), V3(0.075, 0.005, 0.045), bm));
      sc.shapes.push_back(mk_box(-1, 4, V3(0, 0, -5.0), V3(15, 15, 5), bp().plane_mu));        // plane.urdf at the origin
      break;
    }
  }
  if (env_type_full == ENV_GOLF_JOINTS) { sc.obs_dim = 15; sc.act_dim = 6; }
  int ns = (int)sc.shapes.size();
  auto is_free = [&](int i) { return sc.shapes[i].body > 0; };
  auto is_static = [&](int i) { return sc.shapes[i].body < 0; };
  auto is_robot = [&](int i) { return sc.shapes[i].body == 0; };
  for (int a = 0; a < ns; a++) for (int b = a + 1; b < ns; b++)
    if (is_free(a) && is_free(b) && sc.shapes[a].body != sc.shapes[b].body) sc.pairs.push_back({a, b});
  // (Box-v0: door and button are anchored inside the box and never reach the ground plane)
  for (int a = 0; a < ns; a++) for (int b = 0; b < ns; b++)
    if (is_free(a) && is_static(b) && !(env_type == ENV_BOX && sc.shapes[b].link == 4)) sc.pairs.push_back({a, b});
  // Box-v0: the upper arm links (the last four of LINK_ORDER: 4, 3, 2, 1) are tested against the ground plane only, which keeps
  // the pair list within the kernel's 160 (they cannot reach the box without the wrist or the gripper touching it first)
  auto arm_only = [&](int a, int b) { return env_type == ENV_BOX && a >= 13 && !(is_static(b) && sc.shapes[b].link == 4); };
  for (int a = 0; a < ns; a++) for (int b = 0; b < ns; b++) if (is_robot(a) && is_free(b) && !arm_only(a, b)) sc.pairs.push_back({a, b});
  for (int a = 0; a < ns; a++) for (int b = 0; b < ns; b++) if (is_robot(a) && is_static(b) && !arm_only(a, b)) sc.pairs.push_back({a, b});
  return sc;
}

struct FreeBody {
  V3 pos; Q4 quat; V3 lin, ang; bool active = false;
};

struct Contact {
  int sa, sb;          // shape indices (A, B); normal points from B to A
  V3 posA, posB, n; double dist;
  double mu, rest, roll;
};

// ----------------------------------------------------------------------------------------------------------------
// Spatial algebra in WORLD coordinates about the world origin: motion (w, vO), force (nO, f). 6x6 row-major.
struct S6 { double a[6]; };
inline S6 s6(V3 w, V3 v) { S6 r; r.a[0] = w.x; r.a[1] = w.y; r.a[2] = w.z; r.a[3] = v.x; r.a[4] = v.y; r.a[5] = v.z; return r; }
inline V3 top(const S6& s) { return V3(s.a[0], s.a[1], s.a[2]); }
inline V3 bot(const S6& s) { return V3(s.a[3], s.a[4], s.a[5]); }
inline S6 operator+(const S6& x, const S6& y) { S6 r; for (int i = 0; i < 6; i++) r.a[i] = x.a[i] + y.a[i]; return r; }
inline S6 operator-(const S6& x, const S6& y) { S6 r; for (int i = 0; i < 6; i++) r.a[i] = x.a[i] - y.a[i]; return r; }
inline S6 operator*(const S6& x, double s) { S6 r; for (int i = 0; i < 6; i++) r.a[i] = x.a[i] * s; return r; }
inline double dot6(const S6& x, const S6& y) { double s = 0; for (int i = 0; i < 6; i++) s += x.a[i] * y.a[i]; return s; }
inline S6 crm(const S6& v, const S6& m) { return s6(cross(top(v), top(m)), cross(top(v), bot(m)) + cross(bot(v), top(m))); }
inline S6 crf(const S6& v, const S6& f) { return s6(cross(top(v), top(f)) + cross(bot(v), bot(f)), cross(top(v), bot(f))); }
struct M6 { double a[36]; };
inline S6 mul(const M6& A, const S6& x) { S6 r; for (int i = 0; i < 6; i++) { double s = 0; for (int j = 0; j < 6; j++) s += A.a[6 * i + j] * x.a[j]; r.a[i] = s; } return r; }
inline M6 spatial_inertia(double m, V3 c, const M3& Ic) {
  M6 I; std::memset(I.a, 0, sizeof(I.a));
  double cc = dot(c, c);
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) I.a[6 * i + j] = Ic(i, j) + m * ((i == j ? cc : 0.0) - c[i] * c[j]);
  double cx[9] = {0, -c.z, c.y, c.z, 0, -c.x, -c.y, c.x, 0};
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) { I.a[6 * i + 3 + j] = m * cx[3 * i + j]; I.a[6 * (3 + i) + j] = -m * cx[3 * i + j]; }
  for (int i = 0; i < 3; i++) I.a[6 * (3 + i) + 3 + i] = m;
  return I;
}

struct Robot {
  double q[NDOF], qd[NDOF];
  // kinematics (filled by fk())
  V3 p[URG_NLINK]; Q4 quat[URG_NLINK]; M3 R[URG_NLINK]; V3 com[URG_NLINK]; V3 axis[URG_NLINK];
  // ABA cache (filled by aba())
  S6 S[URG_NLINK], v[URG_NLINK], c[URG_NLINK], pA[URG_NLINK], U[URG_NLINK], acc[URG_NLINK];
  M6 IA[URG_NLINK]; double D[URG_NLINK], u[URG_NLINK];
  int dof_of_link[URG_NLINK]; uint32_t anc[URG_NLINK];  // anc: bitmask over dofs of joints on the path to the base
  Q4 jquat[URG_NLINK];
  V3 base;  // position of the base link in the world (robot.py:39-41: base_pos)

  Robot() {
    std::memset(q, 0, sizeof(q)); std::memset(qd, 0, sizeof(qd));
    for (int i = 0; i < URG_NLINK; i++) dof_of_link[i] = -1;
    for (int d = 0; d < NDOF; d++) dof_of_link[MOV[d]] = d;
    for (int i = 0; i < URG_NLINK; i++) {
      int par = URG_LINKS[i].parent;
      anc[i] = par >= 0 ? anc[par] : 0u;
      if (dof_of_link[i] >= 0) anc[i] |= 1u << dof_of_link[i];
      M3 J; for (int k = 0; k < 9; k++) J.m[k] = URG_LINKS[i].jrot[k];
      jquat[i] = mat_to_quat(J);
    }
  }

  // Link frames by chaining quaternions parent -> joint origin -> joint rotation.
  void fk() {
    for (int i = 0; i < URG_NLINK; i++) {
      const UrgLinkModel& L = URG_LINKS[i];
      V3 pp = base; Q4 pq; M3 pR = M3::identity();
      if (L.parent >= 0) { pp = p[L.parent]; pq = quat[L.parent]; pR = R[L.parent]; }
      p[i] = pp + pR * V3(L.jxyz[0], L.jxyz[1], L.jxyz[2]);
      Q4 qi = pq * jquat[i];
      if (L.jtype == 1) qi = qi * axis_angle(V3(L.axis[0], L.axis[1], L.axis[2]), q[dof_of_link[i]]);
      quat[i] = normalized(qi);
      R[i] = quat_to_mat(quat[i]);
      com[i] = p[i] + R[i] * V3(L.com[0], L.com[1], L.com[2]);
      axis[i] = R[i] * V3(L.axis[0], L.axis[1], L.axis[2]);
    }
  }

  // Featherstone ABA, all 19 links, zero applied joint torque (motors are constraint rows). Output qdd[12].
  // External forces: gravity at each link COM, Bullet's velocity-dependent damping, gyroscopic term (row P2).
  void aba(double gz, double* qdd) {
    for (int i = 0; i < URG_NLINK; i++) {
      const UrgLinkModel& L = URG_LINKS[i];
      int d = dof_of_link[i];
      S6 vJ = s6(V3(), V3());
      if (d >= 0) { S[i] = s6(axis[i], cross(p[i], axis[i])); vJ = S[i] * qd[d]; }
      S6 vp = L.parent >= 0 ? v[L.parent] : s6(V3(), V3());
      v[i] = vp + vJ;
      c[i] = crm(v[i], vJ);
      M3 Ic = rot_diag(R[i], V3(L.inertia[0], L.inertia[1], L.inertia[2]));
      IA[i] = spatial_inertia(L.mass, com[i], Ic);
      V3 w = top(v[i]), vc = bot(v[i]) + cross(w, com[i]);
      V3 f = V3(0, 0, L.mass * gz) - vc * (L.mass * (bp().lin_damping + bp().lin_damping * norm(vc)));
      V3 tq = -1.0 * ((Ic * w) * (bp().ang_damping + bp().ang_damping * norm(w)));
      S6 fext = s6(tq + cross(com[i], f), f);
      pA[i] = crf(v[i], mul(IA[i], v[i])) - fext;
    }
    for (int i = URG_NLINK - 1; i >= 0; i--) {
      int par = URG_LINKS[i].parent, d = dof_of_link[i];
      M6 Ia = IA[i]; S6 pa;
      if (d >= 0) {
        U[i] = mul(IA[i], S[i]); D[i] = dot6(S[i], U[i]); u[i] = -dot6(S[i], pA[i]);
        for (int r = 0; r < 6; r++) for (int s = 0; s < 6; s++) Ia.a[6 * r + s] -= U[i].a[r] * U[i].a[s] / D[i];
        pa = pA[i] + mul(Ia, c[i]) + U[i] * (u[i] / D[i]);
      } else {
        pa = pA[i] + mul(Ia, c[i]);
      }
      if (par >= 0) { for (int k = 0; k < 36; k++) IA[par].a[k] += Ia.a[k]; pA[par] = pA[par] + pa; }
    }
    for (int i = 0; i < URG_NLINK; i++) {
      int par = URG_LINKS[i].parent, d = dof_of_link[i];
      S6 ap = (par >= 0 ? acc[par] : s6(V3(), V3())) + c[i];
      if (d >= 0) { qdd[d] = (u[i] - dot6(U[i], ap)) / D[i]; acc[i] = ap + S[i] * qdd[d]; }
      else acc[i] = ap;
    }
  }

  // M^-1 tau with the cached articulated inertias (what Bullet's calcAccelerationDeltasMultiDof does per row).
  void minv_mul(const double* tau, double* out) const {
    S6 pa[URG_NLINK]; double uu[URG_NLINK]; S6 a[URG_NLINK];
    for (int i = 0; i < URG_NLINK; i++) pa[i] = s6(V3(), V3());
    for (int i = URG_NLINK - 1; i >= 0; i--) {
      int par = URG_LINKS[i].parent, d = dof_of_link[i];
      S6 x = pa[i];
      if (d >= 0) { uu[i] = tau[d] - dot6(S[i], pa[i]); x = x + U[i] * (uu[i] / D[i]); }
      if (par >= 0) pa[par] = pa[par] + x;
    }
    for (int i = 0; i < URG_NLINK; i++) {
      int par = URG_LINKS[i].parent, d = dof_of_link[i];
      S6 ap = par >= 0 ? a[par] : s6(V3(), V3());
      if (d >= 0) { out[d] = (uu[i] - dot6(U[i], ap)) / D[i]; a[i] = ap + S[i] * out[d]; }
      else a[i] = ap;
    }
  }

  // Generalized force of a unit force `dir` at world point P (lin=true) or a unit torque `dir` (lin=false) on link L.
  void jac_row(int link, V3 P, V3 dir, bool lin, double* J) const {
    for (int d = 0; d < NDOF; d++) {
      J[d] = 0;
      if (!(anc[link] >> d & 1)) continue;
      int j = MOV[d];
      J[d] = lin ? dot(dir, cross(axis[j], P - p[j])) : dot(dir, axis[j]);
    }
  }
};

// One solver row: up to one robot side and up to two free-body sides.
struct Row {
  bool has_robot = false; double Jr[NDOF], Dr[NDOF];
  int f[2] = {-1, -1}; V3 Jl[2], Ja[2], Dl[2], Da[2];
  double rhs = 0, lo = 0, hi = 0, invd = 0, applied = 0, mu = 0;
  int nidx = -1;  // index of the normal row this friction row is limited by
};

struct World {
  Scene sc;
  Robot rob;
  FreeBody fb[MAXFREE];
  double target[NMOTOR]; bool arm_armed = false, grip_armed = false;
  std::vector<Contact> contacts;
  int last_iters = 0;
  long st_ticks = 0, st_iters = 0, st_rows = 0, st_contacts = 0, st_rowiters = 0;  // profiling counters (tests / design notes)
  long st_hist_rows[192] = {0}, st_hist_nc[64] = {0}, st_hist_nnc[32] = {0}, st_hist_roll[16] = {0};

  explicit World(int env_type) : sc(make_scene(env_type)) { for (int i = 0; i < NMOTOR; i++) target[i] = 0; rob.base = sc.base; }

  // coordinate and velocity of an anchored body's single degree of freedom
  double anchor_q(int b) const {
    const Anchor& A = sc.anchor[b];
    if (A.kind == 1) { const Q4& q = fb[b].quat; return 2.0 * std::atan2(q.x * A.axis.x + q.y * A.axis.y + q.z * A.axis.z, q.w); }
    return dot(fb[b].pos - A.origin, A.axis);
  }
  double anchor_qd(int b) const { return sc.anchor[b].kind == 1 ? dot(fb[b].ang, sc.anchor[b].axis) : dot(fb[b].lin, sc.anchor[b].axis); }

  BoxW world_box(const ShapeDef& s) const {
    BoxW b;
    if (s.body == 0) { b.R = rob.R[s.link] * s.lR; b.c = rob.p[s.link] + rob.R[s.link] * s.lc; }
    else if (s.body > 0) { M3 R = quat_to_mat(fb[s.body - 1].quat); b.R = R * s.lR; b.c = fb[s.body - 1].pos + R * s.lc; }
    else { b.R = s.lR; b.c = s.lc; }
    b.h = s.half; return b;
  }
  bool shape_active(const ShapeDef& s) const { return s.body <= 0 || fb[s.body - 1].active; }

  double contact_margin = 1e300;  // decision_margin() of the last collide()
  void collide() {
    contacts.clear();
    decision_margin() = 1e300;
    for (auto& pr : sc.pairs) {
      const ShapeDef& A = sc.shapes[pr.first]; const ShapeDef& B = sc.shapes[pr.second];
      if (!shape_active(A) || !shape_active(B)) continue;
      double margin = std::fmin(A.cbt, B.cbt);
      CPoint pts[4]; int n = 0;
      BoxW a = world_box(A), b = world_box(B);
      if (A.kind == 2) {   // robot link with mesh geometry against a box or a sphere
        // mid-phase: the link's bounding box (which contains the polytope) with the same margin; only then the polytope
        bool close;
        if (B.kind == 0) close = box_box(a, b, margin, pts) > 0;
        else { V3 oa, na; double d; close = sphere_box(b.c, B.radius, a, margin, oa, na, d); }
        if (close) {
          PolyW pa = poly_world(URG_POLYS[A.poly], rob.p[A.link], rob.R[A.link]);
          if (B.kind == 0) n = poly_box(pa, b, margin, pts);
          else {
            V3 oa, na; double d;
            if (sphere_poly(b.c, B.radius, pa, margin, oa, na, d)) { pts[0].n = -na; pts[0].dist = d; pts[0].posA = oa; pts[0].posB = oa + na * d; n = 1; }
          }
        }
      } else if (A.kind == 0 && B.kind == 0) {
        n = box_box(a, b, margin, pts);
      } else if (A.kind == 1 && B.kind == 0) {
        V3 ob, nb; double d;
        if (sphere_box(a.c, A.radius, b, margin, ob, nb, d)) { pts[0].n = nb; pts[0].dist = d; pts[0].posB = ob; pts[0].posA = ob + nb * d; n = 1; }
      } else if (A.kind == 0 && B.kind == 1) {
        V3 oa, na; double d;
        if (sphere_box(b.c, B.radius, a, margin, oa, na, d)) { pts[0].n = -na; pts[0].dist = d; pts[0].posA = oa; pts[0].posB = oa + na * d; n = 1; }
      } else {
        V3 dv = a.c - b.c; double len = norm(dv); double d = len - A.radius - B.radius;
        note_decision(d - margin);
        if (d < margin) {
          V3 nn = len > 1e-12 ? dv * (1.0 / len) : V3(1, 0, 0);
          pts[0].n = nn; pts[0].dist = d; pts[0].posB = b.c + nn * B.radius; pts[0].posA = pts[0].posB + nn * d; n = 1;
        }
      }
      for (int i = 0; i < n; i++) {
        Contact c; c.sa = pr.first; c.sb = pr.second; c.posA = pts[i].posA; c.posB = pts[i].posB; c.n = pts[i].n; c.dist = pts[i].dist;
        c.mu = std::fmax(-bp().max_friction, std::fmin(bp().max_friction, A.mu * B.mu));
        c.rest = A.rest * B.rest;
        c.roll = std::fmax(-bp().max_friction, std::fmin(bp().max_friction, A.roll * B.mu + B.roll * A.mu));
        if (i >= (int)bp().roll_points) c.roll = 0;  // only the first point of a manifold gets rolling-friction rows [recalled]
        contacts.push_back(c);
      }
    }
    contact_margin = decision_margin();
  }

  // side of a row for one shape: direction dir at point P (lin) or pure torque dir (!lin), scaled by sgn
  void add_side(Row& r, const ShapeDef& s, V3 P, V3 dir, bool lin, double sgn) {
    if (s.body == 0) {
      double J[NDOF]; rob.jac_row(s.link, P, dir, lin, J);
      r.has_robot = true;
      for (int d = 0; d < NDOF; d++) r.Jr[d] = sgn * J[d];
    } else if (s.body > 0) {
      int k = r.f[0] < 0 ? 0 : 1;
      int b = s.body - 1; r.f[k] = b;
      M3 R = quat_to_mat(fb[b].quat);
      V3 inv_i(1.0 / sc.finertia[b].x, 1.0 / sc.finertia[b].y, 1.0 / sc.finertia[b].z);
      M3 Iinv = rot_diag(R, inv_i);
      if (lin) { r.Jl[k] = dir * sgn; r.Ja[k] = cross(P - fb[b].pos, dir) * sgn; }
      else { r.Jl[k] = V3(); r.Ja[k] = dir * sgn; }
      const Anchor& An = sc.anchor[b];
      if (An.kind == 0) { r.Dl[k] = r.Jl[k] * (1.0 / sc.fmass[b]); r.Da[k] = Iinv * r.Ja[k]; }
      else if (An.kind == 1) { r.Dl[k] = V3(); r.Da[k] = An.axis * (dot(An.axis, r.Ja[k]) / An.inertia_axis); }
      else { r.Dl[k] = An.axis * (dot(An.axis, r.Jl[k]) / sc.fmass[b]); r.Da[k] = V3(); }
    }
  }
  void finish_row(Row& r) {
    double den = 0;
    if (r.has_robot) { rob.minv_mul(r.Jr, r.Dr); for (int d = 0; d < NDOF; d++) den += r.Jr[d] * r.Dr[d]; }
    for (int k = 0; k < 2; k++) if (r.f[k] >= 0) den += dot(r.Jl[k], r.Dl[k]) + dot(r.Ja[k], r.Da[k]);
    r.invd = den > 1e-20 ? 1.0 / den : 0.0;  // a row no body can move along (anchored bodies: a tangent across the hinge axis) stays inert
  }
  double row_vel(const Row& r, const double* qd) const {
    double s = 0;
    if (r.has_robot) for (int d = 0; d < NDOF; d++) s += r.Jr[d] * qd[d];
    for (int k = 0; k < 2; k++) if (r.f[k] >= 0) s += dot(r.Jl[k], fb[r.f[k]].lin) + dot(r.Ja[k], fb[r.f[k]].ang);
    return s;
  }

  struct DV { double r[NDOF]; V3 l[MAXFREE], a[MAXFREE]; };
  static double row_dv(const Row& r, const DV& dv) {
    double s = 0;
    if (r.has_robot) for (int d = 0; d < NDOF; d++) s += r.Jr[d] * dv.r[d];
    for (int k = 0; k < 2; k++) if (r.f[k] >= 0) s += dot(r.Jl[k], dv.l[r.f[k]]) + dot(r.Ja[k], dv.a[r.f[k]]);
    return s;
  }
  static void row_apply(const Row& r, DV& dv, double imp) {
    if (r.has_robot) for (int d = 0; d < NDOF; d++) dv.r[d] += r.Dr[d] * imp;
    for (int k = 0; k < 2; k++) if (r.f[k] >= 0) { dv.l[r.f[k]] += r.Dl[k] * imp; dv.a[r.f[k]] += r.Da[k] * imp; }
  }
  // Bullet's resolveSingleConstraintRowGeneric for multibody rows; returns the velocity change (residual) [recalled]
  static double resolve(Row& r, DV& dv) {
    double delta = r.rhs - row_dv(r, dv) * r.invd;
    double sum = r.applied + delta;
    if (sum < r.lo) { delta = r.lo - r.applied; r.applied = r.lo; }
    else if (sum > r.hi) { delta = r.hi - r.applied; r.applied = r.hi; }
    else r.applied = sum;
    row_apply(r, dv, delta);
    return r.invd > 0 ? delta / r.invd : 0.0;
  }

  // One 1/240 s tick of stepSimulation.
  void tick() {
    const double dt = BP_DT;
    rob.fk();
    collide();
    // --- unconstrained velocity update
    double qdd[NDOF];
    rob.aba(sc.gravity, qdd);
    for (int d = 0; d < NDOF; d++) rob.qd[d] = std::fmax(-BP_MAX_COORD_VEL, std::fmin(BP_MAX_COORD_VEL, rob.qd[d] + dt * qdd[d]));
    for (int b = 0; b < sc.nfree; b++) {
      if (!fb[b].active) continue;
      FreeBody& B = fb[b];
      M3 R = quat_to_mat(B.quat);
      if (sc.anchor[b].kind == 1) {        // hinge: gravity torque of the merged COM about the axis, Bullet's angular damping
        const Anchor& An = sc.anchor[b];
        double w = dot(B.ang, An.axis);
        double tq = dot(An.axis, cross(R * An.com, V3(0, 0, sc.fmass[b] * sc.gravity)));
        w += dt * (tq / An.inertia_axis - w * (bp().ang_damping + bp().ang_damping * std::fabs(w)));
        B.ang = An.axis * w; B.lin = V3();
        continue;
      }
      if (sc.anchor[b].kind == 2) {        // slider: gravity and linear damping along the axis
        const Anchor& An = sc.anchor[b];
        double v = dot(B.lin, An.axis);
        v += dt * (An.axis.z * sc.gravity - v * (bp().lin_damping + bp().lin_damping * std::fabs(v)));
        B.lin = An.axis * v; B.ang = V3();
        continue;
      }
      V3 acc = V3(0, 0, sc.gravity) - B.lin * (bp().lin_damping + bp().lin_damping * norm(B.lin));
      V3 wl = tmul(R, B.ang);
      V3 I = sc.finertia[b];
      V3 Iw(I.x * wl.x, I.y * wl.y, I.z * wl.z);
      V3 tq = -1.0 * (Iw * (bp().ang_damping + bp().ang_damping * norm(wl))) - cross(wl, Iw);
      V3 al(tq.x / I.x, tq.y / I.y, tq.z / I.z);
      B.lin += acc * dt; B.ang += (R * al) * dt;
    }
    // --- rows
    std::vector<Row> nc, nr, fr, tr;
    for (int d = 0; d < NDOF; d++) {  // joint limits, lower then upper (row P4)
      const UrgLinkModel& L = URG_LINKS[MOV[d]];
      for (int side = 0; side < 2; side++) {
        double pen = side == 0 ? rob.q[d] - L.lower : L.upper - rob.q[d];
        if (pen > 0) continue;
        Row r; r.has_robot = true; for (int k = 0; k < NDOF; k++) r.Jr[k] = 0;
        r.Jr[d] = side == 0 ? 1.0 : -1.0;
        finish_row(r);
        double rel = row_vel(r, rob.qd);
        r.rhs = (-pen * bp().erp / dt - rel) * r.invd; r.lo = 0; r.hi = BP_LIMIT_MAX_IMPULSE;
        nc.push_back(r);
      }
    }
    for (int m = 0; m < NMOTOR; m++) {  // position motors (row P3); zero-force default motors are no-ops and skipped
      if (m < 6 ? !arm_armed : !grip_armed) continue;
      const UrgLinkModel& L = URG_LINKS[MOV[m]];
      Row r; r.has_robot = true; for (int k = 0; k < NDOF; k++) r.Jr[k] = 0;
      r.Jr[m] = 1.0;
      finish_row(r);
      double want = bp().motor_kp * (target[m] - rob.q[m]) / dt + (1.0 - bp().motor_kd) * rob.qd[m];
      want = std::fmax(-L.velocity, std::fmin(L.velocity, want));
      r.rhs = (want - rob.qd[m]) * r.invd; r.lo = -L.effort * dt; r.hi = L.effort * dt;
      nc.push_back(r);
    }
    for (int g = 0; g < 5; g++) {  // mimic gears (row A5)
      Row r; r.has_robot = true; for (int k = 0; k < NDOF; k++) r.Jr[k] = 0;
      r.Jr[6] = 1.0; r.Jr[GEAR_DOF[g]] = GEAR_RATIO[g];
      finish_row(r);
      double rel = row_vel(r, rob.qd);
      double pos_err = BP_GEAR_ERP * (rob.q[6] + GEAR_RATIO[g] * rob.q[GEAR_DOF[g]]);
      r.rhs = (-pos_err * bp().erp / dt - rel) * r.invd; r.lo = -BP_GEAR_MAX_FORCE * dt; r.hi = BP_GEAR_MAX_FORCE * dt;
      nc.push_back(r);
    }
    // joint rows of the anchored bodies (Box-v0: btMultiBodyJointLimitConstraint of door and button when violated, the
    // button's position motor): solved after the robot's joint rows and before the contact normals, in this order
    std::vector<Row> nb;
    for (int b = 0; b < sc.nfree; b++) {
      const Anchor& An = sc.anchor[b];
      if (An.kind == 0 || !fb[b].active) continue;
      const double qa = anchor_q(b), va = anchor_qd(b);
      for (int side = 0; side < 2; side++) {
        double pen = side == 0 ? qa - An.lower : An.upper - qa;
        if (pen > 0) continue;
        Row r; r.f[0] = b; double sg = side == 0 ? 1.0 : -1.0;
        r.Jl[0] = An.kind == 2 ? An.axis * sg : V3(); r.Ja[0] = An.kind == 1 ? An.axis * sg : V3();
        r.Dl[0] = An.kind == 2 ? r.Jl[0] * (1.0 / sc.fmass[b]) : V3(); r.Da[0] = An.kind == 1 ? r.Ja[0] * (1.0 / An.inertia_axis) : V3();
        finish_row(r);
        r.rhs = (-pen * bp().erp / dt - sg * va) * r.invd; r.lo = 0; r.hi = BP_LIMIT_MAX_IMPULSE;
        nb.push_back(r);
      }
      if (An.motor_force > 0) {
        Row r; r.f[0] = b;
        r.Jl[0] = An.kind == 2 ? An.axis : V3(); r.Ja[0] = An.kind == 1 ? An.axis : V3();
        r.Dl[0] = An.kind == 2 ? r.Jl[0] * (1.0 / sc.fmass[b]) : V3(); r.Da[0] = An.kind == 1 ? r.Ja[0] * (1.0 / An.inertia_axis) : V3();
        finish_row(r);
        double want = bp().motor_kp * (0.0 - qa) / dt + (1.0 - bp().motor_kd) * va;
        r.rhs = (want - va) * r.invd; r.lo = -An.motor_force * dt; r.hi = An.motor_force * dt;
        nb.push_back(r);
      }
    }
    for (size_t ci = 0; ci < contacts.size(); ci++) {  // contacts (rows P5, P6)
      const Contact& c = contacts[ci];
      const ShapeDef& A = sc.shapes[c.sa]; const ShapeDef& B = sc.shapes[c.sb];
      Row n; add_side(n, A, c.posA, c.n, true, 1.0); add_side(n, B, c.posB, c.n, true, -1.0); finish_row(n);
      double rel = row_vel(n, rob.qd);
      double pen = c.dist + bp().linear_slop;
      double rest = 0;
      if (std::fabs(rel) >= bp().restitution_threshold) { rest = c.rest * -rel; if (rest < 0) rest = 0; }
      double pos_e = 0, vel_e = rest - rel;
      if (pen > 0) vel_e -= pen / dt; else pos_e = -pen * bp().erp2 / dt;
      n.rhs = (pos_e + vel_e) * n.invd; n.lo = 0; n.hi = 1e10;
      int nidx = (int)nr.size(); nr.push_back(n);
      V3 t1, t2; plane_space(c.n, t1, t2);
      if (c.roll > 0) {
        for (int k = 0; k < 2; k++) {
          Row t; V3 dir = k == 0 ? t1 : t2;
          add_side(t, A, c.posA, dir, false, 1.0); add_side(t, B, c.posB, dir, false, -1.0); finish_row(t);
          t.rhs = -row_vel(t, rob.qd) * t.invd; t.mu = c.roll; t.nidx = nidx; tr.push_back(t);
        }
      }
      for (int k = 0; k < 2; k++) {
        Row t; V3 dir = k == 0 ? t1 : t2;
        add_side(t, A, c.posA, dir, true, 1.0); add_side(t, B, c.posB, dir, true, -1.0); finish_row(t);
        t.rhs = -row_vel(t, rob.qd) * t.invd; t.mu = c.mu; t.nidx = nidx; fr.push_back(t);
      }
    }
    // --- projected Gauss-Seidel, Bullet's sweep order (row P6)
    DV dv; for (int d = 0; d < NDOF; d++) dv.r[d] = 0;
    int it = 0;
    for (; it < (int)bp().solver_iters; it++) {
      double res = 0;
      int nn = (int)nc.size();
      for (int j = 0; j < nn; j++) { int idx = (it & 1) ? j : nn - 1 - j; double r = resolve(nc[idx], dv); res = std::fmax(res, r * r); }
      for (auto& r : nb) { double x = resolve(r, dv); res = std::fmax(res, x * x); }
      for (auto& r : nr) { double x = resolve(r, dv); res = std::fmax(res, x * x); }
      for (auto& t : tr) {
        double tot = nr[t.nidx].applied;
        if (tot > 0) { t.lo = -t.mu * tot; t.hi = t.mu * tot; double x = resolve(t, dv); res = std::fmax(res, x * x); }
      }
      for (size_t j = 0; j + 1 < fr.size(); j += 2) {  // implicit friction cone over the two tangent rows [recalled]
        Row& a = fr[j]; Row& b = fr[j + 1];
        double lim = a.mu * nr[a.nidx].applied;
        if (bp().friction_cone == 0.0) {  // ablation: independent box limits, row after row
          a.lo = -lim; a.hi = lim; b.lo = -lim; b.hi = lim;
          double x = resolve(a, dv), y = resolve(b, dv);
          res = std::fmax(res, std::fmax(x * x, y * y));
          continue;
        }
        double da = a.rhs - row_dv(a, dv) * a.invd, db = b.rhs - row_dv(b, dv) * b.invd;
        double sa = a.applied + da, sb = b.applied + db;
        if (sa < -lim || sa > lim || sb < -lim || sb > lim) {
          double ang = std::atan2(sa, sb);
          double ca = std::fabs(lim * std::sin(ang)), cb = std::fabs(lim * std::cos(ang));
          if (sa < -ca) { da = -ca - a.applied; a.applied = -ca; } else if (sa > ca) { da = ca - a.applied; a.applied = ca; } else a.applied = sa;
          if (sb < -cb) { db = -cb - b.applied; b.applied = -cb; } else if (sb > cb) { db = cb - b.applied; b.applied = cb; } else b.applied = sb;
        } else { a.applied = sa; b.applied = sb; }
        row_apply(a, dv, da); row_apply(b, dv, db);
        double ra = a.invd > 0 ? da / a.invd : 0.0, rb = b.invd > 0 ? db / b.invd : 0.0;
        res = std::fmax(res, std::fmax(ra * ra, rb * rb));
      }
      if (res <= BP_RESIDUAL_THRESHOLD) { it++; break; }
    }
    last_iters = it;
    { long R = (long)(nc.size() + nb.size() + nr.size() + tr.size() + fr.size()); st_ticks++; st_iters += it; st_rows += R; st_contacts += (long)contacts.size(); st_rowiters += R * it; st_hist_rows[R < 191 ? R : 191]++; st_hist_nc[contacts.size() < 63 ? contacts.size() : 63]++; st_hist_roll[tr.size() / 2 < 15 ? tr.size() / 2 : 15]++; st_hist_nnc[nc.size() < 31 ? nc.size() : 31]++; }
    // --- apply and integrate (semi-implicit Euler; exponential map for free-body orientation)
    for (int d = 0; d < NDOF; d++) {
      rob.qd[d] = std::fmax(-BP_MAX_COORD_VEL, std::fmin(BP_MAX_COORD_VEL, rob.qd[d] + dv.r[d]));
      rob.q[d] += dt * rob.qd[d];
    }
    for (int b = 0; b < sc.nfree; b++) {
      if (!fb[b].active) continue;
      FreeBody& B = fb[b];
      B.lin += dv.l[b]; B.ang += dv.a[b];
      B.pos += B.lin * dt;
      if (sc.anchor[b].kind == 2) continue;
      double wn = norm(B.ang);
      if (wn * dt > BP_ANGULAR_MOTION_THRESHOLD) wn = BP_ANGULAR_MOTION_THRESHOLD / dt;
      V3 ax;
      if (wn < 0.001) ax = B.ang * (0.5 * dt - dt * dt * dt * 0.020833333333 * wn * wn);
      else ax = B.ang * (std::sin(0.5 * wn * dt) / wn);
      B.quat = normalized(Q4(ax.x, ax.y, ax.z, std::cos(0.5 * wn * dt)) * B.quat);
    }
  }

  // calculateInverseKinematics as the reference calls it (robot.py:119-121): DLS with orientation, no null space,
  // seeded from the current joint angles, <= 20 iterations [recalled]. Returns 6 arm targets.
  void ik(V3 tpos, Q4 tq, double* out6) {
    Robot r = rob;
    tq = normalized(tq);  // pybullet moves the target into the base frame through a btTransform, which normalises it [recalled]
    double diff = 1e30;
    for (int it = 0; it < BP_IK_ITERS && diff > BP_IK_RESIDUAL; it++) {
      r.fk();
      V3 ep = r.p[7]; Q4 eq = r.quat[7];
      V3 dp = tpos - ep;
      diff = norm(dp);
      Q4 dq = tq * conj(eq);  // target * current^-1
      double w = std::fmax(-1.0, std::fmin(1.0, dq.w));
      double ang = (double)(float)(2.0 * std::acos(w));
      double s2 = 1.0 - dq.w * dq.w;
      V3 ax = s2 < 10.0 * 2.220446049250313e-16 ? V3(1, 0, 0) : V3(dq.x, dq.y, dq.z) * (1.0 / std::sqrt(s2));
      if (ang > 3.14159265358979323846) ang -= 2.0 * 3.14159265358979323846;
      double an = norm(ax); V3 dr = ax * (ang / an);
      double e[6] = {dp.x, dp.y, dp.z, dr.x, dr.y, dr.z};
      double J[6][6];
      for (int d = 0; d < 6; d++) {
        int j = MOV[d];
        V3 jl = cross(r.axis[j], ep - r.p[j]);
        J[0][d] = jl.x; J[1][d] = jl.y; J[2][d] = jl.z; J[3][d] = r.axis[j].x; J[4][d] = r.axis[j].y; J[5][d] = r.axis[j].z;
      }
      double A[6][7];
      for (int a = 0; a < 6; a++) {
        for (int b = 0; b < 6; b++) { double s = 0; for (int k = 0; k < 6; k++) s += J[k][a] * J[k][b]; A[a][b] = s + (a == b ? bp().ik_damping : 0.0); }
        double s = 0; for (int k = 0; k < 6; k++) s += J[k][a] * e[k]; A[a][6] = s;
      }
      for (int a = 0; a < 6; a++) {  // Gaussian elimination with partial pivoting
        int pv = a; for (int b = a + 1; b < 6; b++) if (std::fabs(A[b][a]) > std::fabs(A[pv][a])) pv = b;
        if (pv != a) for (int k = 0; k < 7; k++) std::swap(A[a][k], A[pv][k]);
        for (int b = a + 1; b < 6; b++) { double f = A[b][a] / A[a][a]; for (int k = a; k < 7; k++) A[b][k] -= f * A[a][k]; }
      }
      double dth[6];
      for (int a = 5; a >= 0; a--) { double s = A[a][6]; for (int b = a + 1; b < 6; b++) s -= A[a][b] * dth[b]; dth[a] = s / A[a][a]; }
      double mx = 0; for (int a = 0; a < 6; a++) mx = std::fmax(mx, std::fabs(dth[a]));
      double sc_ = mx > BP_IK_MAX_STEP ? BP_IK_MAX_STEP / mx : 1.0;
      for (int a = 0; a < 6; a++) r.q[a] += dth[a] * sc_;
    }
    for (int a = 0; a < 6; a++) out6[a] = r.q[a];
  }

  // Union of the world AABBs of links 11 and 16, then midpoint (env_ball_balance_v0.py:249-275; row A7).
  V3 gripper_center() {
    rob.fk();
    V3 lo(1e300, 1e300, 1e300), hi(-1e300, -1e300, -1e300);
    const int ids[2] = {11, 16};
    for (int t = 0; t < 2; t++) {
      const UrgLinkModel& L = URG_LINKS[ids[t]];
      const M3& R = rob.R[ids[t]];
      V3 c = rob.p[ids[t]] + R * V3(L.aabb_center[0], L.aabb_center[1], L.aabb_center[2]);
      for (int i = 0; i < 3; i++) {
        double e = std::fabs(R(i, 0)) * L.aabb_half[0] + std::fabs(R(i, 1)) * L.aabb_half[1] + std::fabs(R(i, 2)) * L.aabb_half[2];
        lo[i] = std::fmin(lo[i], c[i] - e); hi[i] = std::fmax(hi[i], c[i] + e);
      }
    }
    return (lo + hi) * 0.5;
  }
};

}  // namespace urgo

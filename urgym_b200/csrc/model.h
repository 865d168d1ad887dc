// Device-side constant model: the UR5 + Robotiq-85 tree with fixed joints folded into their movable ancestor
// (12 bodies / 12 DoF), per-sub-link data for Bullet's per-link damping, collision primitives and the pair list.
// Built once on the host by build_model() from the generated link table (model_data.h, from
// /root/reference/urgym/urdf/ur5_robotiq_85.urdf via tools/compile_model.py) and uploaded to device memory.
#pragma once
#include <stdint.h>

#include "real.h"

#define UG_NDOF 12
#define UG_NLINK 19
#define UG_MAXFREE 3
#define UG_MAXSHAPES 32
#define UG_MAXPAIRS 160
#define UG_HITWORDS ((UG_MAXPAIRS + 31) / 32)
// Manifold-point capacity of a tick is a per-scene template constant of the kernel (kernels.cuh: EnvNC, 16 / 24 / 32 from
// the measured distributions, tools/contact_hist.py); UG_MAXC is the width of the contact-point export and the largest of them.
#define UG_MAXC 32
#define UG_STATE 96
#define UG_ENV_COUNT 8  // 0 BallBalance 1 TwoBallsBalance 2 CubesPush 3 CubesGrasp 4 Golf 5 CubesPushDiff 6 GolfJoints 7 Box

// Convex polytope of a gripper link whose collision geometry is a mesh (model_data.h: URG_POLYS), in the link frame.
#define UG_NPOLY 5
#define UG_POLY_MAXV 16
#define UG_POLY_MAXF 32
#define UG_POLY_MAXL 8
#define UG_POLY_MAXE 12
struct DevPoly {
  int nv, nf, ne, pad;
  real v[UG_POLY_MAXV][3]; real fn[UG_POLY_MAXF][4];   // vertices; face planes n . x = d (outward)
  uint8_t fcount[UG_POLY_MAXF]; uint8_t floop[UG_POLY_MAXF][UG_POLY_MAXL];  // CCW vertex loop per face
  real ed[UG_POLY_MAXE][3];                             // edge directions used for the edge-edge separating axes
};

struct DevShape {
  int kind;    // 0 box, 1 sphere, 2 convex polytope (robot link; lc / lR / half then hold its bounding box for the mid-phase)
  int poly;    // kind 2: index into DevModel::polys; the polytope lives in the frame of Bullet link `link`
  int body;    // 0 robot, 1..3 free body, -1 static
  int link;    // Bullet link index (robot) or sub-shape id
  int rbody;   // robot shapes: index of the movable body (dof) that carries the link
  real lc[3]; real lR[9];  // local centre / rotation in the carrying body's frame (static: world)
  real half[3]; real radius;
  real mu, rest, roll, cbt, brad;
};

struct DevSub { real mass; real com[3]; real I[6]; };  // I: xx xy xz yy yz zz about the sub-link COM, body frame

struct DevModel {
  int par[UG_NDOF]; uint32_t anc[UG_NDOF]; uint32_t sub[UG_NDOF]; int axis[UG_NDOF];
  real jpos[UG_NDOF][3]; real jrot[UG_NDOF][9];
  real mass[UG_NDOF]; real com[UG_NDOF][3]; real I[UG_NDOF][6];
  int nsub[UG_NDOF]; DevSub subl[UG_NDOF][3];
  real lower[UG_NDOF], upper[UG_NDOF], effort[UG_NDOF], velocity[UG_NDOF];
  int link_body[UG_NLINK]; real link_lpos[UG_NLINK][3]; real link_lrot[UG_NLINK][9];
  real link_jquat[UG_NLINK][4]; real link_jxyz[UG_NLINK][3]; int link_parent[UG_NLINK]; int link_dof[UG_NLINK]; int link_axis[UG_NLINK];
  real aabb_c[2][3], aabb_h[2][3];  // collision AABBs of links 11 and 16 in their link frames
  int env_type, nfree, nshapes, npairs, obs_dim, act_dim, max_steps; real gravity;
  int generic_sweeps;  // 1: contact-free ticks take the generic sweeps too (URGYM_GENERIC_SWEEPS=1 at create; comparison runs and tests)
  real fmass[UG_MAXFREE]; real finertia[UG_MAXFREE][3];
  // A body slot is a free rigid body (fanchor 0) or a world-anchored 1-DoF body of Box-v0's second articulated body
  // (skew-box-button.urdf: both joints hang off the fixed base): 1 = hinge about the world axis faxis through the body origin
  // forigin, 2 = slider along faxis from forigin. Generalized inverse inertia: faxis faxis^T / finertia_axis (angular) or
  // faxis faxis^T / fmass (linear). flower / fupper: joint limits; fcom: hinge body's merged COM (gravity torque); fmotor:
  // force limit of the slider's position motor to coordinate 0 (env_box_v0.py:176), 0 = none.
  int fanchor[UG_MAXFREE]; real faxis[UG_MAXFREE][3], forigin[UG_MAXFREE][3], fcom[UG_MAXFREE][3];
  real finertia_axis[UG_MAXFREE], flower[UG_MAXFREE], fupper[UG_MAXFREE], fmotor[UG_MAXFREE];
  DevShape shapes[UG_MAXSHAPES]; uint8_t pairs[UG_MAXPAIRS][2];
  // broad-phase bound per pair: the pair can only produce manifold points (separating-axis slack <= margin on all 15 axes)
  // while the centre distance (static partner: distance from the centre to the static box) is <= cullr (model_build.cpp)
  real cullr[UG_MAXPAIRS];
  DevPoly polys[UG_NPOLY];
};

// Host: fills *m for env_type 0..UG_ENV_COUNT-1. Returns 0 or -1.
int build_model(int env_type, DevModel* m);

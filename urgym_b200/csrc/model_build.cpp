// Host-side one-off model compile ("a host-side, one-off model compile into a constant table").
// Folds the seven fixed joints of /root/reference/urgym/urdf/ur5_robotiq_85.urdf into their movable ancestors and
// lays out the collision primitives and pair list per env type (env constructors, e.g.
// /root/reference/urgym/envs/env_ball_balance_v0.py:69-71,177-203; env_golf_v0.py:87,289-358).
#include <cmath>
#include <cstring>
#include <vector>

#include "model.h"
#include "model_data.h"
#include "params.h"

namespace {

struct D3 { double v[3]; };
struct DM { double m[9]; };
DM mm(const DM& a, const DM& b) { DM c; for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += a.m[3 * i + k] * b.m[3 * k + j]; c.m[3 * i + j] = s; } return c; }
D3 mv(const DM& a, const D3& x) { D3 r; for (int i = 0; i < 3; i++) r.v[i] = a.m[3 * i] * x.v[0] + a.m[3 * i + 1] * x.v[1] + a.m[3 * i + 2] * x.v[2]; return r; }
DM ident() { DM r; for (int i = 0; i < 9; i++) r.m[i] = (i % 4 == 0); return r; }

const int MOV[UG_NDOF] = {1, 2, 3, 4, 5, 6, 9, 11, 13, 14, 16, 18};
const int LINK_ORDER[17] = {12, 17, 11, 16, 10, 15, 9, 14, 13, 18, 8, 6, 5, 4, 3, 2, 1};

void mat_to_quat(const DM& R, double* q) {
  double tr = R.m[0] + R.m[4] + R.m[8];
  if (tr > 0) { double s = std::sqrt(tr + 1.0); q[3] = s * 0.5; s = 0.5 / s; q[0] = (R.m[7] - R.m[5]) * s; q[1] = (R.m[2] - R.m[6]) * s; q[2] = (R.m[3] - R.m[1]) * s; }
  else {
    int i = R.m[0] < R.m[4] ? (R.m[4] < R.m[8] ? 2 : 1) : (R.m[0] < R.m[8] ? 2 : 0);
    int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = std::sqrt(R.m[4 * i] - R.m[4 * j] - R.m[4 * k] + 1.0);
    q[i] = s * 0.5; s = 0.5 / s;
    q[3] = (R.m[3 * k + j] - R.m[3 * j + k]) * s; q[j] = (R.m[3 * j + i] + R.m[3 * i + j]) * s; q[k] = (R.m[3 * k + i] + R.m[3 * i + k]) * s;
  }
}

DevShape mk_box(int body, int link, const double* lc, const double* half, double mu, double rest = 0, double roll = 0) {
  DevShape s; std::memset(&s, 0, sizeof(s));
  s.kind = 0; s.body = body; s.link = link; s.rbody = -1;
  for (int i = 0; i < 3; i++) { s.lc[i] = (real)lc[i]; s.half[i] = (real)half[i]; }
  for (int i = 0; i < 9; i++) s.lR[i] = (i % 4 == 0);
  double n = std::sqrt(half[0] * half[0] + half[1] * half[1] + half[2] * half[2]);
  s.mu = (real)mu; s.rest = (real)rest; s.roll = (real)roll; s.cbt = (real)(UGP_CONTACT_BREAKING * n); s.brad = (real)n;
  return s;
}
DevShape mk_sphere(int body, int link, double r, double mu, double rest = 0, double roll = 0) {
  double z[3] = {0, 0, 0}, h[3] = {r, r, r};
  DevShape s = mk_box(body, link, z, h, mu, rest, roll);
  s.kind = 1; s.radius = (real)r; s.cbt = (real)(UGP_CONTACT_BREAKING * r); s.brad = (real)r;
  return s;
}
void box_inertia(double m, const double* h, real* out) {
  double lx = 2 * h[0], ly = 2 * h[1], lz = 2 * h[2];
  out[0] = (real)(m / 12 * (ly * ly + lz * lz)); out[1] = (real)(m / 12 * (lx * lx + lz * lz)); out[2] = (real)(m / 12 * (lx * lx + ly * ly));
}

}  // namespace

int build_model(int env_type_full, DevModel* m) {
  if (env_type_full < 0 || env_type_full >= UG_ENV_COUNT) return -1;
  // CubesPushDiff (5) and GolfJoints (6) live in their parents' worlds (env_cubes_push_diff_v0.py, env_golf_joints_v0.py)
  const int env_type = env_type_full == 5 ? 2 : (env_type_full == 6 ? 4 : env_type_full);
  std::memset(m, 0, sizeof(*m));
  const double base[3] = {0, env_type == 7 ? 0.5 : 0.0, 0};   // robot base: env_box_v0.py:55 (0, 0.5, 0), every other env the origin
  int dof_of_link[UG_NLINK];
  for (int i = 0; i < UG_NLINK; i++) dof_of_link[i] = -1;
  for (int d = 0; d < UG_NDOF; d++) dof_of_link[MOV[d]] = d;
  // carrying body and body-frame-local transform of every link
  int carrier[UG_NLINK]; DM lrot[UG_NLINK]; D3 lpos[UG_NLINK];
  for (int i = 0; i < UG_NLINK; i++) {
    const UrgLinkModel& L = URG_LINKS[i];
    DM jr; for (int k = 0; k < 9; k++) jr.m[k] = L.jrot[k];
    D3 jx = {{L.jxyz[0], L.jxyz[1], L.jxyz[2]}};
    if (dof_of_link[i] >= 0) { carrier[i] = dof_of_link[i]; lrot[i] = ident(); lpos[i] = D3{{0, 0, 0}}; }
    else {
      int p = L.parent;
      DM pr = p >= 0 ? lrot[p] : ident(); D3 pp = p >= 0 ? lpos[p] : D3{{0, 0, 0}};
      carrier[i] = p >= 0 ? carrier[p] : -1;
      D3 o = mv(pr, jx);
      lpos[i] = D3{{pp.v[0] + o.v[0], pp.v[1] + o.v[1], pp.v[2] + o.v[2]}};
      lrot[i] = mm(pr, jr);
    }
    m->link_body[i] = carrier[i];
    for (int k = 0; k < 3; k++) m->link_lpos[i][k] = (real)lpos[i].v[k];
    for (int k = 0; k < 9; k++) m->link_lrot[i][k] = (real)lrot[i].m[k];
    double q[4]; mat_to_quat(jr, q);
    for (int k = 0; k < 4; k++) m->link_jquat[i][k] = (real)q[k];
    for (int k = 0; k < 3; k++) m->link_jxyz[i][k] = (real)L.jxyz[k];
    m->link_parent[i] = L.parent; m->link_dof[i] = dof_of_link[i];
    m->link_axis[i] = L.axis[0] > 0.5 ? 0 : (L.axis[1] > 0.5 ? 1 : 2);
  }
  for (int d = 0; d < UG_NDOF; d++) {
    int li = MOV[d]; const UrgLinkModel& L = URG_LINKS[li];
    int pl = L.parent;
    int pb = pl >= 0 ? carrier[pl] : -1;
    m->par[d] = pb;
    m->anc[d] = (pb >= 0 ? m->anc[pb] : 0u) | (1u << d);
    DM pr = pl >= 0 ? lrot[pl] : ident(); D3 pp = pl >= 0 ? lpos[pl] : D3{{0, 0, 0}};
    D3 jx = {{L.jxyz[0], L.jxyz[1], L.jxyz[2]}}; D3 o = mv(pr, jx);
    DM jr; for (int k = 0; k < 9; k++) jr.m[k] = L.jrot[k];
    DM r = mm(pr, jr);
    for (int k = 0; k < 3; k++) m->jpos[d][k] = (real)(pp.v[k] + o.v[k] + (pb < 0 ? base[k] : 0.0));
    for (int k = 0; k < 9; k++) m->jrot[d][k] = (real)r.m[k];
    if (!((L.axis[0] == 1 && L.axis[1] == 0 && L.axis[2] == 0) || (L.axis[0] == 0 && L.axis[1] == 1 && L.axis[2] == 0) || (L.axis[0] == 0 && L.axis[1] == 0 && L.axis[2] == 1))) return -1;
    m->axis[d] = L.axis[0] == 1 ? 0 : (L.axis[1] == 1 ? 1 : 2);
    m->lower[d] = (real)L.lower; m->upper[d] = (real)L.upper; m->effort[d] = (real)L.effort; m->velocity[d] = (real)L.velocity;
  }
  for (int d = 0; d < UG_NDOF; d++) for (int e = 0; e < UG_NDOF; e++) if (m->anc[e] >> d & 1) m->sub[d] |= 1u << e;
  // merged mass properties and per-sub-link data
  for (int d = 0; d < UG_NDOF; d++) {
    double M = 0, c[3] = {0, 0, 0};
    std::vector<int> subs;
    for (int i = 0; i < UG_NLINK; i++) if (carrier[i] == d) subs.push_back(i);
    if (subs.size() > 3) return -1;
    m->nsub[d] = (int)subs.size();
    double sc[3][3], sI[3][9];
    for (size_t s = 0; s < subs.size(); s++) {
      int i = subs[s]; const UrgLinkModel& L = URG_LINKS[i];
      D3 cl = {{L.com[0], L.com[1], L.com[2]}}; D3 cb = mv(lrot[i], cl);
      for (int k = 0; k < 3; k++) { sc[s][k] = lpos[i].v[k] + cb.v[k]; c[k] += L.mass * sc[s][k]; }
      M += L.mass;
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) { double v = 0; for (int k = 0; k < 3; k++) v += lrot[i].m[3 * a + k] * L.inertia[k] * lrot[i].m[3 * b + k]; sI[s][3 * a + b] = v; }
      DevSub& S = m->subl[d][s];
      S.mass = (real)L.mass; for (int k = 0; k < 3; k++) S.com[k] = (real)sc[s][k];
      S.I[0] = (real)sI[s][0]; S.I[1] = (real)sI[s][1]; S.I[2] = (real)sI[s][2]; S.I[3] = (real)sI[s][4]; S.I[4] = (real)sI[s][5]; S.I[5] = (real)sI[s][8];
    }
    for (int k = 0; k < 3; k++) c[k] /= M;
    double I[9] = {0};
    for (size_t s = 0; s < subs.size(); s++) {
      double ms = URG_LINKS[subs[s]].mass, r[3] = {sc[s][0] - c[0], sc[s][1] - c[1], sc[s][2] - c[2]};
      double rr = r[0] * r[0] + r[1] * r[1] + r[2] * r[2];
      for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) I[3 * a + b] += sI[s][3 * a + b] + ms * ((a == b ? rr : 0.0) - r[a] * r[b]);
    }
    m->mass[d] = (real)M; for (int k = 0; k < 3; k++) m->com[d][k] = (real)c[k];
    m->I[d][0] = (real)I[0]; m->I[d][1] = (real)I[1]; m->I[d][2] = (real)I[2]; m->I[d][3] = (real)I[4]; m->I[d][4] = (real)I[5]; m->I[d][5] = (real)I[8];
  }
  const int aabb_links[2] = {11, 16};
  for (int t = 0; t < 2; t++) for (int k = 0; k < 3; k++) { m->aabb_c[t][k] = (real)URG_LINKS[aabb_links[t]].aabb_center[k]; m->aabb_h[t][k] = (real)URG_LINKS[aabb_links[t]].aabb_half[k]; }

  // ---- scene (same shape and pair order as the oracle's make_scene; the order is part of the spec because the
  // contact cap keeps the first UG_MAXC manifold points)
  m->env_type = env_type_full; m->gravity = (real)-9.8;
  int ns = 0;
  for (int k = 0; k < 17; k++) {
    int li = LINK_ORDER[k]; const UrgLinkModel& L = URG_LINKS[li];
    DM br; for (int i = 0; i < 9; i++) br.m[i] = L.box_rot[i];
    D3 bc = {{L.box_center[0], L.box_center[1], L.box_center[2]}};
    D3 c = mv(lrot[li], bc); DM R = mm(lrot[li], br);
    double lc[3] = {lpos[li].v[0] + c.v[0], lpos[li].v[1] + c.v[1], lpos[li].v[2] + c.v[2]};
    DevShape s = mk_box(0, li, lc, L.box_half, 0.5);
    for (int i = 0; i < 9; i++) s.lR[i] = (real)R.m[i];
    s.rbody = carrier[li];
    if (URG_LINK_POLY[li] >= 0) { s.kind = 2; s.poly = URG_LINK_POLY[li]; }  // mesh geometry: convex polytope, the box is its bound
    m->shapes[ns++] = s;
  }
  const double table_c[3] = {0, 0, -0.25}, table_h[3] = {0.5, 1.0, 0.25}, zero[3] = {0, 0, 0};
  // plane.urdf at z = -0.5 under the table (env_ball_balance_v0.py:70, env_cubes_push_v0.py:77, env_golf_v0.py:77): 30 x 30 x 10
  // box collider 5 m below its origin, lateral friction 1 [recalled]; static id 2
  const double ground_c[3] = {0, 0, -0.5 - 5.0}, ground_h[3] = {15, 15, 5};
  const double paddle_h[3] = {0.05, 0.1, 0.005}, cube_h[3] = {0.02, 0.02, 0.02};
  switch (env_type) {
    case 0:
      m->nfree = 2; m->obs_dim = 6; m->act_dim = 3; m->max_steps = 500;
      m->fmass[0] = (real)0.1; box_inertia(0.1, paddle_h, m->finertia[0]);
      m->fmass[1] = (real)0.01; for (int k = 0; k < 3; k++) m->finertia[1][k] = (real)(0.4 * 0.01 * 0.02 * 0.02);
      m->shapes[ns++] = mk_box(1, -1, zero, paddle_h, 0.5);
      m->shapes[ns++] = mk_sphere(2, -1, 0.02, 0.5);
      m->shapes[ns++] = mk_box(-1, 0, table_c, table_h, 0.5);
      m->shapes[ns++] = mk_box(-1, 2, ground_c, ground_h, 1.0);
      break;
    case 1:
      m->nfree = 3; m->obs_dim = 9; m->act_dim = 3; m->max_steps = 500;
      m->fmass[0] = (real)0.1; box_inertia(0.1, paddle_h, m->finertia[0]);
      for (int b = 1; b < 3; b++) { m->fmass[b] = (real)0.01; for (int k = 0; k < 3; k++) m->finertia[b][k] = (real)(0.4 * 0.01 * 0.003 * 0.003); }
      m->shapes[ns++] = mk_box(1, -1, zero, paddle_h, 0.5);
      m->shapes[ns++] = mk_sphere(2, -1, 0.003, 0.5);
      m->shapes[ns++] = mk_sphere(3, -1, 0.003, 0.5);
      m->shapes[ns++] = mk_box(-1, 0, table_c, table_h, 0.5);
      m->shapes[ns++] = mk_box(-1, 2, ground_c, ground_h, 1.0);
      break;
    case 2:
      m->nfree = 2; m->obs_dim = 9; m->act_dim = 3; m->max_steps = 50;
      for (int b = 0; b < 2; b++) { m->fmass[b] = (real)1.0; box_inertia(1.0, cube_h, m->finertia[b]); m->shapes[ns++] = mk_box(b + 1, -1, zero, cube_h, 1.0); }
      m->shapes[ns++] = mk_box(-1, 0, table_c, table_h, 0.5);
      m->shapes[ns++] = mk_box(-1, 2, ground_c, ground_h, 1.0);
      break;
    case 3: {
      m->nfree = 1; m->obs_dim = 17; m->act_dim = 9; m->max_steps = 20; m->gravity = (real)-10.0;
      m->fmass[0] = (real)1.0; box_inertia(1.0, cube_h, m->finertia[0]);
      m->shapes[ns++] = mk_box(1, -1, zero, cube_h, 1.0);
      const double pc[3] = {0, 0, -5.0}, ph[3] = {15, 15, 5};
      m->shapes[ns++] = mk_box(-1, 0, pc, ph, 1.0);
      break;
    }
    case 4: {
      m->nfree = 2; m->obs_dim = 9; m->act_dim = 3; m->max_steps = 100;
      double zc = UGP_CLUB_ZC;
      const double bh[3] = {0.015, 0.015, 0.075}, th[3] = {0.05, 0.015, 0.015}, dh[3] = {0.04, 0.04, 0.005};
      real ib[3], it[3]; box_inertia(1.0, bh, ib); box_inertia(1.0, th, it);
      double idx = 1.0 / 12 * 0.01 * 0.01 + 0.25 * 0.04 * 0.04, idz = 0.5 * 0.04 * 0.04;
      double d0 = 0 - zc, d1 = -0.09 - zc, d2 = 0.08 - zc, pa = d0 * d0 + d1 * d1 + d2 * d2;
      m->fmass[0] = (real)3.0;
      m->finertia[0][0] = (real)((double)ib[0] + it[0] + idx + pa); m->finertia[0][1] = (real)((double)ib[1] + it[1] + idx + pa); m->finertia[0][2] = (real)((double)ib[2] + it[2] + idz);
      m->fmass[1] = (real)0.01; for (int k = 0; k < 3; k++) m->finertia[1][k] = (real)(0.4 * 0.01 * 0.015 * 0.015);
      const double c0[3] = {0, 0, d0}, c1[3] = {0, 0, d1}, c2[3] = {0, 0, d2};
      m->shapes[ns++] = mk_box(1, -1, c0, bh, 10.0, 0.1);
      m->shapes[ns++] = mk_box(1, 0, c1, th, 0.5);
      m->shapes[ns++] = mk_box(1, 1, c2, dh, 0.5);
      m->shapes[ns++] = mk_sphere(2, -1, 0.015, 0.5, 0.1, 10.0);
      m->shapes[ns++] = mk_box(-1, 0, table_c, table_h, 0.5);
      const double pc[3] = {0, -0.6, 0.02};
      m->shapes[ns++] = mk_box(-1, 1, pc, URG_PITCH_HALF, 0.5);
      m->shapes[ns++] = mk_box(-1, 2, ground_c, ground_h, 1.0);
      break;
    }
    case 7: {  // Box-v0: env_box_v0.py:55-85, urgym/urdf/skew-box-button.urdf; same shape order as the oracle's make_scene
      m->nfree = 2; m->obs_dim = 7; m->act_dim = 8; m->max_steps = 200; m->gravity = (real)-10.0;
      const double mh1 = 0.05, mh3 = 0.1, md = 0.1 + 2 * mh1 + mh3;
      m->fanchor[0] = 1; m->faxis[0][1] = (real)1; m->forigin[0][0] = (real)0.075; m->forigin[0][2] = (real)0.1; m->flower[0] = (real)0; m->fupper[0] = (real)2;
      m->finertia_axis[0] = (real)(0.5 + 2 * mh1 * (0.13 * 0.13 + 0.025 * 0.025) + mh3 * (0.13 * 0.13 + 0.05 * 0.05));
      m->fmass[0] = (real)md; for (int k = 0; k < 3; k++) m->finertia[0][k] = (real)1;
      m->fcom[0][0] = (real)(-0.13 * (2 * mh1 + mh3) / md); m->fcom[0][2] = (real)((2 * mh1 * 0.025 + mh3 * 0.05) / md);
      m->fanchor[1] = 2; m->faxis[1][2] = (real)1; m->forigin[1][2] = (real)0.07; m->flower[1] = (real)-0.05; m->fupper[1] = (real)0; m->fmotor[1] = (real)1;
      m->fmass[1] = (real)0.01; for (int k = 0; k < 3; k++) m->finertia[1][k] = (real)1;
      const double c0[3] = {-0.076, 0, 0}, h0[3] = {0.08, 0.1, 0.005}, c1[3] = {-0.13, 0.05, 0.025}, c2[3] = {-0.13, -0.05, 0.025}, hp[3] = {0.005, 0.005, 0.025};
      const double c3[3] = {-0.13, 0, 0.05}, h3[3] = {0.01, 0.05, 0.01}, hb[3] = {0.02, 0.02, 0.025};
      m->shapes[ns++] = mk_box(1, 0, c0, h0, 0.5); m->shapes[ns++] = mk_box(1, 1, c1, hp, 0.5); m->shapes[ns++] = mk_box(1, 2, c2, hp, 0.5);
      m->shapes[ns++] = mk_box(1, 3, c3, h3, 1.0); m->shapes[ns++] = mk_box(2, 0, zero, hb, 0.5);
      const double w1[3] = {0.075, 0, 0.045}, w2[3] = {-0.075, 0, 0.045}, wh[3] = {0.005, 0.1, 0.045}, w3[3] = {0, 0.1, 0.045}, w4[3] = {0, -0.1, 0.045}, wg[3] = {0.075, 0.005, 0.045};
      m->shapes[ns++] = mk_box(-1, 0, w1, wh, 0.5); m->shapes[ns++] = mk_box(-1, 1, w2, wh, 0.5);
      m->shapes[ns++] = mk_box(-1, 2, w3, wg, 0.5); m->shapes[ns++] = mk_box(-1, 3, w4, wg, 0.5);
      const double pc[3] = {0, 0, -5.0}, ph[3] = {15, 15, 5};
      m->shapes[ns++] = mk_box(-1, 4, pc, ph, 1.0);
      break;
    }
  }
  if (env_type_full == 6) { m->obs_dim = 15; m->act_dim = 6; }   // env_golf_joints_v0.py:60-63
  m->nshapes = ns;
  static_assert(URG_NPOLY == UG_NPOLY && URG_POLY_MAXV == UG_POLY_MAXV && URG_POLY_MAXF == UG_POLY_MAXF && URG_POLY_MAXL == UG_POLY_MAXL &&
                URG_POLY_MAXE == UG_POLY_MAXE, "model_data.h and model.h disagree on the polytope capacities");
  for (int k = 0; k < UG_NPOLY; k++) {
    const UrgPoly& P = URG_POLYS[k]; DevPoly& D = m->polys[k];
    D.nv = P.nv; D.nf = P.nf; D.ne = P.ne;
    for (int i = 0; i < P.nv; i++) for (int c = 0; c < 3; c++) D.v[i][c] = (real)P.v[i][c];
    for (int f = 0; f < P.nf; f++) {
      for (int c = 0; c < 4; c++) D.fn[f][c] = (real)P.fn[f][c];
      D.fcount[f] = (uint8_t)P.fcount[f];
      for (int c = 0; c < P.fcount[f]; c++) D.floop[f][c] = (uint8_t)P.floop[f][c];
    }
    for (int e = 0; e < P.ne; e++) for (int c = 0; c < 3; c++) D.ed[e][c] = (real)P.ed[e][c];
  }
  int np = 0;
  auto add = [&](int a, int b) { if (np < UG_MAXPAIRS) { m->pairs[np][0] = (uint8_t)a; m->pairs[np][1] = (uint8_t)b; } np++; };
  for (int a = 0; a < ns; a++) for (int b = a + 1; b < ns; b++) if (m->shapes[a].body > 0 && m->shapes[b].body > 0 && m->shapes[a].body != m->shapes[b].body) add(a, b);
  // Box-v0 (pair budget): door and button never reach the ground plane (static id 4); the upper arm links (the last four robot
  // shapes: 4, 3, 2, 1) are tested against the ground plane only — same rules as the oracle's make_scene
  auto plane7 = [&](int b) { return env_type == 7 && m->shapes[b].body < 0 && m->shapes[b].link == 4; };
  auto arm_only = [&](int a, int b) { return env_type == 7 && a >= 13 && !plane7(b); };
  for (int a = 0; a < ns; a++) for (int b = 0; b < ns; b++) if (m->shapes[a].body > 0 && m->shapes[b].body < 0 && !plane7(b)) add(a, b);
  for (int a = 0; a < ns; a++) for (int b = 0; b < ns; b++) if (m->shapes[a].body == 0 && m->shapes[b].body > 0 && !arm_only(a, b)) add(a, b);
  for (int a = 0; a < ns; a++) for (int b = 0; b < ns; b++) if (m->shapes[a].body == 0 && m->shapes[b].body < 0 && !arm_only(a, b)) add(a, b);
  if (np > UG_MAXPAIRS) return -1;
  m->npairs = np;
  // Broad-phase bound per pair. The narrowphase reports a box-box manifold when the separating-axis slack is <= margin on
  // all 15 axes, which can hold at a true distance above the margin, so the bounding spheres alone are not a safe cull.
  // With A's three face axes a_i: |dc . a_i| <= hA_i + pB_i + margin, pB_i = sum_j hB_j |b_j . a_i|, |pB| <= sqrt(3) rB
  // (Cauchy-Schwarz), hence |dc| <= rA + sqrt(3) rB + sqrt(3) margin, and the same with A and B exchanged. Against a static
  // (world-aligned) box the vector from the box to A's centre obeys |e| <= sqrt(3) (rA + margin). Spheres have exact distances.
  const double r3 = std::sqrt(3.0);
  for (int p = 0; p < np; p++) {
    const DevShape& A = m->shapes[m->pairs[p][0]]; const DevShape& B = m->shapes[m->pairs[p][1]];
    const double mg = std::fmin((double)A.cbt, (double)B.cbt), ra = A.brad, rb = B.brad;
    double bound;
    if (B.body < 0) bound = A.kind == 1 ? ra + mg : r3 * (ra + mg);
    else if (A.kind == 1 || B.kind == 1) bound = ra + rb + mg;
    else bound = ra + rb + (r3 - 1.0) * std::fmin(ra, rb) + r3 * mg;
    m->cullr[p] = (real)(bound * (1.0 + 1e-6));  // a hair of slack: the device evaluates the distance in `real`
  }
  return 0;
}

// TEST INFRASTRUCTURE — part of the CPU oracle (see oracle/README.md).
//
// Narrowphase for the primitive pairs the five in-scope envs produce (paddle / ball / cube / club / table / plane /
// pitch / robot-link boxes). Restates, in fp64, the behaviour pybullet's collision stage shows to the envs
// (SURVEY.md §8a row P5): box-box by 15-axis SAT + incident-face clipping with at most four manifold points
// [recalled: btBoxBoxDetector, after ODE], sphere-box and sphere-sphere closest points, and contacts reported while
// the separation is below the pair's contact-breaking threshold (so "touching" is slightly fatter than penetration,
// SURVEY.md §8a row A8). Stateless: contacts are a pure function of the poses.
#pragma once
#include "urgo_math.h"

namespace urgo {

// Smallest |x| over every inequality `x > 0` the narrowphase has decided since the caller last reset it (separating-axis
// tests against the margin, clip keep tests, sphere distance tests): how close the contact SET of a tick came to being a
// different one. Tests read it to tell an fp32 threshold flip from a disagreement (tests/test_gpu_parity.py).
inline double& decision_margin() { static thread_local double m = 1e300; return m; }
inline void note_decision(double x) { double a = std::fabs(x); if (a < decision_margin()) decision_margin() = a; }

struct BoxW { V3 c; M3 R; V3 h; };
// normal n points from B to A; posA = posB + n * dist; dist < 0 is penetration
struct CPoint { V3 posA, posB, n; double dist; };

inline double edge_bias(double s) { return s < 0 ? s * 1.05 : s / 1.05; }

// Sutherland-Hodgman clip of a polygon (x,y,z per vertex; z carried along) against |x|<=hx, |y|<=hy.
inline int clip_rect(const double (*in)[3], int n, double hx, double hy, double (*out)[3]) {
  double bufA[16][3], bufB[16][3];
  for (int i = 0; i < n; i++) for (int k = 0; k < 3; k++) bufA[i][k] = in[i][k];
  double (*src)[3] = bufA; double (*dst)[3] = bufB;
  for (int e = 0; e < 4; e++) {
    int ax = e >> 1; double sgn = (e & 1) ? -1.0 : 1.0; double lim = ax == 0 ? hx : hy;
    int m = 0;
    for (int i = 0; i < n; i++) {
      const double* P = src[i]; const double* Q = src[(i + 1) % n];
      double dp = lim - sgn * P[ax], dq = lim - sgn * Q[ax];  // >= 0 inside
      if (dp >= 0) { for (int k = 0; k < 3; k++) dst[m][k] = P[k]; m++; }
      if ((dp >= 0) != (dq >= 0)) {
        double t = dp / (dp - dq);
        for (int k = 0; k < 3; k++) dst[m][k] = P[k] + t * (Q[k] - P[k]);
        m++;
      }
    }
    n = m;
    double (*tmp)[3] = src; src = dst; dst = tmp;
    if (n == 0) break;
  }
  for (int i = 0; i < n; i++) for (int k = 0; k < 3; k++) out[i][k] = src[i][k];
  return n;
}

inline int box_box(const BoxW& A, const BoxW& B, double margin, CPoint out[4]) {
  V3 p = B.c - A.c;
  V3 pp = tmul(A.R, p);
  double R[3][3], Q[3][3];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) { R[i][j] = dot(A.R.col(i), B.R.col(j)); Q[i][j] = std::fabs(R[i][j]); }
  double best = -1e300; int code = 0; bool invert = false; V3 nrm;
  for (int i = 0; i < 3; i++) {  // A's face axes
    double e1 = pp[i];
    double e2 = A.h[i] + B.h[0] * Q[i][0] + B.h[1] * Q[i][1] + B.h[2] * Q[i][2];
    double s = std::fabs(e1) - e2;
    note_decision(s - margin);
    if (s > margin) return 0;
    if (s > best) { best = s; nrm = A.R.col(i); invert = e1 < 0; code = 1 + i; }
  }
  for (int j = 0; j < 3; j++) {  // B's face axes
    double e1 = dot(B.R.col(j), p);
    double e2 = A.h[0] * Q[0][j] + A.h[1] * Q[1][j] + A.h[2] * Q[2][j] + B.h[j];
    double s = std::fabs(e1) - e2;
    note_decision(s - margin);
    if (s > margin) return 0;
    if (s > best) { best = s; nrm = B.R.col(j); invert = e1 < 0; code = 4 + j; }
  }
  for (int i = 0; i < 3; i++) {  // edge x edge axes, in A's frame
    int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
    for (int j = 0; j < 3; j++) {
      int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      V3 L;  // e_i x (column j of R)
      L[i] = 0; L[i1] = -R[i2][j]; L[i2] = R[i1][j];
      double l = norm(L);
      if (l <= 1e-9) continue;
      double e1 = dot(pp, L);
      double e2 = A.h[i1] * Q[i2][j] + A.h[i2] * Q[i1][j] + B.h[j1] * Q[i][j2] + B.h[j2] * Q[i][j1];
      double s = (std::fabs(e1) - e2) / l;
      note_decision(s - margin);
      if (s > margin) return 0;
      if (edge_bias(s) > best) { best = s; nrm = A.R * (L * (1.0 / l)); invert = e1 < 0; code = 7 + 3 * i + j; }
    }
  }
  if (invert) nrm = -nrm;  // now points from A to B
  double depth = -best;
  if (code >= 7) {
    int i = (code - 7) / 3, j = (code - 7) % 3;
    V3 pa = A.c, pb = B.c;
    for (int k = 0; k < 3; k++) {
      pa += A.R.col(k) * ((dot(nrm, A.R.col(k)) > 0 ? 1.0 : -1.0) * A.h[k]);
      pb -= B.R.col(k) * ((dot(nrm, B.R.col(k)) > 0 ? 1.0 : -1.0) * B.h[k]);
    }
    V3 ua = A.R.col(i), ub = B.R.col(j);
    V3 d = pb - pa;
    double uaub = dot(ua, ub), q1 = dot(ua, d), q2 = -dot(ub, d);
    double den = 1.0 - uaub * uaub, alpha = 0, beta = 0;
    if (den > 1e-4) { alpha = (q1 + uaub * q2) / den; beta = (uaub * q1 + q2) / den; }
    pa += ua * alpha; pb += ub * beta;
    out[0].n = -nrm; out[0].dist = -depth; out[0].posB = pb; out[0].posA = pb + out[0].n * out[0].dist;
    return 1;
  }
  const BoxW& ref = code <= 3 ? A : B;
  const BoxW& inc = code <= 3 ? B : A;
  V3 nr = code <= 3 ? nrm : -nrm;  // out of the reference face, towards the incident box
  int k = (code - 1) % 3, c1 = (k + 1) % 3, c2 = (k + 2) % 3;
  int la = 0; double lav = -1;
  for (int j = 0; j < 3; j++) { double v = std::fabs(dot(nr, inc.R.col(j))); if (v > lav) { lav = v; la = j; } }
  int a1 = (la + 1) % 3, a2 = (la + 2) % 3;
  double sg = dot(nr, inc.R.col(la)) > 0 ? -1.0 : 1.0;
  V3 fc = inc.c + inc.R.col(la) * (sg * inc.h[la]) - ref.c;
  V3 e1v = inc.R.col(a1) * inc.h[a1], e2v = inc.R.col(a2) * inc.h[a2];
  V3 quad[4] = {fc - e1v - e2v, fc + e1v - e2v, fc + e1v + e2v, fc - e1v + e2v};
  V3 rc1 = ref.R.col(c1), rc2 = ref.R.col(c2);
  double poly[4][3], clipped[16][3];
  for (int v = 0; v < 4; v++) { poly[v][0] = dot(quad[v], rc1); poly[v][1] = dot(quad[v], rc2); poly[v][2] = dot(quad[v], nr); }
  int n = clip_rect(poly, 4, ref.h[c1], ref.h[c2], clipped);
  int keep[16], m = 0; double dep[16];
  for (int v = 0; v < n; v++) {
    double dv = ref.h[k] - clipped[v][2];
    note_decision(dv + margin);
    if (dv >= -margin) { keep[m] = v; dep[m] = dv; m++; }
  }
  if (m == 0) return 0;
  int sel[4], cnt = m;
  if (m > 4) {
    int i0 = 0;
    for (int v = 1; v < m; v++) if (dep[v] > dep[i0]) i0 = v;
    for (int j = 0; j < 4; j++) sel[j] = (i0 + (j * m) / 4) % m;
    cnt = 4;
  } else {
    for (int j = 0; j < m; j++) sel[j] = j;
  }
  for (int j = 0; j < cnt; j++) {
    const double* c = clipped[keep[sel[j]]];
    V3 pt = ref.c + rc1 * c[0] + rc2 * c[1] + nr * c[2];
    double dv = dep[sel[j]];
    out[j].n = -nrm; out[j].dist = -dv;
    if (code <= 3) { out[j].posB = pt; out[j].posA = pt + out[j].n * out[j].dist; }
    else { out[j].posA = pt; out[j].posB = pt - out[j].n * out[j].dist; }
  }
  return cnt;
}

// Sphere against box. Returns the box-surface point, the normal from box to sphere and the signed distance.
inline bool sphere_box(V3 sc, double r, const BoxW& B, double margin, V3& on_box, V3& n_box_to_sphere, double& dist) {
  V3 l = tmul(B.R, sc - B.c);
  V3 cl(std::fmin(std::fmax(l.x, -B.h.x), B.h.x), std::fmin(std::fmax(l.y, -B.h.y), B.h.y), std::fmin(std::fmax(l.z, -B.h.z), B.h.z));
  V3 d = l - cl;
  double len2 = dot(d, d);
  V3 nl;
  if (len2 > 1e-24) {
    double len = std::sqrt(len2);
    nl = d * (1.0 / len);
    dist = len - r;
  } else {  // centre inside the box: push out through the nearest face
    int ax = 0; double md = B.h[0] - std::fabs(l[0]);
    for (int i = 1; i < 3; i++) { double v = B.h[i] - std::fabs(l[i]); if (v < md) { md = v; ax = i; } }
    nl = V3(0, 0, 0); nl[ax] = l[ax] >= 0 ? 1.0 : -1.0;
    cl = l; cl[ax] = nl[ax] * B.h[ax];
    dist = -md - r;
  }
  note_decision(dist - margin);
  if (!(dist < margin)) return false;
  on_box = B.c + B.R * cl;
  n_box_to_sphere = B.R * nl;
  return true;
}

}  // namespace urgo

// TEST INFRASTRUCTURE — part of the CPU oracle (see oracle/README.md).
//
// Narrowphase for the primitive pairs the five in-scope envs produce (paddle / ball / cube / club / table / plane /
// pitch / robot-link boxes). Restates, in fp64, the behaviour pybullet's collision stage shows to the envs
// (SURVEY.md §8a row P5): box-box by 15-axis SAT + incident-face clipping with at most four manifold points
// [recalled: btBoxBoxDetector, after ODE], sphere-box and sphere-sphere closest points, and contacts reported while
// the separation is below the pair's contact-breaking threshold (so "touching" is slightly fatter than penetration,
// SURVEY.md §8a row A8). Stateless: contacts are a pure function of the poses.
#pragma once
#include "model_data.h"
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


// ----------------------------------------------------------------------------------------------------------------
// Convex polytope (a gripper link whose URDF collision geometry is a mesh: Bullet collides the convex hull of the mesh
// vertices [recalled: BulletUrdfImporter]) against a box or a sphere. Separating-axis test over the polytope's face
// normals, the box's face normals and the cross products of their edge directions, then clipping of the incident face
// against the side planes of the reference face: the published polyhedral contact-clipping scheme (Bullet's
// btPolyhedralContactClipping / ODE's box-box generalised to face lists), with the same acceptance rule as box_box
// above (points within `margin` of the reference face are kept, at most four, spread from the deepest).
struct PolyW { int nv, nf, ne; V3 v[URG_POLY_MAXV]; V3 fn[URG_POLY_MAXF]; double fd[URG_POLY_MAXF]; V3 ed[URG_POLY_MAXE]; const UrgPoly* src; };

inline PolyW poly_world(const UrgPoly& P, V3 origin, const M3& R) {
  PolyW w; w.nv = P.nv; w.nf = P.nf; w.ne = P.ne; w.src = &P;
  for (int i = 0; i < P.nv; i++) w.v[i] = origin + R * V3(P.v[i][0], P.v[i][1], P.v[i][2]);
  for (int f = 0; f < P.nf; f++) { w.fn[f] = R * V3(P.fn[f][0], P.fn[f][1], P.fn[f][2]); w.fd[f] = P.fn[f][3] + dot(w.fn[f], origin); }
  for (int e = 0; e < P.ne; e++) w.ed[e] = R * V3(P.ed[e][0], P.ed[e][1], P.ed[e][2]);
  return w;
}

// Clips the convex polygon `poly` (n <= 16 vertices) against the half-space m . (p - o) >= 0.
inline int clip_halfspace(V3* poly, int n, V3 o, V3 m) {
  V3 out[24]; int k = 0;
  for (int i = 0; i < n; i++) {
    V3 P = poly[i], Q = poly[(i + 1) % n];
    double dp = dot(m, P - o), dq = dot(m, Q - o);
    if (dp >= 0) out[k++] = P;
    if ((dp >= 0) != (dq >= 0)) { double t = dp / (dp - dq); out[k++] = P + (Q - P) * t; }
  }
  if (k > 16) k = 16;
  for (int i = 0; i < k; i++) poly[i] = out[i];
  return k;
}

inline int keep_four(const V3* pts, const double* dist, int m, int* sel) {
  if (m <= 4) { for (int j = 0; j < m; j++) sel[j] = j; return m; }
  int i0 = 0;
  for (int v = 1; v < m; v++) if (-dist[v] > -dist[i0]) i0 = v;   // deepest first
  for (int j = 0; j < 4; j++) sel[j] = (i0 + (j * m) / 4) % m;
  return 4;
}

// A = polytope (world), B = box. Normal n of the output points from B to A, posA = posB + n * dist.
inline int poly_box(const PolyW& A, const BoxW& B, double margin, CPoint out[4]) {
  V3 bax[3] = {B.R.col(0), B.R.col(1), B.R.col(2)};
  double best = -1e300; int kind = 0, idx = 0, sgn = 1; V3 nBA;  // kind 0: face of A, 1: face of B (idx axis, sgn side), 2: edge pair
  for (int f = 0; f < A.nf; f++) {                 // faces of A: lowest point of B above the face plane
    const V3 n = A.fn[f];
    double sB = dot(n, B.c) - (B.h.x * std::fabs(dot(n, bax[0])) + B.h.y * std::fabs(dot(n, bax[1])) + B.h.z * std::fabs(dot(n, bax[2])));
    double s = sB - A.fd[f];
    note_decision(s - margin);
    if (s > margin) return 0;
    if (s > best) { best = s; kind = 0; idx = f; nBA = -n; }
  }
  double lo[3], hi[3];
  for (int j = 0; j < 3; j++) {                    // faces of B: extent of A along the box axis, relative to the box centre
    lo[j] = 1e300; hi[j] = -1e300;
    for (int i = 0; i < A.nv; i++) { double p = dot(bax[j], A.v[i] - B.c); lo[j] = std::fmin(lo[j], p); hi[j] = std::fmax(hi[j], p); }
    double sp = lo[j] - B.h[j], sm = -hi[j] - B.h[j];   // A beyond the +j face / beyond the -j face
    note_decision(sp - margin); note_decision(sm - margin);
    if (sp > margin || sm > margin) return 0;
    if (sp > best) { best = sp; kind = 1; idx = j; sgn = 1; nBA = bax[j]; }
    if (sm > best) { best = sm; kind = 1; idx = j; sgn = -1; nBA = -bax[j]; }
  }
  int ea = 0, eb = 0;
  for (int e = 0; e < A.ne; e++)
    for (int j = 0; j < 3; j++) {                  // edge x edge
      V3 L = cross(A.ed[e], bax[j]); double l = norm(L);
      if (l <= 1e-9) continue;
      L = L * (1.0 / l);
      double mnA = 1e300, mxA = -1e300;
      for (int i = 0; i < A.nv; i++) { double p = dot(L, A.v[i] - B.c); mnA = std::fmin(mnA, p); mxA = std::fmax(mxA, p); }
      double rB = B.h.x * std::fabs(dot(L, bax[0])) + B.h.y * std::fabs(dot(L, bax[1])) + B.h.z * std::fabs(dot(L, bax[2]));
      double s1 = mnA - rB, s2 = -mxA - rB;        // A on the +L side of B / on the -L side
      double sx = std::fmax(s1, s2);
      note_decision(sx - margin);
      if (sx > margin) return 0;
      if (edge_bias(sx) > best) { best = sx; kind = 2; ea = e; eb = j; nBA = s1 >= s2 ? L : -L; }
    }
  if (kind == 2) {                                 // one point: closest points of the two supporting edges
    // supporting feature of each body towards the other: mean of the vertices within 10 um of the extreme
    double exA = 1e300; for (int i = 0; i < A.nv; i++) exA = std::fmin(exA, dot(nBA, A.v[i]));
    V3 pa; int ca = 0; for (int i = 0; i < A.nv; i++) if (dot(nBA, A.v[i]) <= exA + 1e-5) { pa += A.v[i]; ca++; }
    pa = pa * (1.0 / ca);
    V3 pb = B.c;
    for (int k = 0; k < 3; k++) if (k != eb) pb += bax[k] * ((dot(nBA, bax[k]) > 0 ? 1.0 : -1.0) * B.h[k]);
    V3 ua = A.ed[ea], ub = bax[eb], d = pb - pa;
    double uaub = dot(ua, ub), q1 = dot(ua, d), q2 = -dot(ub, d), den = 1.0 - uaub * uaub, beta = 0;
    if (den > 1e-4) beta = (uaub * q1 + q2) / den;
    pb += ub * beta;
    out[0].n = nBA; out[0].dist = best; out[0].posB = pb; out[0].posA = pb + nBA * best;
    return 1;
  }
  V3 poly[24]; int n = 0; V3 rn; double rd;         // reference plane rn . p = rd (outward)
  if (kind == 0) {                                 // reference face on A; incident = the box face most against it
    rn = A.fn[idx]; rd = A.fd[idx];
    int la = 0; double lav = -1; for (int j = 0; j < 3; j++) { double v = std::fabs(dot(rn, bax[j])); if (v > lav) { lav = v; la = j; } }
    int a1 = (la + 1) % 3, a2 = (la + 2) % 3;
    double sg = dot(rn, bax[la]) > 0 ? -1.0 : 1.0;
    V3 fc = B.c + bax[la] * (sg * B.h[la]), e1 = bax[a1] * B.h[a1], e2 = bax[a2] * B.h[a2];
    poly[0] = fc - e1 - e2; poly[1] = fc + e1 - e2; poly[2] = fc + e1 + e2; poly[3] = fc - e1 + e2; n = 4;
    const int cnt = A.src->fcount[idx]; const int* loop = A.src->floop[idx];
    for (int k = 0; k < cnt && n > 0; k++) {
      V3 o = A.v[loop[k]], e = A.v[loop[(k + 1) % cnt]] - o;
      n = clip_halfspace(poly, n, o, cross(rn, e));
    }
  } else {                                         // reference face on B; incident = A's face most against it
    rn = bax[idx] * (double)sgn; rd = dot(rn, B.c) + B.h[idx];
    int fi = 0; double fv = 1e300; for (int f = 0; f < A.nf; f++) { double v = dot(A.fn[f], rn); if (v < fv) { fv = v; fi = f; } }
    const int cnt = A.src->fcount[fi]; const int* loop = A.src->floop[fi];
    for (int k = 0; k < cnt; k++) poly[n++] = A.v[loop[k]];
    int c1 = (idx + 1) % 3, c2 = (idx + 2) % 3;
    for (int t = 0; t < 4 && n > 0; t++) {
      V3 ax = bax[t < 2 ? c1 : c2] * ((t & 1) ? -1.0 : 1.0);     // keep ax . (p - c) <= h
      n = clip_halfspace(poly, n, B.c + ax * B.h[t < 2 ? c1 : c2], -ax);
    }
  }
  V3 kept[16]; double dist[16]; int m = 0;
  for (int i = 0; i < n; i++) {
    double dv = dot(rn, poly[i]) - rd;              // signed distance above the reference face
    note_decision(margin - dv);
    if (dv <= margin) { kept[m] = poly[i]; dist[m] = dv; m++; }
  }
  if (m == 0) return 0;
  int sel[4]; int cnt = keep_four(kept, dist, m, sel);
  for (int j = 0; j < cnt; j++) {
    V3 pt = kept[sel[j]]; double dv = dist[sel[j]];
    out[j].n = nBA; out[j].dist = dv;
    if (kind == 0) { out[j].posB = pt; out[j].posA = pt + nBA * dv; }     // the point is on B (incident), A's face is the reference
    else { out[j].posA = pt; out[j].posB = pt - nBA * dv; }
  }
  return cnt;
}

// Closest point of the polytope to a sphere centre: returns like sphere_box (point on the polytope, normal towards the sphere).
inline bool sphere_poly(V3 sc, double r, const PolyW& A, double margin, V3& on_poly, V3& n_to_sphere, double& dist) {
  double mx = -1e300; int fm = 0;
  for (int f = 0; f < A.nf; f++) { double s = dot(A.fn[f], sc) - A.fd[f]; if (s > mx) { mx = s; fm = f; } }
  if (mx <= 0) {  // centre inside: out through the nearest face
    n_to_sphere = A.fn[fm]; on_poly = sc - A.fn[fm] * mx; dist = mx - r;
  } else {
    double bd = 1e300; V3 bp;
    for (int f = 0; f < A.nf; f++) {
      double s = dot(A.fn[f], sc) - A.fd[f];
      if (s <= 0) continue;
      V3 q = sc - A.fn[f] * s;                       // projection on the face plane
      const int cnt = A.src->fcount[f]; const int* loop = A.src->floop[f];
      bool inside = true; double be = 1e300; V3 bq;
      for (int k = 0; k < cnt; k++) {
        V3 o = A.v[loop[k]], e = A.v[loop[(k + 1) % cnt]] - o;
        if (dot(cross(A.fn[f], e), q - o) < 0) inside = false;
        double t = dot(sc - o, e) / dot(e, e); t = std::fmin(1.0, std::fmax(0.0, t));
        V3 c = o + e * t; double d2 = dot(sc - c, sc - c);
        if (d2 < be) { be = d2; bq = c; }
      }
      V3 cp = inside ? q : bq; double d2 = inside ? s * s : be;
      if (d2 < bd) { bd = d2; bp = cp; }
    }
    double len = std::sqrt(bd);
    on_poly = bp; n_to_sphere = (sc - bp) * (1.0 / len); dist = len - r;
  }
  note_decision(dist - margin);
  return dist < margin;
}

}  // namespace urgo

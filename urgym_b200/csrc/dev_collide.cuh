// fp32 narrowphase for the primitive pairs of the five envs: box-box (15-axis SAT, incident-face clipping, <= 4
// manifold points), sphere-box, sphere-sphere. Contacts are reported while the separation is below the pair's
// contact-breaking threshold, as pybullet's getContactPoints does. Stateless.
#pragma once
#include "dev_math.cuh"
#include "model.h"

struct BoxW { f3 c; m3 R; f3 h; };
struct CPt { f3 posA, posB, n; real dist; };  // n from B to A

UG_DEV real edge_bias(real s) { return s < 0._r ? s * 1.05_r : s / 1.05_r; }

// Sutherland-Hodgman clip of a convex polygon (x,y,z per vertex) against |x|<=hx, |y|<=hy. The callers clip a quad: each of
// the four half-planes adds at most one vertex, so the buffers hold 8 vertices (local memory: keep it small).
__device__ __noinline__ int clip_rect(real (*bufA)[3], int n, real hx, real hy, real (*bufB)[3], int* which) {
  real (*src)[3] = bufA; real (*dst)[3] = bufB;
  int flip = 0;
  for (int e = 0; e < 4; e++) {
    int ax = e >> 1; real sgn = (e & 1) ? -1.0_r : 1.0_r; real lim = ax == 0 ? hx : hy;
    int m = 0;
    for (int i = 0; i < n; i++) {
      const real* P = src[i]; const real* Q = src[(i + 1 == n) ? 0 : i + 1];
      real dp = lim - sgn * P[ax], dq = lim - sgn * Q[ax];
      if (dp >= 0._r) { dst[m][0] = P[0]; dst[m][1] = P[1]; dst[m][2] = P[2]; m++; }
      if ((dp >= 0._r) != (dq >= 0._r)) {
        real t = dp / (dp - dq);
        dst[m][0] = P[0] + t * (Q[0] - P[0]); dst[m][1] = P[1] + t * (Q[1] - P[1]); dst[m][2] = P[2] + t * (Q[2] - P[2]);
        m++;
      }
    }
    n = m;
    real (*tmp)[3] = src; src = dst; dst = tmp; flip ^= 1;
    if (n == 0) break;
  }
  *which = flip;  // 0: result in bufA, 1: result in bufB
  return n;
}

__device__ __noinline__ int box_box(const BoxW& A, const BoxW& B, real margin, CPt* out) {
  f3 p = B.c - A.c;
  f3 pp = tmulv(A.R, p);
  real R[3][3], Q[3][3];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) { R[i][j] = dot(col(A.R, i), col(B.R, j)); Q[i][j] = rabs(R[i][j]); }
  real best = -1e30_r; int code = 0; bool invert = false; f3 nrm = mk3(0, 0, 1);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    real e1 = comp(pp, i);
    real e2 = comp(A.h, i) + B.h.x * Q[i][0] + B.h.y * Q[i][1] + B.h.z * Q[i][2];
    real s = rabs(e1) - e2;
    if (s > margin) return 0;
    if (s > best) { best = s; nrm = col(A.R, i); invert = e1 < 0._r; code = 1 + i; }
  }
#pragma unroll
  for (int j = 0; j < 3; j++) {
    real e1 = dot(col(B.R, j), p);
    real e2 = A.h.x * Q[0][j] + A.h.y * Q[1][j] + A.h.z * Q[2][j] + comp(B.h, j);
    real s = rabs(e1) - e2;
    if (s > margin) return 0;
    if (s > best) { best = s; nrm = col(B.R, j); invert = e1 < 0._r; code = 4 + j; }
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
#pragma unroll
    for (int j = 0; j < 3; j++) {
      const int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      f3 L = mk3(0, 0, 0);
      setc(L, i1, -R[i2][j]); setc(L, i2, R[i1][j]);
      real l = norm3(L);
      if (l <= 1e-9_r) continue;
      real e1 = dot(pp, L);
      real e2 = comp(A.h, i1) * Q[i2][j] + comp(A.h, i2) * Q[i1][j] + comp(B.h, j1) * Q[i][j2] + comp(B.h, j2) * Q[i][j1];
      real s = (rabs(e1) - e2) / l;
      if (s > margin) return 0;
      if (edge_bias(s) > best) { best = s; nrm = mulv(A.R, L * (1.0_r / l)); invert = e1 < 0._r; code = 7 + 3 * i + j; }
    }
  }
  if (invert) nrm = -nrm;
  real depth = -best;
  if (code >= 7) {
    int i = (code - 7) / 3, j = (code - 7) % 3;
    f3 pa = A.c, pb = B.c;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      pa = pa + col(A.R, k) * ((dot(nrm, col(A.R, k)) > 0._r ? 1.0_r : -1.0_r) * comp(A.h, k));
      pb = pb - col(B.R, k) * ((dot(nrm, col(B.R, k)) > 0._r ? 1.0_r : -1.0_r) * comp(B.h, k));
    }
    f3 ua = col(A.R, i), ub = col(B.R, j), d = pb - pa;
    real uaub = dot(ua, ub), q1 = dot(ua, d), q2 = -dot(ub, d), den = 1.0_r - uaub * uaub, beta = 0._r;
    if (den > 1e-4_r) beta = (uaub * q1 + q2) / den;
    pb = pb + ub * beta;
    out[0].n = -nrm; out[0].dist = -depth; out[0].posB = pb; out[0].posA = pb + out[0].n * out[0].dist;
    return 1;
  }
  const BoxW& ref = code <= 3 ? A : B;
  const BoxW& inc = code <= 3 ? B : A;
  f3 nr = code <= 3 ? nrm : -nrm;
  int k = (code - 1) % 3, c1 = (k + 1) % 3, c2 = (k + 2) % 3;
  int la = 0; real lav = -1._r;
#pragma unroll
  for (int j = 0; j < 3; j++) { real v = rabs(dot(nr, col(inc.R, j))); if (v > lav) { lav = v; la = j; } }
  int a1 = (la + 1) % 3, a2 = (la + 2) % 3;
  real sg = dot(nr, col(inc.R, la)) > 0._r ? -1.0_r : 1.0_r;
  f3 fc = inc.c + col(inc.R, la) * (sg * comp(inc.h, la)) - ref.c;
  f3 e1v = col(inc.R, a1) * comp(inc.h, a1), e2v = col(inc.R, a2) * comp(inc.h, a2);
  f3 rc1 = col(ref.R, c1), rc2 = col(ref.R, c2);
  real bufA[8][3], bufB[8][3];
  {
    f3 quad[4] = {fc - e1v - e2v, fc + e1v - e2v, fc + e1v + e2v, fc - e1v + e2v};
#pragma unroll
    for (int v = 0; v < 4; v++) { bufA[v][0] = dot(quad[v], rc1); bufA[v][1] = dot(quad[v], rc2); bufA[v][2] = dot(quad[v], nr); }
  }
  int which;
  int n = clip_rect(bufA, 4, comp(ref.h, c1), comp(ref.h, c2), bufB, &which);
  real (*cl)[3] = which ? bufB : bufA;
  int keep[8]; real dep[8]; int m = 0;
  real hk = comp(ref.h, k);
  for (int v = 0; v < n; v++) { real dv = hk - cl[v][2]; if (dv >= -margin) { keep[m] = v; dep[m] = dv; m++; } }
  if (m == 0) return 0;
  int cnt = m < 4 ? m : 4, i0 = 0;
  if (m > 4) for (int v = 1; v < m; v++) if (dep[v] > dep[i0]) i0 = v;
  for (int j = 0; j < cnt; j++) {
    int s = m > 4 ? (i0 + (j * m) / 4) % m : j;
    const real* c = cl[keep[s]];
    f3 pt = ref.c + rc1 * c[0] + rc2 * c[1] + nr * c[2];
    real dv = dep[s];
    out[j].n = -nrm; out[j].dist = -dv;
    if (code <= 3) { out[j].posB = pt; out[j].posA = pt + out[j].n * out[j].dist; }
    else { out[j].posA = pt; out[j].posB = pt - out[j].n * out[j].dist; }
  }
  return cnt;
}

UG_DEV bool sphere_box(f3 sc, real r, const BoxW& B, real margin, f3& on_box, f3& n_bs, real& dist) {
  f3 l = tmulv(B.R, sc - B.c);
  f3 cl = mk3(rmin(rmax(l.x, -B.h.x), B.h.x), rmin(rmax(l.y, -B.h.y), B.h.y), rmin(rmax(l.z, -B.h.z), B.h.z));
  f3 d = l - cl;
  real len2 = dot(d, d);
  f3 nl;
  if (len2 > 1e-24_r) { real len = rsqrt_(len2); nl = d * (1.0_r / len); dist = len - r; }
  else {
    int ax = 0; real md = B.h.x - rabs(l.x);
    real v1 = B.h.y - rabs(l.y); if (v1 < md) { md = v1; ax = 1; }
    real v2 = B.h.z - rabs(l.z); if (v2 < md) { md = v2; ax = 2; }
    nl = mk3(0, 0, 0); real sgn = comp(l, ax) >= 0._r ? 1.0_r : -1.0_r; setc(nl, ax, sgn);
    cl = l; setc(cl, ax, sgn * comp(B.h, ax));
    dist = -md - r;
  }
  if (!(dist < margin)) return false;
  on_box = B.c + mulv(B.R, cl);
  n_bs = mulv(B.R, nl);
  return true;
}

// ---------------------------------------------------------------------------------------------------------------
// Convex polytope (gripper link with mesh geometry; Bullet collides the convex hull of the mesh vertices) against a box
// or a sphere: separating axes over the polytope's face normals, the box's face normals and the cross products of their
// edge directions; then the incident face is clipped against the side planes of the reference face (polyhedral contact
// clipping, the box-box scheme above with face lists). Same acceptance rule as box_box: points within `margin` of the
// reference face, at most four, spread from the deepest. o / R: world frame of the link that carries the polytope.
UG_DEV int clip_halfspace(f3* poly, int n, f3 o, f3 m) {
  f3 out[24]; int k = 0;
  for (int i = 0; i < n; i++) {
    f3 P = poly[i], Q = poly[(i + 1 == n) ? 0 : i + 1];
    real dp = dot(m, P - o), dq = dot(m, Q - o);
    if (dp >= 0._r) out[k++] = P;
    if ((dp >= 0._r) != (dq >= 0._r)) { real t = dp / (dp - dq); out[k++] = P + (Q - P) * t; }
  }
  if (k > 16) k = 16;
  for (int i = 0; i < k; i++) poly[i] = out[i];
  return k;
}

__device__ __noinline__ int poly_box(const DevPoly& P, f3 o, const m3& R, const BoxW& B, real margin, CPt* out) {
  f3 v[UG_POLY_MAXV];
  const int nv = P.nv, nf = P.nf, ne = P.ne;
#pragma unroll 1
  for (int i = 0; i < nv; i++) v[i] = o + mulv(R, ld3(P.v[i]));
  const f3 bax[3] = {col(B.R, 0), col(B.R, 1), col(B.R, 2)};
  real best = -1e30_r; int kind = 0, idx = 0, sgn = 1; f3 nBA = mk3(0, 0, 1);
#pragma unroll 1
  for (int f = 0; f < nf; f++) {                   // faces of the polytope: lowest point of the box above the face plane
    const f3 n = mulv(R, ld3(P.fn[f]));
    const real fd = P.fn[f][3] + dot(n, o);
    const real sB = dot(n, B.c) - (B.h.x * rabs(dot(n, bax[0])) + B.h.y * rabs(dot(n, bax[1])) + B.h.z * rabs(dot(n, bax[2])));
    const real s = sB - fd;
    if (s > margin) return 0;
    if (s > best) { best = s; kind = 0; idx = f; nBA = -n; }
  }
#pragma unroll 1
  for (int j = 0; j < 3; j++) {                    // faces of the box: extent of the polytope along the box axis
    real lo = 1e30_r, hi = -1e30_r;
#pragma unroll 1
    for (int i = 0; i < nv; i++) { real p = dot(bax[j], v[i] - B.c); lo = rmin(lo, p); hi = rmax(hi, p); }
    const real hj = comp(B.h, j), sp = lo - hj, sm = -hi - hj;
    if (sp > margin || sm > margin) return 0;
    if (sp > best) { best = sp; kind = 1; idx = j; sgn = 1; nBA = bax[j]; }
    if (sm > best) { best = sm; kind = 1; idx = j; sgn = -1; nBA = -bax[j]; }
  }
  int ea = 0, eb = 0;
#pragma unroll 1
  for (int e = 0; e < ne; e++) {
    const f3 ed = mulv(R, ld3(P.ed[e]));
#pragma unroll 1
    for (int j = 0; j < 3; j++) {                  // edge x edge
      f3 L = cross(ed, bax[j]); const real l = norm3(L);
      if (l <= 1e-9_r) continue;
      L = L * (1.0_r / l);
      real mnA = 1e30_r, mxA = -1e30_r;
#pragma unroll 1
      for (int i = 0; i < nv; i++) { real p = dot(L, v[i] - B.c); mnA = rmin(mnA, p); mxA = rmax(mxA, p); }
      const real rB = B.h.x * rabs(dot(L, bax[0])) + B.h.y * rabs(dot(L, bax[1])) + B.h.z * rabs(dot(L, bax[2]));
      const real s1 = mnA - rB, s2 = -mxA - rB, sx = rmax(s1, s2);
      if (sx > margin) return 0;
      if (edge_bias(sx) > best) { best = sx; kind = 2; ea = e; eb = j; nBA = s1 >= s2 ? L : -L; }
    }
  }
  if (kind == 2) {                                 // one point: closest points of the two supporting edges
    real exA = 1e30_r;
    for (int i = 0; i < nv; i++) exA = rmin(exA, dot(nBA, v[i]));
    f3 pa = mk3(0, 0, 0); int ca = 0;
    for (int i = 0; i < nv; i++) if (dot(nBA, v[i]) <= exA + 1e-5_r) { pa = pa + v[i]; ca++; }
    pa = pa * (1.0_r / (real)ca);
    f3 pb = B.c;
#pragma unroll
    for (int k = 0; k < 3; k++) if (k != eb) pb = pb + bax[k] * ((dot(nBA, bax[k]) > 0._r ? 1.0_r : -1.0_r) * comp(B.h, k));
    const f3 ua = mulv(R, ld3(P.ed[ea])), ub = bax[eb], d = pb - pa;
    const real uaub = dot(ua, ub), q1 = dot(ua, d), q2 = -dot(ub, d), den = 1.0_r - uaub * uaub;
    real beta = 0._r;
    if (den > 1e-4_r) beta = (uaub * q1 + q2) / den;
    pb = pb + ub * beta;
    out[0].n = nBA; out[0].dist = best; out[0].posB = pb; out[0].posA = pb + nBA * best;
    return 1;
  }
  f3 poly[24]; int n = 0; f3 rn; real rd;
  if (kind == 0) {                                 // reference face on the polytope; incident = the box face most against it
    rn = mulv(R, ld3(P.fn[idx])); rd = P.fn[idx][3] + dot(rn, o);
    int la = 0; real lav = -1._r;
#pragma unroll
    for (int j = 0; j < 3; j++) { real t = rabs(dot(rn, bax[j])); if (t > lav) { lav = t; la = j; } }
    const int a1 = (la + 1) % 3, a2 = (la + 2) % 3;
    const real sg = dot(rn, bax[la]) > 0._r ? -1.0_r : 1.0_r;
    const f3 fc = B.c + bax[la] * (sg * comp(B.h, la)), e1 = bax[a1] * comp(B.h, a1), e2 = bax[a2] * comp(B.h, a2);
    poly[0] = fc - e1 - e2; poly[1] = fc + e1 - e2; poly[2] = fc + e1 + e2; poly[3] = fc - e1 + e2; n = 4;
    const int cnt = P.fcount[idx];
    for (int k = 0; k < cnt && n > 0; k++) {
      const f3 a = v[P.floop[idx][k]], e = v[P.floop[idx][(k + 1 == cnt) ? 0 : k + 1]] - a;
      n = clip_halfspace(poly, n, a, cross(rn, e));
    }
  } else {                                         // reference face on the box; incident = the polytope's face most against it
    rn = bax[idx] * (real)sgn; rd = dot(rn, B.c) + comp(B.h, idx);
    int fi = 0; real fv = 1e30_r;
#pragma unroll 1
    for (int f = 0; f < nf; f++) { real t = dot(mulv(R, ld3(P.fn[f])), rn); if (t < fv) { fv = t; fi = f; } }
    const int cnt = P.fcount[fi];
    for (int k = 0; k < cnt; k++) poly[n++] = v[P.floop[fi][k]];
    const int c1 = (idx + 1) % 3, c2 = (idx + 2) % 3;
    for (int t = 0; t < 4 && n > 0; t++) {
      const int c = t < 2 ? c1 : c2;
      const f3 ax = bax[c] * ((t & 1) ? -1.0_r : 1.0_r);
      n = clip_halfspace(poly, n, B.c + ax * comp(B.h, c), -ax);
    }
  }
  real dist[16]; int m = 0;
  for (int i = 0; i < n; i++) {
    const real dv = dot(rn, poly[i]) - rd;
    if (dv <= margin) { poly[m] = poly[i]; dist[m] = dv; m++; }
  }
  if (m == 0) return 0;
  int cnt = m < 4 ? m : 4, i0 = 0;
  if (m > 4) for (int t = 1; t < m; t++) if (-dist[t] > -dist[i0]) i0 = t;
  for (int j = 0; j < cnt; j++) {
    const int sl = m > 4 ? (i0 + (j * m) / 4) % m : j;
    const f3 pt = poly[sl]; const real dv = dist[sl];
    out[j].n = nBA; out[j].dist = dv;
    if (kind == 0) { out[j].posB = pt; out[j].posA = pt + nBA * dv; }
    else { out[j].posA = pt; out[j].posB = pt - nBA * dv; }
  }
  return cnt;
}

// ---------------------------------------------------------------------------------------------------------------
// poly_box by four lanes. A tick of the ball / golf scenes runs 4..10 polytope pairs (finger links against the paddle or
// the club), each 4.5 k instructions of separating-axis tests plus the clipping when one lane does it, with its vertex and
// clip buffers in local memory: measured 190 k cycles per tick with one warp per SM and 600 k with thirteen (the buffers
// fall out of the small L1 next to 225 KB of shared memory), i.e. more than the PGS sweeps (profiles/r02_latency_chains.txt).
// Here the four lanes [4g, 4g + 4) of the warp take one pair: each computes every fourth axis separation (the same
// expressions as poly_box, operand for operand) into a shared-memory table, then all four scan the table in poly_box's
// order with poly_box's comparisons, so the selected axis is the one poly_box selects; lane 4g clips in shared memory.
// Buffers: `sc` = UG_PB_SCRATCH reals of shared memory per group: [0,48) world vertices, later the result (count is the
// return value; 4 x {posA, posB, n, dist}); [48,192) two clip buffers of 24 vertices, during the axis tests the table of
// separations [48,122) and of edge-direction signs [122,158); [192,208) the kept distances.
#define UG_PB_SCRATCH 208
#define UG_PB_GROUPS 8
UG_DEV int clip_halfspace_sm(const real* src, int n, f3 o, f3 m, real* dst) {
  int k = 0;
  for (int i = 0; i < n; i++) {
    const f3 P = ld3(src + 3 * i), Q = ld3(src + 3 * ((i + 1 == n) ? 0 : i + 1));
    const real dp = dot(m, P - o), dq = dot(m, Q - o);
    if (dp >= 0._r) { st3(dst + 3 * k, P); k++; }
    if ((dp >= 0._r) != (dq >= 0._r)) { const real t = dp / (dp - dq); st3(dst + 3 * k, P + (Q - P) * t); k++; }
  }
  return k > 16 ? 16 : k;
}

__device__ __noinline__ int poly_box_coop(const DevPoly& P, f3 o, const m3& R, const BoxW& B, real margin, real* sc, int sub, unsigned gmask) {
  real* const vs = sc; real* const sep = sc + 48; real* const sgnf = sc + 122; real* const dist = sc + 192;
  const int nv = P.nv, nf = P.nf, ne = P.ne;
#pragma unroll 1
  for (int i = sub; i < nv; i += 4) st3(vs + 3 * i, o + mulv(R, ld3(P.v[i])));
  const f3 bax[3] = {col(B.R, 0), col(B.R, 1), col(B.R, 2)};
  __syncwarp(gmask);
  // ---- the table of separations: [0,nf) polytope faces, [nf,nf+6) box faces (+, - per axis), [nf+6, nf+6+3ne) edge pairs
#pragma unroll 1
  for (int f = sub; f < nf; f += 4) {
    const f3 n = mulv(R, ld3(P.fn[f]));
    const real fd = P.fn[f][3] + dot(n, o);
    const real sB = dot(n, B.c) - (B.h.x * rabs(dot(n, bax[0])) + B.h.y * rabs(dot(n, bax[1])) + B.h.z * rabs(dot(n, bax[2])));
    sep[f] = sB - fd;
  }
  if (sub < 3) {
    const int j = sub;
    const f3 bj = j == 0 ? bax[0] : (j == 1 ? bax[1] : bax[2]);
    real lo = 1e30_r, hi = -1e30_r;
#pragma unroll 1
    for (int i = 0; i < nv; i++) { real p = dot(bj, ld3(vs + 3 * i) - B.c); lo = rmin(lo, p); hi = rmax(hi, p); }
    const real hj = comp(B.h, j);
    sep[nf + 2 * j] = lo - hj; sep[nf + 2 * j + 1] = -hi - hj;
  }
#pragma unroll 1
  for (int t = sub; t < 3 * ne; t += 4) {
    const int e = t / 3, j = t - 3 * e;
    const f3 ed = mulv(R, ld3(P.ed[e]));
    const f3 bj = j == 0 ? bax[0] : (j == 1 ? bax[1] : bax[2]);
    f3 L = cross(ed, bj); const real l = norm3(L);
    real sx = -1e30_r, fl = 0._r;            // degenerate pair of directions: never separating, never selected
    if (!(l <= 1e-9_r)) {
      L = L * (1.0_r / l);
      real mnA = 1e30_r, mxA = -1e30_r;
#pragma unroll 1
      for (int i = 0; i < nv; i++) { real p = dot(L, ld3(vs + 3 * i) - B.c); mnA = rmin(mnA, p); mxA = rmax(mxA, p); }
      const real rB = B.h.x * rabs(dot(L, bax[0])) + B.h.y * rabs(dot(L, bax[1])) + B.h.z * rabs(dot(L, bax[2]));
      const real s1 = mnA - rB, s2 = -mxA - rB;
      sx = rmax(s1, s2); fl = s1 >= s2 ? 1._r : 0._r;
    }
    sep[nf + 6 + t] = sx; sgnf[t] = fl;
  }
  __syncwarp(gmask);
  // ---- every lane scans the table in poly_box's order
  real best = -1e30_r; int kind = 0, idx = 0, sgn = 1, tsel = 0;
#pragma unroll 1
  for (int f = 0; f < nf; f++) { const real sv = sep[f]; if (sv > margin) return 0; if (sv > best) { best = sv; kind = 0; idx = f; } }
#pragma unroll 1
  for (int c = 0; c < 6; c++) { const real sv = sep[nf + c]; if (sv > margin) return 0; if (sv > best) { best = sv; kind = 1; idx = c >> 1; sgn = (c & 1) ? -1 : 1; } }
#pragma unroll 1
  for (int t = 0; t < 3 * ne; t++) { const real sv = sep[nf + 6 + t]; if (sv > margin) return 0; if (edge_bias(sv) > best) { best = sv; kind = 2; tsel = t; } }
  const real sel_flag = sgnf[tsel];
  // kind 1 needs the polytope's face most against the reference face: same split, table in sep[0,nf)
  f3 rn = mk3(0, 0, 1);
  if (kind == 1) rn = (idx == 0 ? bax[0] : (idx == 1 ? bax[1] : bax[2])) * (real)sgn;
  __syncwarp(gmask);                         // everyone has read the table
  if (kind == 1) {
#pragma unroll 1
    for (int f = sub; f < nf; f += 4) sep[f] = dot(mulv(R, ld3(P.fn[f])), rn);
  }
  __syncwarp(gmask);
  int cnt = 0;
  if (sub == 0) {
    real* const res = vs;                    // the vertices are dead once the clip has read them (kind 0 reads them; see below)
    if (kind == 2) {                         // one point: closest points of the two supporting edges
      const int ea = tsel / 3, eb = tsel - 3 * ea;
      f3 L = cross(mulv(R, ld3(P.ed[ea])), eb == 0 ? bax[0] : (eb == 1 ? bax[1] : bax[2]));
      L = L * (1.0_r / norm3(L));
      const f3 nBA = sel_flag != 0._r ? L : -L;
      real exA = 1e30_r;
      for (int i = 0; i < nv; i++) exA = rmin(exA, dot(nBA, ld3(vs + 3 * i)));
      f3 pa = mk3(0, 0, 0); int ca = 0;
      for (int i = 0; i < nv; i++) if (dot(nBA, ld3(vs + 3 * i)) <= exA + 1e-5_r) { pa = pa + ld3(vs + 3 * i); ca++; }
      pa = pa * (1.0_r / (real)ca);
      f3 pb = B.c;
#pragma unroll
      for (int k = 0; k < 3; k++) if (k != eb) pb = pb + bax[k] * ((dot(nBA, bax[k]) > 0._r ? 1.0_r : -1.0_r) * comp(B.h, k));
      const f3 ua = mulv(R, ld3(P.ed[ea])), ub = eb == 0 ? bax[0] : (eb == 1 ? bax[1] : bax[2]), d = pb - pa;
      const real uaub = dot(ua, ub), q1 = dot(ua, d), q2 = -dot(ub, d), den = 1.0_r - uaub * uaub;
      real beta = 0._r;
      if (den > 1e-4_r) beta = (uaub * q1 + q2) / den;
      pb = pb + ub * beta;
      st3(res + 0, pb + nBA * best); st3(res + 3, pb); st3(res + 6, nBA); res[9] = best;
      cnt = 1;
    } else {
      real* bufA = sc + 48; real* bufB = sc + 120;
      int n = 0; real rd; f3 nBA;
      if (kind == 0) {                       // reference face on the polytope; incident = the box face most against it
        rn = mulv(R, ld3(P.fn[idx])); rd = P.fn[idx][3] + dot(rn, o); nBA = -rn;
        int la = 0; real lav = -1._r;
#pragma unroll
        for (int j = 0; j < 3; j++) { real t = rabs(dot(rn, bax[j])); if (t > lav) { lav = t; la = j; } }
        const int a1 = (la + 1) % 3, a2 = (la + 2) % 3;
        const f3 bl = la == 0 ? bax[0] : (la == 1 ? bax[1] : bax[2]), b1 = a1 == 0 ? bax[0] : (a1 == 1 ? bax[1] : bax[2]), b2 = a2 == 0 ? bax[0] : (a2 == 1 ? bax[1] : bax[2]);
        const real sg = dot(rn, bl) > 0._r ? -1.0_r : 1.0_r;
        const f3 fc = B.c + bl * (sg * comp(B.h, la)), e1 = b1 * comp(B.h, a1), e2 = b2 * comp(B.h, a2);
        st3(bufA + 0, fc - e1 - e2); st3(bufA + 3, fc + e1 - e2); st3(bufA + 6, fc + e1 + e2); st3(bufA + 9, fc - e1 + e2); n = 4;
        const int fcnt = P.fcount[idx];
        for (int k = 0; k < fcnt && n > 0; k++) {
          const f3 a = ld3(vs + 3 * P.floop[idx][k]), e = ld3(vs + 3 * P.floop[idx][(k + 1 == fcnt) ? 0 : k + 1]) - a;
          n = clip_halfspace_sm(bufA, n, a, cross(rn, e), bufB);
          real* t = bufA; bufA = bufB; bufB = t;
        }
      } else {                               // reference face on the box; incident = the polytope's face most against it
        nBA = rn; rd = dot(rn, B.c) + comp(B.h, idx);
        int fi = 0; real fv = 1e30_r;
#pragma unroll 1
        for (int f = 0; f < nf; f++) { const real t = sep[f]; if (t < fv) { fv = t; fi = f; } }
        const int fcnt = P.fcount[fi];
        real* inc = sc + 120;                // the table still occupies [48, 48 + nf): load the face into the other buffer
        for (int k = 0; k < fcnt; k++) st3(inc + 3 * k, ld3(vs + 3 * P.floop[fi][k]));
        n = fcnt; bufA = sc + 120; bufB = sc + 48;
        const int c1 = (idx + 1) % 3, c2 = (idx + 2) % 3;
        for (int t = 0; t < 4 && n > 0; t++) {
          const int c = t < 2 ? c1 : c2;
          const f3 ax = (c == 0 ? bax[0] : (c == 1 ? bax[1] : bax[2])) * ((t & 1) ? -1.0_r : 1.0_r);
          n = clip_halfspace_sm(bufA, n, B.c + ax * comp(B.h, c), -ax, bufB);
          real* tt = bufA; bufA = bufB; bufB = tt;
        }
      }
      int m = 0;
      for (int i = 0; i < n; i++) {
        const f3 pt = ld3(bufA + 3 * i);
        const real dv = dot(rn, pt) - rd;
        if (dv <= margin) { st3(bufA + 3 * m, pt); dist[m] = dv; m++; }
      }
      if (m > 0) {
        cnt = m < 4 ? m : 4; int i0 = 0;
        if (m > 4) for (int t = 1; t < m; t++) if (-dist[t] > -dist[i0]) i0 = t;
        for (int j = 0; j < cnt; j++) {
          const int sl = m > 4 ? (i0 + (j * m) / 4) % m : j;
          const f3 pt = ld3(bufA + 3 * sl); const real dv = dist[sl];
          real* r = res + 10 * j;
          if (kind == 0) { st3(r, pt + nBA * dv); st3(r + 3, pt); } else { st3(r, pt); st3(r + 3, pt - nBA * dv); }
          st3(r + 6, nBA); r[9] = dv;
        }
      }
    }
  }
  return cnt;
}

// Closest point of the polytope to a sphere centre, like sphere_box: point on the polytope, normal towards the sphere.
__device__ __noinline__ bool sphere_poly(f3 sc, real r, const DevPoly& P, f3 o, const m3& R, real margin, f3& on_poly, f3& n_ps, real& dist) {
  const int nf = P.nf;
  real mx = -1e30_r; int fm = 0;
#pragma unroll 1
  for (int f = 0; f < nf; f++) { const f3 n = mulv(R, ld3(P.fn[f])); real s = dot(n, sc) - (P.fn[f][3] + dot(n, o)); if (s > mx) { mx = s; fm = f; } }
  if (mx <= 0._r) {
    const f3 n = mulv(R, ld3(P.fn[fm]));
    n_ps = n; on_poly = sc - n * mx; dist = mx - r;
  } else {
    real bd = 1e30_r; f3 bp = sc;
#pragma unroll 1
    for (int f = 0; f < nf; f++) {
      const f3 n = mulv(R, ld3(P.fn[f]));
      const real s = dot(n, sc) - (P.fn[f][3] + dot(n, o));
      if (s <= 0._r) continue;
      const f3 q = sc - n * s;
      const int cnt = P.fcount[f];
      bool inside = true; real be = 1e30_r; f3 bq = q;
      for (int k = 0; k < cnt; k++) {
        const f3 a = o + mulv(R, ld3(P.v[P.floop[f][k]])), e = o + mulv(R, ld3(P.v[P.floop[f][(k + 1 == cnt) ? 0 : k + 1]])) - a;
        if (dot(cross(n, e), q - a) < 0._r) inside = false;
        real t = dot(sc - a, e) / dot(e, e); t = rmin(1.0_r, rmax(0._r, t));
        const f3 c = a + e * t; const real d2 = dot(sc - c, sc - c);
        if (d2 < be) { be = d2; bq = c; }
      }
      const f3 cp = inside ? q : bq; const real d2 = inside ? s * s : be;
      if (d2 < bd) { bd = d2; bp = cp; }
    }
    const real len = rsqrt_(bd);
    on_poly = bp; n_ps = (sc - bp) * (1.0_r / len); dist = len - r;
  }
  return dist < margin;
}

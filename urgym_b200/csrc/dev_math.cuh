// Device helpers in the build's scalar type `real`: 3-vectors, row-major 3x3, quaternions (x,y,z,w), Euler conventions
// of the reference's pybullet calls (getEulerFromQuaternion / getQuaternionFromEuler, env_ball_balance_v0.py:114-116,139).
//
// `real` is float in the product build (liburgym_b200.so) and double in the parity build (liburgym_b200_f64.so,
// -DUG_F64): the same sources, literals (`_r`), intrinsics (r* wrappers) and shared-memory layout in the other width.
// The parity build exists so that integer outputs (settle-loop trip counts, flags, contact-pair lists) can be compared
// with the fp64 CPU oracle without fp32 threshold flips; it drops the approximate intrinsics.
#pragma once
#include <cuda_runtime.h>

#include "real.h"

#define UG_DEV __device__ __forceinline__

__host__ __device__ constexpr real operator""_r(long double v) { return (real)v; }

#ifdef UG_F64
typedef double4 real4;
UG_DEV real4 make_real4(real x, real y, real z, real w) { return make_double4(x, y, z, w); }
UG_DEV real rmin(real a, real b) { return fmin(a, b); }
UG_DEV real rmax(real a, real b) { return fmax(a, b); }
UG_DEV real rabs(real a) { return fabs(a); }
UG_DEV real rsqrt_(real a) { return sqrt(a); }
UG_DEV real rrsqrt(real a) { return 1.0 / sqrt(a); }
UG_DEV real rsin(real a) { return sin(a); }
UG_DEV real rcos(real a) { return cos(a); }
UG_DEV void rsincos(real a, real* s, real* c) { sincos(a, s, c); }
UG_DEV void rsincos_fast(real a, real* s, real* c) { sincos(a, s, c); }
UG_DEV real ratan2(real a, real b) { return atan2(a, b); }
UG_DEV real rasin(real a) { return asin(a); }
UG_DEV real rtanh(real a) { return tanh(a); }
UG_DEV real rfma(real a, real b, real c) { return fma(a, b, c); }
UG_DEV real rsqrt_ftz(real x) { return 1.0 / sqrt(x); }
#else
typedef float4 real4;
UG_DEV real4 make_real4(real x, real y, real z, real w) { return make_float4(x, y, z, w); }
UG_DEV real rmin(real a, real b) { return fminf(a, b); }
UG_DEV real rmax(real a, real b) { return fmaxf(a, b); }
UG_DEV real rabs(real a) { return fabsf(a); }
UG_DEV real rsqrt_(real a) { return sqrtf(a); }
UG_DEV real rrsqrt(real a) { return rsqrtf(a); }
UG_DEV real rsin(real a) { return sinf(a); }
UG_DEV real rcos(real a) { return cosf(a); }
UG_DEV void rsincos(real a, real* s, real* c) { sincosf(a, s, c); }
UG_DEV void rsincos_fast(real a, real* s, real* c) { __sincosf(a, s, c); }
UG_DEV real ratan2(real a, real b) { return atan2f(a, b); }
UG_DEV real rasin(real a) { return asinf(a); }
UG_DEV real rtanh(real a) { return tanhf(a); }
UG_DEV real rfma(real a, real b, real c) { return fmaf(a, b, c); }
UG_DEV real rsqrt_ftz(real x) { real r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
#endif

struct f3 { real x, y, z; };
UG_DEV f3 mk3(real x, real y, real z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
UG_DEV f3 ld3(const real* p) { return mk3(p[0], p[1], p[2]); }
UG_DEV void st3(real* p, f3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
UG_DEV f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
UG_DEV f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
UG_DEV f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
UG_DEV f3 operator*(f3 a, real s) { return mk3(a.x * s, a.y * s, a.z * s); }
UG_DEV f3 operator*(real s, f3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
UG_DEV real dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
UG_DEV f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
UG_DEV real norm3(f3 a) { return rsqrt_(dot(a, a)); }
UG_DEV real comp(f3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
UG_DEV void setc(f3& a, int i, real v) { if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v; }

struct m3 { real m[9]; };
UG_DEV m3 ldm(const real* p) { m3 r;
#pragma unroll
  for (int i = 0; i < 9; i++) r.m[i] = p[i]; return r; }
UG_DEV void stm(real* p, const m3& a) {
#pragma unroll
  for (int i = 0; i < 9; i++) p[i] = a.m[i]; }
UG_DEV f3 col(const m3& a, int c) { return mk3(a.m[c], a.m[3 + c], a.m[6 + c]); }
UG_DEV f3 mulv(const m3& a, f3 v) { return mk3(a.m[0] * v.x + a.m[1] * v.y + a.m[2] * v.z, a.m[3] * v.x + a.m[4] * v.y + a.m[5] * v.z, a.m[6] * v.x + a.m[7] * v.y + a.m[8] * v.z); }
UG_DEV f3 tmulv(const m3& a, f3 v) { return mk3(a.m[0] * v.x + a.m[3] * v.y + a.m[6] * v.z, a.m[1] * v.x + a.m[4] * v.y + a.m[7] * v.z, a.m[2] * v.x + a.m[5] * v.y + a.m[8] * v.z); }
UG_DEV m3 mulm(const m3& a, const m3& b) { m3 c;
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) c.m[3 * i + j] = a.m[3 * i] * b.m[j] + a.m[3 * i + 1] * b.m[3 + j] + a.m[3 * i + 2] * b.m[6 + j];
  return c; }
// symmetric 3x3 stored xx xy xz yy yz zz
UG_DEV f3 symv(const real* s, f3 v) { return mk3(s[0] * v.x + s[1] * v.y + s[2] * v.z, s[1] * v.x + s[3] * v.y + s[4] * v.z, s[2] * v.x + s[4] * v.y + s[5] * v.z); }
// R S R^T for symmetric S (6 floats) -> symmetric (6 floats)
UG_DEV void rot_sym(const m3& R, const real* s, real* o) {
  f3 c0 = mulv(R, mk3(s[0], s[1], s[2])), c1 = mulv(R, mk3(s[1], s[3], s[4])), c2 = mulv(R, mk3(s[2], s[4], s[5]));  // R S (columns)
  // (R S) R^T: element (i,j) = sum_k (RS)[i][k] R[j][k]
  o[0] = c0.x * R.m[0] + c1.x * R.m[1] + c2.x * R.m[2];
  o[1] = c0.x * R.m[3] + c1.x * R.m[4] + c2.x * R.m[5];
  o[2] = c0.x * R.m[6] + c1.x * R.m[7] + c2.x * R.m[8];
  o[3] = c0.y * R.m[3] + c1.y * R.m[4] + c2.y * R.m[5];
  o[4] = c0.y * R.m[6] + c1.y * R.m[7] + c2.y * R.m[8];
  o[5] = c0.z * R.m[6] + c1.z * R.m[7] + c2.z * R.m[8];
}

struct q4 { real x, y, z, w; };
UG_DEV q4 mkq(real x, real y, real z, real w) { q4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
UG_DEV q4 qmul(q4 a, q4 b) {
  return mkq(a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
             a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z);
}
UG_DEV q4 qconj(q4 q) { return mkq(-q.x, -q.y, -q.z, q.w); }
UG_DEV q4 qnormalize(q4 q) { real n = rrsqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w); return mkq(q.x * n, q.y * n, q.z * n, q.w * n); }
UG_DEV m3 quat_to_mat(q4 q) {
  real d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w, s = 2.0_r / d;
  real xs = q.x * s, ys = q.y * s, zs = q.z * s, wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  real xx = q.x * xs, xy = q.x * ys, xz = q.x * zs, yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  m3 R; R.m[0] = 1.0_r - (yy + zz); R.m[1] = xy - wz; R.m[2] = xz + wy; R.m[3] = xy + wz; R.m[4] = 1.0_r - (xx + zz); R.m[5] = yz - wx;
  R.m[6] = xz - wy; R.m[7] = yz + wx; R.m[8] = 1.0_r - (xx + yy); return R;
}
UG_DEV q4 mat_to_quat(const m3& R) {
  real tr = R.m[0] + R.m[4] + R.m[8];
  if (tr > 0.0_r) { real s = rsqrt_(tr + 1.0_r); real w = s * 0.5_r; s = 0.5_r / s; return mkq((R.m[7] - R.m[5]) * s, (R.m[2] - R.m[6]) * s, (R.m[3] - R.m[1]) * s, w); }
  if (R.m[0] >= R.m[4] && R.m[0] >= R.m[8]) { real s = rsqrt_(R.m[0] - R.m[4] - R.m[8] + 1.0_r); real x = s * 0.5_r; s = 0.5_r / s; return mkq(x, (R.m[3] + R.m[1]) * s, (R.m[6] + R.m[2]) * s, (R.m[7] - R.m[5]) * s); }
  if (R.m[4] >= R.m[8]) { real s = rsqrt_(R.m[4] - R.m[8] - R.m[0] + 1.0_r); real y = s * 0.5_r; s = 0.5_r / s; return mkq((R.m[1] + R.m[3]) * s, y, (R.m[7] + R.m[5]) * s, (R.m[2] - R.m[6]) * s); }
  real s = rsqrt_(R.m[8] - R.m[0] - R.m[4] + 1.0_r); real z = s * 0.5_r; s = 0.5_r / s; return mkq((R.m[2] + R.m[6]) * s, (R.m[5] + R.m[7]) * s, z, (R.m[3] - R.m[1]) * s);
}
UG_DEV q4 quat_axis(int axis, real ang) { real s, c; rsincos(0.5_r * ang, &s, &c); return mkq(axis == 0 ? s : 0._r, axis == 1 ? s : 0._r, axis == 2 ? s : 0._r, c); }
UG_DEV q4 quat_from_euler(real r, real p, real y) {
  real sr, cr, sp, cp, sy, cy; rsincos(0.5_r * r, &sr, &cr); rsincos(0.5_r * p, &sp, &cp); rsincos(0.5_r * y, &sy, &cy);
  return qnormalize(mkq(sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy));
}
UG_DEV f3 euler_from_quat(q4 q) {
  real sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, squ = q.w * q.w;
  real sarg = -2.0_r * (q.x * q.z - q.w * q.y);
  const real PI = 3.14159265358979323846_r;
  if (sarg <= -0.99999_r) return mk3(0._r, -0.5_r * PI, 2.0_r * ratan2(q.x, -q.y));
  if (sarg >= 0.99999_r) return mk3(0._r, 0.5_r * PI, 2.0_r * ratan2(-q.x, q.y));
  return mk3(ratan2(2.0_r * (q.y * q.z + q.w * q.x), squ - sqx - sqy + sqz), rasin(sarg), ratan2(2.0_r * (q.x * q.y + q.w * q.z), squ + sqx - sqy - sqz));
}
UG_DEV void plane_space(f3 n, f3& p, f3& q) {
  if (rabs(n.z) > 0.7071067811865475244_r) { real a = n.y * n.y + n.z * n.z, k = rrsqrt(a); p = mk3(0._r, -n.z * k, n.y * k); q = mk3(a * k, -n.x * p.z, n.x * p.y); }
  else { real a = n.x * n.x + n.y * n.y, k = rrsqrt(a); p = mk3(-n.y * k, n.x * k, 0._r); q = mk3(-n.z * p.y, n.z * p.x, a * k); }
}
// 32-bit shared-memory addressing for the solver's hot loops (one IADD + LDS per operand, no generic-address math)
UG_DEV unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
#ifdef UG_F64
UG_DEV real lds(unsigned a) { real v; asm("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a)); return v; }
#else
UG_DEV real lds(unsigned a) { real v; asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
#endif
#define UG_RB ((unsigned)sizeof(real))  // bytes per scalar in shared-memory byte offsets
// ---- lane groups: G consecutive lanes of a warp run one env (G = 32: warp per env; G = 16: two envs per warp).
// All intra-env communication goes through these helpers, so the two halves of a warp never wait for each other.
template <int G>
struct Gp {
  int l;          // lane within the group
  unsigned mask;  // lanes of this group within the warp
  int shift;      // first lane of the group within the warp
  __device__ __forceinline__ void sync() const { __syncwarp(mask); }
  __device__ __forceinline__ real shfl(real v, int src) const { return __shfl_sync(mask, v, src, G); }
  __device__ __forceinline__ int shfl(int v, int src) const { return __shfl_sync(mask, v, src, G); }
  __device__ __forceinline__ unsigned ballot(bool p) const { return (__ballot_sync(mask, p) >> shift) & (G == 32 ? 0xffffffffu : ((1u << (G & 31)) - 1u)); }
  __device__ __forceinline__ bool any(bool p) const { return ballot(p) != 0u; }
  __device__ __forceinline__ real sum(real v) const {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, G);
    return v;
  }
  __device__ __forceinline__ real max(real v) const {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = rmax(v, __shfl_xor_sync(mask, v, o, G));
    return v;
  }
};
template <int G>
__device__ __forceinline__ Gp<G> make_gp() {
  Gp<G> g; const int lane = threadIdx.x & 31;
  g.l = lane & (G - 1); g.shift = lane & ~(G - 1);
  g.mask = G == 32 ? 0xffffffffu : (((1u << (G & 31)) - 1u) << g.shift);
  return g;
}

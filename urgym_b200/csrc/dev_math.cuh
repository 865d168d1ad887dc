// fp32 device helpers: 3-vectors, row-major 3x3, quaternions (x,y,z,w), Euler conventions of the reference's
// pybullet calls (getEulerFromQuaternion / getQuaternionFromEuler, env_ball_balance_v0.py:114-116,139).
#pragma once
#include <cuda_runtime.h>

#define UG_DEV __device__ __forceinline__

struct f3 { float x, y, z; };
UG_DEV f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
UG_DEV f3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
UG_DEV void st3(float* p, f3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
UG_DEV f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
UG_DEV f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
UG_DEV f3 operator-(f3 a) { return mk3(-a.x, -a.y, -a.z); }
UG_DEV f3 operator*(f3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
UG_DEV f3 operator*(float s, f3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
UG_DEV float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
UG_DEV f3 cross(f3 a, f3 b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
UG_DEV float norm3(f3 a) { return sqrtf(dot(a, a)); }
UG_DEV float comp(f3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }
UG_DEV void setc(f3& a, int i, float v) { if (i == 0) a.x = v; else if (i == 1) a.y = v; else a.z = v; }

struct m3 { float m[9]; };
UG_DEV m3 ldm(const float* p) { m3 r;
#pragma unroll
  for (int i = 0; i < 9; i++) r.m[i] = p[i]; return r; }
UG_DEV void stm(float* p, const m3& a) {
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
UG_DEV f3 symv(const float* s, f3 v) { return mk3(s[0] * v.x + s[1] * v.y + s[2] * v.z, s[1] * v.x + s[3] * v.y + s[4] * v.z, s[2] * v.x + s[4] * v.y + s[5] * v.z); }
// R S R^T for symmetric S (6 floats) -> symmetric (6 floats)
UG_DEV void rot_sym(const m3& R, const float* s, float* o) {
  f3 c0 = mulv(R, mk3(s[0], s[1], s[2])), c1 = mulv(R, mk3(s[1], s[3], s[4])), c2 = mulv(R, mk3(s[2], s[4], s[5]));  // R S (columns)
  // (R S) R^T: element (i,j) = sum_k (RS)[i][k] R[j][k]
  o[0] = c0.x * R.m[0] + c1.x * R.m[1] + c2.x * R.m[2];
  o[1] = c0.x * R.m[3] + c1.x * R.m[4] + c2.x * R.m[5];
  o[2] = c0.x * R.m[6] + c1.x * R.m[7] + c2.x * R.m[8];
  o[3] = c0.y * R.m[3] + c1.y * R.m[4] + c2.y * R.m[5];
  o[4] = c0.y * R.m[6] + c1.y * R.m[7] + c2.y * R.m[8];
  o[5] = c0.z * R.m[6] + c1.z * R.m[7] + c2.z * R.m[8];
}

UG_DEV float rsqrt_ftz(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

struct q4 { float x, y, z, w; };
UG_DEV q4 mkq(float x, float y, float z, float w) { q4 r; r.x = x; r.y = y; r.z = z; r.w = w; return r; }
UG_DEV q4 qmul(q4 a, q4 b) {
  return mkq(a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
             a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z);
}
UG_DEV q4 qconj(q4 q) { return mkq(-q.x, -q.y, -q.z, q.w); }
UG_DEV q4 qnormalize(q4 q) { float n = rsqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w); return mkq(q.x * n, q.y * n, q.z * n, q.w * n); }
UG_DEV m3 quat_to_mat(q4 q) {
  float d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w, s = 2.0f / d;
  float xs = q.x * s, ys = q.y * s, zs = q.z * s, wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  float xx = q.x * xs, xy = q.x * ys, xz = q.x * zs, yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  m3 R; R.m[0] = 1.0f - (yy + zz); R.m[1] = xy - wz; R.m[2] = xz + wy; R.m[3] = xy + wz; R.m[4] = 1.0f - (xx + zz); R.m[5] = yz - wx;
  R.m[6] = xz - wy; R.m[7] = yz + wx; R.m[8] = 1.0f - (xx + yy); return R;
}
UG_DEV q4 mat_to_quat(const m3& R) {
  float tr = R.m[0] + R.m[4] + R.m[8];
  if (tr > 0.0f) { float s = sqrtf(tr + 1.0f); float w = s * 0.5f; s = 0.5f / s; return mkq((R.m[7] - R.m[5]) * s, (R.m[2] - R.m[6]) * s, (R.m[3] - R.m[1]) * s, w); }
  if (R.m[0] >= R.m[4] && R.m[0] >= R.m[8]) { float s = sqrtf(R.m[0] - R.m[4] - R.m[8] + 1.0f); float x = s * 0.5f; s = 0.5f / s; return mkq(x, (R.m[3] + R.m[1]) * s, (R.m[6] + R.m[2]) * s, (R.m[7] - R.m[5]) * s); }
  if (R.m[4] >= R.m[8]) { float s = sqrtf(R.m[4] - R.m[8] - R.m[0] + 1.0f); float y = s * 0.5f; s = 0.5f / s; return mkq((R.m[1] + R.m[3]) * s, y, (R.m[7] + R.m[5]) * s, (R.m[2] - R.m[6]) * s); }
  float s = sqrtf(R.m[8] - R.m[0] - R.m[4] + 1.0f); float z = s * 0.5f; s = 0.5f / s; return mkq((R.m[2] + R.m[6]) * s, (R.m[5] + R.m[7]) * s, z, (R.m[3] - R.m[1]) * s);
}
UG_DEV q4 quat_axis(int axis, float ang) { float s, c; sincosf(0.5f * ang, &s, &c); return mkq(axis == 0 ? s : 0.f, axis == 1 ? s : 0.f, axis == 2 ? s : 0.f, c); }
UG_DEV q4 quat_from_euler(float r, float p, float y) {
  float sr, cr, sp, cp, sy, cy; sincosf(0.5f * r, &sr, &cr); sincosf(0.5f * p, &sp, &cp); sincosf(0.5f * y, &sy, &cy);
  return qnormalize(mkq(sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy, cr * cp * cy + sr * sp * sy));
}
UG_DEV f3 euler_from_quat(q4 q) {
  float sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, squ = q.w * q.w;
  float sarg = -2.0f * (q.x * q.z - q.w * q.y);
  const float PI = 3.14159265358979323846f;
  if (sarg <= -0.99999f) return mk3(0.f, -0.5f * PI, 2.0f * atan2f(q.x, -q.y));
  if (sarg >= 0.99999f) return mk3(0.f, 0.5f * PI, 2.0f * atan2f(-q.x, q.y));
  return mk3(atan2f(2.0f * (q.y * q.z + q.w * q.x), squ - sqx - sqy + sqz), asinf(sarg), atan2f(2.0f * (q.x * q.y + q.w * q.z), squ + sqx - sqy - sqz));
}
UG_DEV void plane_space(f3 n, f3& p, f3& q) {
  if (fabsf(n.z) > 0.7071067811865475244f) { float a = n.y * n.y + n.z * n.z, k = rsqrtf(a); p = mk3(0.f, -n.z * k, n.y * k); q = mk3(a * k, -n.x * p.z, n.x * p.y); }
  else { float a = n.x * n.x + n.y * n.y, k = rsqrtf(a); p = mk3(-n.y * k, n.x * k, 0.f); q = mk3(-n.z * p.y, n.z * p.x, a * k); }
}
// 32-bit shared-memory addressing for the solver's hot loops (one IADD + LDS per operand, no generic-address math)
UG_DEV unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
UG_DEV float lds(unsigned a) { float v; asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
// ---- lane groups: G consecutive lanes of a warp run one env (G = 32: warp per env; G = 16: two envs per warp).
// All intra-env communication goes through these helpers, so the two halves of a warp never wait for each other.
template <int G>
struct Gp {
  int l;          // lane within the group
  unsigned mask;  // lanes of this group within the warp
  int shift;      // first lane of the group within the warp
  __device__ __forceinline__ void sync() const { __syncwarp(mask); }
  __device__ __forceinline__ float shfl(float v, int src) const { return __shfl_sync(mask, v, src, G); }
  __device__ __forceinline__ int shfl(int v, int src) const { return __shfl_sync(mask, v, src, G); }
  __device__ __forceinline__ unsigned ballot(bool p) const { return (__ballot_sync(mask, p) >> shift) & (G == 32 ? 0xffffffffu : ((1u << (G & 31)) - 1u)); }
  __device__ __forceinline__ bool any(bool p) const { return ballot(p) != 0u; }
  __device__ __forceinline__ float sum(float v) const {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o, G);
    return v;
  }
  __device__ __forceinline__ float max(float v) const {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(mask, v, o, G));
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

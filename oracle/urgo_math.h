// TEST INFRASTRUCTURE — part of the CPU oracle (see oracle/README.md). fp64 vector / quaternion / spatial helpers.
#pragma once
#include <cmath>

namespace urgo {

struct V3 {
  double x, y, z;
  V3() : x(0), y(0), z(0) {}
  V3(double a, double b, double c) : x(a), y(b), z(c) {}
  double& operator[](int i) { return (&x)[i]; }
  double operator[](int i) const { return (&x)[i]; }
};
inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, double s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(double s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3& operator+=(V3& a, V3 b) { a = a + b; return a; }
inline V3& operator-=(V3& a, V3 b) { a = a - b; return a; }
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }

// Row-major 3x3. Columns are the body axes expressed in the parent/world frame.
struct M3 {
  double m[9];
  double& operator()(int r, int c) { return m[3 * r + c]; }
  double operator()(int r, int c) const { return m[3 * r + c]; }
  V3 col(int c) const { return V3(m[c], m[3 + c], m[6 + c]); }
  static M3 identity() { M3 r; for (int i = 0; i < 9; i++) r.m[i] = (i % 4 == 0); return r; }
  static M3 zero() { M3 r; for (int i = 0; i < 9; i++) r.m[i] = 0; return r; }
};
inline V3 operator*(const M3& A, V3 v) {
  return V3(A.m[0] * v.x + A.m[1] * v.y + A.m[2] * v.z, A.m[3] * v.x + A.m[4] * v.y + A.m[5] * v.z,
            A.m[6] * v.x + A.m[7] * v.y + A.m[8] * v.z);
}
inline V3 tmul(const M3& A, V3 v) {  // A^T v
  return V3(A.m[0] * v.x + A.m[3] * v.y + A.m[6] * v.z, A.m[1] * v.x + A.m[4] * v.y + A.m[7] * v.z,
            A.m[2] * v.x + A.m[5] * v.y + A.m[8] * v.z);
}
inline M3 operator*(const M3& A, const M3& B) {
  M3 C;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) C.m[3 * i + j] = A.m[3 * i] * B.m[j] + A.m[3 * i + 1] * B.m[3 + j] + A.m[3 * i + 2] * B.m[6 + j];
  return C;
}
inline M3 transpose(const M3& A) {
  M3 C;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) C.m[3 * i + j] = A.m[3 * j + i];
  return C;
}
inline M3 operator+(const M3& A, const M3& B) { M3 C; for (int i = 0; i < 9; i++) C.m[i] = A.m[i] + B.m[i]; return C; }
// R diag(d) R^T
inline M3 rot_diag(const M3& R, V3 d) {
  M3 C;
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) C.m[3 * i + j] = R.m[3 * i] * d.x * R.m[3 * j] + R.m[3 * i + 1] * d.y * R.m[3 * j + 1] + R.m[3 * i + 2] * d.z * R.m[3 * j + 2];
  return C;
}

// Quaternion (x, y, z, w), Hamilton product, as pybullet reports them.
struct Q4 {
  double x, y, z, w;
  Q4() : x(0), y(0), z(0), w(1) {}
  Q4(double a, double b, double c, double d) : x(a), y(b), z(c), w(d) {}
};
inline Q4 operator*(Q4 a, Q4 b) {
  return Q4(a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
            a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x, a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z);
}
inline Q4 conj(Q4 q) { return Q4(-q.x, -q.y, -q.z, q.w); }
inline Q4 normalized(Q4 q) {
  double n = std::sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  return Q4(q.x / n, q.y / n, q.z / n, q.w / n);
}
inline M3 quat_to_mat(Q4 q) {  // btMatrix3x3::setRotation
  double d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
  double s = 2.0 / d;
  double xs = q.x * s, ys = q.y * s, zs = q.z * s;
  double wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  double xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
  double yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  M3 R;
  R.m[0] = 1.0 - (yy + zz); R.m[1] = xy - wz; R.m[2] = xz + wy;
  R.m[3] = xy + wz; R.m[4] = 1.0 - (xx + zz); R.m[5] = yz - wx;
  R.m[6] = xz - wy; R.m[7] = yz + wx; R.m[8] = 1.0 - (xx + yy);
  return R;
}
inline Q4 mat_to_quat(const M3& R) {  // btMatrix3x3::getRotation
  double tr = R.m[0] + R.m[4] + R.m[8];
  double t[4];
  if (tr > 0.0) {
    double s = std::sqrt(tr + 1.0);
    t[3] = s * 0.5;
    s = 0.5 / s;
    t[0] = (R(2, 1) - R(1, 2)) * s;
    t[1] = (R(0, 2) - R(2, 0)) * s;
    t[2] = (R(1, 0) - R(0, 1)) * s;
  } else {
    int i = R(0, 0) < R(1, 1) ? (R(1, 1) < R(2, 2) ? 2 : 1) : (R(0, 0) < R(2, 2) ? 2 : 0);
    int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = std::sqrt(R(i, i) - R(j, j) - R(k, k) + 1.0);
    t[i] = s * 0.5;
    s = 0.5 / s;
    t[3] = (R(k, j) - R(j, k)) * s;
    t[j] = (R(j, i) + R(i, j)) * s;
    t[k] = (R(k, i) + R(i, k)) * s;
  }
  return Q4(t[0], t[1], t[2], t[3]);
}
inline Q4 axis_angle(V3 a, double ang) {
  double s = std::sin(0.5 * ang) / norm(a);
  return Q4(a.x * s, a.y * s, a.z * s, std::cos(0.5 * ang));
}

// pybullet getQuaternionFromEuler: roll-pitch-yaw about fixed X, Y, Z [recalled]; call sites env_ball_balance_v0.py:114-116
inline Q4 quat_from_euler(double r, double p, double y) {
  double phi = r * 0.5, the = p * 0.5, psi = y * 0.5;
  Q4 q(std::sin(phi) * std::cos(the) * std::cos(psi) - std::cos(phi) * std::sin(the) * std::sin(psi),
       std::cos(phi) * std::sin(the) * std::cos(psi) + std::sin(phi) * std::cos(the) * std::sin(psi),
       std::cos(phi) * std::cos(the) * std::sin(psi) - std::sin(phi) * std::sin(the) * std::cos(psi),
       std::cos(phi) * std::cos(the) * std::cos(psi) + std::sin(phi) * std::sin(the) * std::sin(psi));
  return normalized(q);
}
// pybullet getEulerFromQuaternion, including its gimbal-lock branch [recalled]
inline V3 euler_from_quat(Q4 q) {
  double sqx = q.x * q.x, sqy = q.y * q.y, sqz = q.z * q.z, squ = q.w * q.w;
  double sarg = -2.0 * (q.x * q.z - q.w * q.y);
  const double PI = 3.14159265358979323846;
  if (sarg <= -0.99999) return V3(0.0, -0.5 * PI, 2.0 * std::atan2(q.x, -q.y));
  if (sarg >= 0.99999) return V3(0.0, 0.5 * PI, 2.0 * std::atan2(-q.x, q.y));
  return V3(std::atan2(2.0 * (q.y * q.z + q.w * q.x), squ - sqx - sqy + sqz), std::asin(sarg),
            std::atan2(2.0 * (q.x * q.y + q.w * q.z), squ + sqx - sqy - sqz));
}

// btPlaneSpace1: two tangents for a unit normal [recalled]
inline void plane_space(V3 n, V3& p, V3& q) {
  if (std::fabs(n.z) > 0.7071067811865475244) {
    double a = n.y * n.y + n.z * n.z;
    double k = 1.0 / std::sqrt(a);
    p = V3(0, -n.z * k, n.y * k);
    q = V3(a * k, -n.x * p.z, n.x * p.y);
  } else {
    double a = n.x * n.x + n.y * n.y;
    double k = 1.0 / std::sqrt(a);
    p = V3(-n.y * k, n.x * k, 0);
    q = V3(-n.z * p.y, n.z * p.x, a * k);
  }
}

}  // namespace urgo

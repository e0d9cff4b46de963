// TEST INFRASTRUCTURE ONLY — part of the CPU oracle (see oracle/README.md).
// Minimal FP64 3-vector / rigid-pose arithmetic standing in for the Eigen types the reference uses
// (Eigen::Vector3d, Eigen::Isometry3d; e.g. trajopt/src/collision_terms.cpp:215-223).
#pragma once
#include <cmath>

namespace orc
{
struct V3
{
  double x{ 0 }, y{ 0 }, z{ 0 };
  V3() = default;
  V3(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
  double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
  double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};
inline V3 operator+(V3 a, V3 b) { return { a.x + b.x, a.y + b.y, a.z + b.z }; }
inline V3 operator-(V3 a, V3 b) { return { a.x - b.x, a.y - b.y, a.z - b.z }; }
inline V3 operator-(V3 a) { return { -a.x, -a.y, -a.z }; }
inline V3 operator*(double s, V3 a) { return { s * a.x, s * a.y, s * a.z }; }
inline V3 operator*(V3 a, double s) { return { s * a.x, s * a.y, s * a.z }; }
inline V3& operator+=(V3& a, V3 b)
{
  a = a + b;
  return a;
}
inline double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return { a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x }; }
inline double norm2(V3 a) { return dot(a, a); }
inline double norm(V3 a) { return std::sqrt(dot(a, a)); }

// Rigid transform: rotation rows r[0..2], translation t.  Memory layout of the 12-double pose of
// include/trajopt_cuda.h: [r00 r01 r02 tx  r10 r11 r12 ty  r20 r21 r22 tz].
struct Pose
{
  double R[3][3]{ { 1, 0, 0 }, { 0, 1, 0 }, { 0, 0, 1 } };
  V3 t;
  static Pose from12(const double* p)
  {
    Pose T;
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        T.R[i][j] = p[4 * i + j];
      T.t[i] = p[4 * i + 3];
    }
    return T;
  }
  void to12(double* p) const
  {
    for (int i = 0; i < 3; ++i)
    {
      for (int j = 0; j < 3; ++j)
        p[4 * i + j] = R[i][j];
      p[4 * i + 3] = t[i];
    }
  }
  V3 rot(V3 v) const
  {
    return { R[0][0] * v.x + R[0][1] * v.y + R[0][2] * v.z, R[1][0] * v.x + R[1][1] * v.y + R[1][2] * v.z,
             R[2][0] * v.x + R[2][1] * v.y + R[2][2] * v.z };
  }
  V3 rotT(V3 v) const  // R^T v
  {
    return { R[0][0] * v.x + R[1][0] * v.y + R[2][0] * v.z, R[0][1] * v.x + R[1][1] * v.y + R[2][1] * v.z,
             R[0][2] * v.x + R[1][2] * v.y + R[2][2] * v.z };
  }
  V3 apply(V3 v) const { return rot(v) + t; }
  V3 applyInv(V3 p) const { return rotT(p - t); }
};

inline Pose operator*(const Pose& A, const Pose& B)
{
  Pose C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      C.R[i][j] = A.R[i][0] * B.R[0][j] + A.R[i][1] * B.R[1][j] + A.R[i][2] * B.R[2][j];
  C.t = A.apply(B.t);
  return C;
}

// sin / cos of a joint angle.  The reference calls libm through Eigen::AngleAxisd (KDL / tesseract FK); libm's and
// CUDA's results differ in the last bit for a fraction of the arguments, and GJK on a curved surface amplifies such
// a difference to 1e-5 in the contact normal.  The checker and the checked therefore evaluate ONE operation sequence
// (Cody-Waite reduction by pi/2 in three parts + the classic minimax polynomials on [-pi/4, pi/4], every operation
// rounded on its own; built with -ffp-contract=off), restated here from the same published constants as
// trajopt-1_b200/csrc/tc_math.cuh tc_sincos; absolute error <= 2e-16 for |x| <= 1e5, libm beyond.
inline void sincosShared(double x, double& sn, double& cs)
{
  if (!(std::fabs(x) <= 1.0e5))
  {
    sn = std::sin(x);
    cs = std::cos(x);
    return;
  }
  const double fn = std::floor(x * 6.36619772367581382433e-01 + 0.5);
  double r = x - fn * 1.57079632673412561417e+00;
  r = r - fn * 6.07710050630396597660e-11;
  r = r - fn * 2.02226624879595063154e-21;
  const double z = r * r;
  double ps = z * 1.58969099521155010221e-10;
  ps = z * (-2.50507602534068634195e-08 + ps);
  ps = z * (2.75573137070700676789e-06 + ps);
  ps = z * (-1.98412698298579493134e-04 + ps);
  ps = z * (8.33333333332248946124e-03 + ps);
  ps = -1.66666666666666324348e-01 + ps;
  const double s = r + (r * z) * ps;
  double pc = z * -1.13596475577881948265e-11;
  pc = z * (2.08757232129817482790e-09 + pc);
  pc = z * (-2.75573143513906633035e-07 + pc);
  pc = z * (2.48015872894767294178e-05 + pc);
  pc = z * (-1.38888888888741095749e-03 + pc);
  pc = 4.16666666666666019037e-02 + pc;
  const double c = (1.0 - 0.5 * z) + (z * z) * pc;
  const int q = static_cast<int>(static_cast<long long>(fn) & 3);
  sn = (q == 0) ? s : ((q == 1) ? c : ((q == 2) ? -s : -c));
  cs = (q == 0) ? c : ((q == 1) ? -s : ((q == 2) ? -c : s));
}

// Rodrigues rotation about a unit axis
inline Pose axisAngle(V3 a, double q)
{
  Pose T;
  double c, s;
  sincosShared(q, s, c);
  const double v = 1 - c;
  T.R[0][0] = c + a.x * a.x * v;
  T.R[0][1] = a.x * a.y * v - a.z * s;
  T.R[0][2] = a.x * a.z * v + a.y * s;
  T.R[1][0] = a.y * a.x * v + a.z * s;
  T.R[1][1] = c + a.y * a.y * v;
  T.R[1][2] = a.y * a.z * v - a.x * s;
  T.R[2][0] = a.z * a.x * v - a.y * s;
  T.R[2][1] = a.z * a.y * v + a.x * s;
  T.R[2][2] = c + a.z * a.z * v;
  return T;
}
}  // namespace orc

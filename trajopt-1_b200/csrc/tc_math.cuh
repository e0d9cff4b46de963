// tc_math.cuh — 3-vector / rigid-pose arithmetic shared by every kernel of libtrajopt_cuda.so.
//
// Everything here is `TC_HD` so that the numerics of the hot path (support mappings, GJK, EPA,
// continuous-contact bookkeeping, gradient rows) can also be compiled by the host compiler into the
// unit-test harness tests/hostcheck (never into the product library's call path).
//
// The reference computes in FP64 through Eigen::Isometry3d / Eigen::VectorXd
// (e.g. trajopt/src/collision_terms.cpp:215-223, trajopt_common/src/collision_utils.cpp:128-159);
// `Real` is double on the parity path and float in the stated-tolerance FP32 mode.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define TC_HD __host__ __device__ __forceinline__
#define TC_HDN __host__ __device__
#else
#define TC_HD inline
#define TC_HDN
#endif

namespace tc
{
template <class R>
struct Vec3
{
  R x, y, z;
  TC_HD Vec3() : x(0), y(0), z(0) {}
  TC_HD Vec3(R x_, R y_, R z_) : x(x_), y(y_), z(z_) {}
  TC_HD R get(int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
template <class R>
TC_HD Vec3<R> operator+(Vec3<R> a, Vec3<R> b)
{
  return Vec3<R>(a.x + b.x, a.y + b.y, a.z + b.z);
}
template <class R>
TC_HD Vec3<R> operator-(Vec3<R> a, Vec3<R> b)
{
  return Vec3<R>(a.x - b.x, a.y - b.y, a.z - b.z);
}
template <class R>
TC_HD Vec3<R> operator-(Vec3<R> a)
{
  return Vec3<R>(-a.x, -a.y, -a.z);
}
template <class R>
TC_HD Vec3<R> operator*(R s, Vec3<R> a)
{
  return Vec3<R>(s * a.x, s * a.y, s * a.z);
}
template <class R>
TC_HD R dot(Vec3<R> a, Vec3<R> b)
{
  return a.x * b.x + a.y * b.y + a.z * b.z;
}
template <class R>
TC_HD Vec3<R> cross(Vec3<R> a, Vec3<R> b)
{
  return Vec3<R>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
template <class R>
TC_HD R norm2(Vec3<R> a)
{
  return dot(a, a);
}
TC_HD double tc_sqrt(double v) { return sqrt(v); }
TC_HD float tc_sqrt(float v) { return sqrtf(v); }
TC_HD double tc_abs(double v) { return fabs(v); }
TC_HD float tc_abs(float v) { return fabsf(v); }
template <class R>
TC_HD R norm(Vec3<R> a)
{
  return tc_sqrt(dot(a, a));
}
template <class R>
TC_HD R tc_max(R a, R b)
{
  return a > b ? a : b;
}
template <class R>
TC_HD R tc_min(R a, R b)
{
  return a < b ? a : b;
}

// Rigid transform, row-major rotation r[3*i+j] and translation t.  The 12-double pose of
// include/trajopt_cuda.h is [r00 r01 r02 tx r10 r11 r12 ty r20 r21 r22 tz].
template <class R>
struct Pose
{
  R r[9];
  Vec3<R> t;
  TC_HD Pose()
  {
    r[0] = 1, r[1] = 0, r[2] = 0, r[3] = 0, r[4] = 1, r[5] = 0, r[6] = 0, r[7] = 0, r[8] = 1;
  }
  template <class S>
  TC_HD static Pose from12(const S* p)
  {
    Pose T;
    for (int i = 0; i < 3; ++i)
    {
      T.r[3 * i + 0] = static_cast<R>(p[4 * i + 0]);
      T.r[3 * i + 1] = static_cast<R>(p[4 * i + 1]);
      T.r[3 * i + 2] = static_cast<R>(p[4 * i + 2]);
    }
    T.t = Vec3<R>(static_cast<R>(p[3]), static_cast<R>(p[7]), static_cast<R>(p[11]));
    return T;
  }
  template <class S>
  TC_HD void to12(S* p) const
  {
    for (int i = 0; i < 3; ++i)
    {
      p[4 * i + 0] = static_cast<S>(r[3 * i + 0]);
      p[4 * i + 1] = static_cast<S>(r[3 * i + 1]);
      p[4 * i + 2] = static_cast<S>(r[3 * i + 2]);
    }
    p[3] = static_cast<S>(t.x);
    p[7] = static_cast<S>(t.y);
    p[11] = static_cast<S>(t.z);
  }
  TC_HD Vec3<R> rot(Vec3<R> v) const
  {
    return Vec3<R>(r[0] * v.x + r[1] * v.y + r[2] * v.z, r[3] * v.x + r[4] * v.y + r[5] * v.z,
                   r[6] * v.x + r[7] * v.y + r[8] * v.z);
  }
  TC_HD Vec3<R> rotT(Vec3<R> v) const
  {
    return Vec3<R>(r[0] * v.x + r[3] * v.y + r[6] * v.z, r[1] * v.x + r[4] * v.y + r[7] * v.z,
                   r[2] * v.x + r[5] * v.y + r[8] * v.z);
  }
  TC_HD Vec3<R> apply(Vec3<R> v) const { return rot(v) + t; }
  TC_HD Vec3<R> applyInv(Vec3<R> p) const { return rotT(p - t); }
};

template <class R>
TC_HD Pose<R> compose(const Pose<R>& A, const Pose<R>& B)
{
  Pose<R> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      C.r[3 * i + j] = A.r[3 * i + 0] * B.r[0 + j] + A.r[3 * i + 1] * B.r[3 + j] + A.r[3 * i + 2] * B.r[6 + j];
  C.t = A.apply(B.t);
  return C;
}

// sin and cos of a joint angle with ONE operation sequence for every compiler and target: Cody-Waite reduction by
// pi/2 in three parts, then the classic degree-13 / degree-14 minimax polynomials on [-pi/4, pi/4].  Every product and
// sum is rounded on its own (no FMA contraction: intrinsics on the device, -ffp-contract=off on the host), so the
// device FK, the host-compiled check harness and the CPU checker (which restates the same sequence) produce
// bit-identical poses from the same joint values.  libm's and CUDA's own sin/cos differ from each other in the last
// bit for a fraction of the arguments, which is enough to send GJK on a curved surface down another path.
// Absolute error <= 2e-16 for |x| <= 1e5 (0.9 ulp at 1; checked against 200-bit arithmetic); beyond that, libm.
#if defined(__CUDA_ARCH__)
#define TC_MUL(a, b) __dmul_rn((a), (b))
#define TC_ADD(a, b) __dadd_rn((a), (b))
#define TC_SUB(a, b) __dsub_rn((a), (b))
#else
#define TC_MUL(a, b) ((a) * (b))
#define TC_ADD(a, b) ((a) + (b))
#define TC_SUB(a, b) ((a) - (b))
#endif
TC_HD void tc_sincos(double x, double* sn, double* cs)
{
  if (!(fabs(x) <= 1.0e5))
  {
#if defined(__CUDA_ARCH__)
    sincos(x, sn, cs);
#else
    *sn = sin(x);
    *cs = cos(x);
#endif
    return;
  }
  const double INV_PIO2 = 6.36619772367581382433e-01;
  const double P1 = 1.57079632673412561417e+00;  // first 33 bits of pi/2
  const double P2 = 6.07710050630396597660e-11;  // next 33 bits
  const double P3 = 2.02226624879595063154e-21;  // pi/2 - P1 - P2
  const double fn = floor(TC_ADD(TC_MUL(x, INV_PIO2), 0.5));
  double r = TC_SUB(x, TC_MUL(fn, P1));
  r = TC_SUB(r, TC_MUL(fn, P2));
  r = TC_SUB(r, TC_MUL(fn, P3));
  const double z = TC_MUL(r, r);
  double ps = TC_MUL(z, 1.58969099521155010221e-10);
  ps = TC_MUL(z, TC_ADD(-2.50507602534068634195e-08, ps));
  ps = TC_MUL(z, TC_ADD(2.75573137070700676789e-06, ps));
  ps = TC_MUL(z, TC_ADD(-1.98412698298579493134e-04, ps));
  ps = TC_MUL(z, TC_ADD(8.33333333332248946124e-03, ps));
  ps = TC_ADD(-1.66666666666666324348e-01, ps);
  const double s = TC_ADD(r, TC_MUL(TC_MUL(r, z), ps));
  double pc = TC_MUL(z, -1.13596475577881948265e-11);
  pc = TC_MUL(z, TC_ADD(2.08757232129817482790e-09, pc));
  pc = TC_MUL(z, TC_ADD(-2.75573143513906633035e-07, pc));
  pc = TC_MUL(z, TC_ADD(2.48015872894767294178e-05, pc));
  pc = TC_MUL(z, TC_ADD(-1.38888888888741095749e-03, pc));
  pc = TC_ADD(4.16666666666666019037e-02, pc);
  const double c = TC_ADD(TC_SUB(1.0, TC_MUL(0.5, z)), TC_MUL(TC_MUL(z, z), pc));
  const int q = static_cast<int>(static_cast<long long>(fn) & 3);
  *sn = (q == 0) ? s : ((q == 1) ? c : ((q == 2) ? -s : -c));
  *cs = (q == 0) ? c : ((q == 1) ? -s : ((q == 2) ? -c : s));
}

// C = A * Rot(axis, q)   (axis unit, Rodrigues)
TC_HD Pose<double> rotateAbout(const Pose<double>& A, Vec3<double> a, double q)
{
  double s, c;
  tc_sincos(q, &s, &c);
  const double v = 1 - c;
  Pose<double> Rq;
  Rq.r[0] = c + a.x * a.x * v;
  Rq.r[1] = a.x * a.y * v - a.z * s;
  Rq.r[2] = a.x * a.z * v + a.y * s;
  Rq.r[3] = a.y * a.x * v + a.z * s;
  Rq.r[4] = c + a.y * a.y * v;
  Rq.r[5] = a.y * a.z * v - a.x * s;
  Rq.r[6] = a.z * a.x * v - a.y * s;
  Rq.r[7] = a.z * a.y * v + a.x * s;
  Rq.r[8] = c + a.z * a.z * v;
  Pose<double> C = compose(A, Rq);
  C.t = A.t;
  return C;
}
}  // namespace tc

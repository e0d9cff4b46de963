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

// C = A * Rot(axis, q)   (axis unit, Rodrigues)
TC_HD Pose<double> rotateAbout(const Pose<double>& A, Vec3<double> a, double q)
{
  double s, c;
#if defined(__CUDA_ARCH__)
  sincos(q, &s, &c);
#else
  s = sin(q);
  c = cos(q);
#endif
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

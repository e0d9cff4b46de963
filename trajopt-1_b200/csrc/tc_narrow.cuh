// tc_narrow.cuh — per-thread convex-convex signed distance (GJK + EPA) with swept-hull support.
//
// Stands where the reference calls contact_manager_->contactTest(...)
//   trajopt/src/collision_terms.cpp:667,889,1134,1157
//   trajopt_ifopt/src/constraints/collision/discrete_collision_evaluators.cpp:166
//   trajopt_ifopt/src/constraints/collision/continuous_collision_evaluators.cpp:228,244,449
// i.e. tesseract_collision's Bullet convex-convex narrow phase (GJK distance, EPA penetration) and its
// CastHullShape support rule for the continuous managers [upstream of the reference tree].
//
// GPU mapping: ONE THREAD PER (unit, shape pair) item; the 32 lanes of a warp always hold items of
// the SAME shape pair (the broadphase buckets items per shape pair), so every hull-vertex read in the
// support scan is a shared-memory broadcast and the scan trip count is warp-uniform.  A swept (cast)
// shape scans its vertices once for both end poses.  The simplex lives in registers: all indexing is
// static after unrolling, sub-simplex compaction is done with selects.
#pragma once
#include "tc_math.cuh"

namespace tc
{
enum
{
  SH_SPHERE = 0,
  SH_HULL = 1,
  SH_CAPSULE = 2,
  SH_CYLINDER = 3,
  SH_CONE = 4
};

// hull vertex as stored on the device: 32 B (FP64) so one 16-byte and one 8-byte broadcast load fetch it
template <class R>
struct alignas(16) Vtx
{
  R x, y, z, w;
};

// Collision shape in its LINK frame (shape_pose of the description already applied).
struct ShapeD
{
  int32_t type, link, shape_id, vert_off, vert_cnt, active_index;
  int32_t lut_off, lut_bytes;  // support look-up table of a hull in the scene's LUT blob (bytes, 16-aligned); -1 = none
  double radius, half_len;  // sphere/capsule: inflation radius of the core; cylinder/cone: solid radius
  double center[3], axis[3];
  double aabb_c[3], aabb_h[3];
  double brad;  // bounding radius of the core about `center`
};

// ---------------------------------------------------------------------------------------------
// Support look-up table of a hull.  Directions are binned on a cube map (6 faces x LUT_G x LUT_G cells); a cell
// lists, in ascending index order, every vertex that can attain the maximum of d.v for SOME direction d of the cell
// (built conservatively on the host, tc_scene.hpp buildSupportLut: a vertex is left out only when one other vertex
// beats it by a margin over the whole cone around the cell).  Scanning the list with the same compare therefore
// returns the vertex a scan of all vertices returns, ties included.  Layout: uint16 off[LUT_CELLS + 1], uint8 cand[].
// ---------------------------------------------------------------------------------------------
enum
{
  LUT_G = 8,
  LUT_CELLS = 6 * LUT_G * LUT_G,
  LUT_MIN_VERTS = 13,  // smaller hulls are scanned in full
  LUT_MAX_VERTS = 255
};

// cube-map cell of direction (x, y, z); -1 when the direction cannot be binned (zero / non-finite in float)
TC_HD int lutCellF(float x, float y, float z)
{
  const float ax = x < 0 ? -x : x, ay = y < 0 ? -y : y, az = z < 0 ? -z : z;
  int m;
  float mj, u, v;
  if (ax >= ay && ax >= az)
    m = 0, mj = x, u = y, v = z;
  else if (ay >= az)
    m = 1, mj = y, u = z, v = x;
  else
    m = 2, mj = z, u = x, v = y;
  const float am = mj < 0 ? -mj : mj;
  if (!(am > 1e-30f) || !(am < 1e30f))
    return -1;
  const float inv = (0.5f * LUT_G) / am;
  int iu = static_cast<int>(u * inv + 0.5f * LUT_G), iv = static_cast<int>(v * inv + 0.5f * LUT_G);
  iu = iu < 0 ? 0 : (iu > LUT_G - 1 ? LUT_G - 1 : iu);
  iv = iv < 0 ? 0 : (iv > LUT_G - 1 ? LUT_G - 1 : iv);
  return ((2 * m + (mj < 0 ? 1 : 0)) * LUT_G + iv) * LUT_G + iu;
}

// Work counter of the support queries of one item: low half = vertices of the queried hulls (what the reference's
// plain scan of every vertex touches: the ALGORITHMIC figure of SURVEY.md 8(d)), high half = vertices actually compared
// (the look-up table leaves ~5 % of a hull) -> roofline.frac and roofline.frac_executed of bench.py
typedef unsigned long long ScanCount;
TC_HD void countScan(ScanCount& s, int full, int compared)
{
  s += static_cast<unsigned long long>(static_cast<unsigned>(full)) | (static_cast<unsigned long long>(static_cast<unsigned>(compared)) << 32);
}

template <class R>
struct ShapeCore
{
  int type, nverts;
  const Vtx<R>* verts;
  const uint16_t* lut;  // support look-up table (see above) or null
  R radius, half_len;
  Vec3<R> center, axis;
  TC_HD R coreRadius() const { return (type == SH_SPHERE || type == SH_CAPSULE) ? radius : R(0); }
};

template <class R>
TC_HD ShapeCore<R> makeCore(const ShapeD& s, const Vtx<R>* verts, const void* lut = nullptr)
{
  ShapeCore<R> c;
  c.type = s.type;
  c.nverts = s.vert_cnt;
  c.verts = verts;
  c.lut = static_cast<const uint16_t*>(lut);
  c.radius = static_cast<R>(s.radius);
  c.half_len = static_cast<R>(s.half_len);
  c.center = Vec3<R>(static_cast<R>(s.center[0]), static_cast<R>(s.center[1]), static_cast<R>(s.center[2]));
  c.axis = Vec3<R>(static_cast<R>(s.axis[0]), static_cast<R>(s.axis[1]), static_cast<R>(s.axis[2]));
  return c;
}

template <class R>
TC_HD R lowestReal();
template <>
TC_HD double lowestReal<double>()
{
  return -1.7976931348623157e308;
}
template <>
TC_HD float lowestReal<float>()
{
  return -3.402823466e38f;
}

template <class R>
struct Tol;
template <>
struct Tol<double>
{
  TC_HD static double gjkTol() { return 1e-10; }    // absolute gap |v| - v.w/|v| at which GJK stops [m]
  TC_HD static double gjkTouch() { return 1e-7; }   // core distance below which the pair goes to EPA [m]
  TC_HD static double dupRel() { return 1e-28; }
  TC_HD static double epaTol() { return 1e-10; }
  TC_HD static double epsLenRel() { return 1e-9; }
  TC_HD static double visRel() { return 1e-14; }
  TC_HD static double tiny() { return 1e-300; }
  TC_HD static double degRel() { return 1e-14; }
  TC_HD static double centerTiny() { return 1e-24; }
};
template <>
struct Tol<float>
{
  TC_HD static float gjkTol() { return 2e-6f; }
  TC_HD static float gjkTouch() { return 2e-5f; }
  TC_HD static float dupRel() { return 1e-12f; }
  TC_HD static float epaTol() { return 2e-6f; }
  TC_HD static float epsLenRel() { return 1e-5f; }
  TC_HD static float visRel() { return 1e-6f; }
  TC_HD static float tiny() { return 1e-30f; }
  TC_HD static float degRel() { return 1e-6f; }
  TC_HD static float centerTiny() { return 1e-12f; }
};

// support of a primitive core (not a hull) in link-frame direction d.
// Tie band: on the SIDE of a capsule / cylinder the contact normal is perpendicular to the axis, and on the side of a
// cone it makes exactly the half-angle with it, so in exact arithmetic the end-cap decision below is a tie, and in
// floating point it is decided by the last bit of d (a coin flip that moves the support point by the whole length of
// the shape; calculateContinuousData turns that into an O(1e-3) change of cc_time).  |al| within 1e-12 |d| of the tie
// is therefore taken as the tie itself, resolved one fixed way (capsule / cylinder: the +axis end, cone: the base
// rim).  The support value loses at most 1e-12 |d| x length, far below GJK's 1e-10 m stop.
template <class R>
TC_HD R supportTie();
template <>
TC_HD double supportTie<double>()
{
  return 1e-12;
}
template <>
TC_HD float supportTie<float>()
{
  return 1e-5f;
}

template <class R>
TC_HD Vec3<R> primitiveSupport(const ShapeCore<R>& s, Vec3<R> d)
{
  if (s.type == SH_SPHERE)
    return s.center;
  const R al = dot(d, s.axis);
  const R dn = norm(d);
  const R tie = supportTie<R>() * dn;
  if (s.type == SH_CAPSULE)
    return s.center + ((al >= -tie) ? s.half_len : -s.half_len) * s.axis;
  const Vec3<R> rad = d - al * s.axis;
  const R rn = norm(rad);
  if (s.type == SH_CYLINDER)
  {
    Vec3<R> p = s.center + ((al >= -tie) ? s.half_len : -s.half_len) * s.axis;
    if (rn > Tol<R>::tiny())
      p = p + (s.radius / rn) * rad;
    return p;
  }
  // cone: apex at +half_len, base disc at -half_len
  const R h = 2 * s.half_len;
  const R sin_a = s.radius / tc_sqrt(s.radius * s.radius + h * h);
  if (al > dn * sin_a + tie)
    return s.center + s.half_len * s.axis;
  Vec3<R> p = s.center - s.half_len * s.axis;
  if (rn > Tol<R>::tiny())
    p = p + (s.radius / rn) * rad;
  return p;
}

// Out-of-line pieces of the support function.  The narrow-phase kernels are bound by instruction fetch (their hot
// code is several times the 32 KB L1.5 instruction cache), so everything that is not the common case — a hull without
// a look-up table, a primitive — is a call, and the inlined part of Placed::support stays small.
#define TC_UNLIKELY(x) __builtin_expect(!!(x), 0)
#if defined(__CUDACC__)
#define TC_NI __host__ __device__ __noinline__
#else
#define TC_NI
#endif

// full scan of a hull for one direction (d0) or, swept, for the directions in both end poses in one pass over the
// vertices; first maximum in index order.  Returns b0 | b1 << 16.
template <class R>
TC_NI int hullScanFull(const Vtx<R>* verts, int n, Vec3<R> d0, Vec3<R> d1, bool cast)
{
  R best0 = lowestReal<R>(), best1 = lowestReal<R>();
  int b0 = 0, b1 = 0;
  if (!cast)
  {
    for (int i = 0; i < n; ++i)
    {
      const Vtx<R> q = verts[i];
      const R v = d0.x * q.x + d0.y * q.y + d0.z * q.z;
      if (v > best0)
      {
        best0 = v;
        b0 = i;
      }
    }
    return b0;
  }
  for (int i = 0; i < n; ++i)
  {
    const Vtx<R> q = verts[i];
    const R v0 = d0.x * q.x + d0.y * q.y + d0.z * q.z;
    const R v1 = d1.x * q.x + d1.y * q.y + d1.z * q.z;
    if (v0 > best0)
    {
      best0 = v0;
      b0 = i;
    }
    if (v1 > best1)
    {
      best1 = v1;
      b1 = i;
    }
  }
  return b0 | (b1 << 16);
}

template <class R, class P>
TC_NI Vec3<R> primitiveSupportPlaced(ShapeCore<R> sh, P p0, P p1, bool cast, Vec3<R> d)
{
  const Vec3<R> w0 = p0.apply(primitiveSupport(sh, p0.rotT(d)));
  if (!cast)
    return w0;
  const Vec3<R> w1 = p1.apply(primitiveSupport(sh, p1.rotT(d)));
  return (dot(d, w0) > dot(d, w1)) ? w0 : w1;
}

// A shape placed in the world, optionally swept between two poses (convex hull of both placements):
// support(d) = better of the supports at pose0 and pose1.
// P = how the two poses are held: Pose<R> (values, i.e. registers) or a view of poses parked in shared memory
// (tc_narrow.cu SmemPose: three poses are 72 of a GJK thread's 168 registers)
template <class R, class P = Pose<R>>
struct Placed
{
  ShapeCore<R> sh;
  P p0, p1;
  bool cast;

  // `scanned` counts the vertices of the hull per queried pose (the work of the plain scan the reference's
  // convex-hull support does), whether or not the look-up table let most of them be skipped
  TC_HD Vec3<R> support(Vec3<R> d, ScanCount& scanned) const
  {
    if (TC_UNLIKELY(sh.type != SH_HULL))
      return primitiveSupportPlaced(sh, p0, p1, cast, d);
    const Vec3<R> d0 = p0.rotT(d);
    const Vec3<R> d1 = cast ? p1.rotT(d) : d0;
    int b0 = 0, b1 = 0, ncmp = 0;
    bool done = false;
    if (sh.lut)
    {
      // look-up table path: only the vertices that can be extreme for directions of d's cube-map cell are compared
      const uint8_t* cand = reinterpret_cast<const uint8_t*>(sh.lut + (LUT_CELLS + 1));
      const int c0 = lutCellF(static_cast<float>(d0.x), static_cast<float>(d0.y), static_cast<float>(d0.z));
      const int c1 = cast ? lutCellF(static_cast<float>(d1.x), static_cast<float>(d1.y), static_cast<float>(d1.z)) : c0;
      if (c0 >= 0 && c1 >= 0)
      {
        R best = lowestReal<R>();
        const int k0 = sh.lut[c0], e0 = sh.lut[c0 + 1];
        ncmp = e0 - k0;
#pragma unroll 1
        for (int k = k0; k < e0; ++k)
        {
          const int i = cand[k];
          const Vtx<R> q = sh.verts[i];
          const R val = d0.x * q.x + d0.y * q.y + d0.z * q.z;
          if (val > best)
          {
            best = val;
            b0 = i;
          }
        }
        if (cast)
        {
          best = lowestReal<R>();
          const int k1 = sh.lut[c1], e1 = sh.lut[c1 + 1];
          ncmp += e1 - k1;
#pragma unroll 1
          for (int k = k1; k < e1; ++k)
          {
            const int i = cand[k];
            const Vtx<R> q = sh.verts[i];
            const R val = d1.x * q.x + d1.y * q.y + d1.z * q.z;
            if (val > best)
            {
              best = val;
              b1 = i;
            }
          }
        }
        done = true;
      }
    }
    if (TC_UNLIKELY(!done))
    {
      const int packed = hullScanFull(sh.verts, sh.nverts, d0, d1, cast);
      b0 = packed & 65535;
      b1 = packed >> 16;
      ncmp = cast ? 2 * sh.nverts : sh.nverts;
    }
    const Vtx<R> q0 = sh.verts[b0];
    const Vec3<R> w0 = p0.apply(Vec3<R>(q0.x, q0.y, q0.z));
    if (!cast)
    {
      countScan(scanned, sh.nverts, ncmp);
      return w0;
    }
    countScan(scanned, 2 * sh.nverts, ncmp);
    const Vtx<R> q1 = sh.verts[b1];
    const Vec3<R> w1 = p1.apply(Vec3<R>(q1.x, q1.y, q1.z));
    return (dot(d, w0) > dot(d, w1)) ? w0 : w1;
  }
  TC_HD Vec3<R> center() const
  {
    const Vec3<R> c0 = p0.apply(sh.center);
    return cast ? R(0.5) * (c0 + p1.apply(sh.center)) : c0;
  }
};

// ---------------------------------------------------------------------------------------------
// simplex in registers: points 0..n-1, each the Minkowski-difference point w and its witness a on A
// (the witness on B is a - w)
// ---------------------------------------------------------------------------------------------
template <class R>
struct Simplex
{
  int n;
  Vec3<R> w[4], a[4];
  R lam[4];
};

// closest point of triangle (a,b,c) to the origin, barycentric weights out
template <class R>
TC_HD Vec3<R> closestTriangle(Vec3<R> a, Vec3<R> b, Vec3<R> c, R& la, R& lb, R& lc)
{
  const Vec3<R> ab = b - a, ac = c - a;
  const R d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= 0 && d2 <= 0)
  {
    la = 1, lb = 0, lc = 0;
    return a;
  }
  const R d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= 0 && d4 <= d3)
  {
    la = 0, lb = 1, lc = 0;
    return b;
  }
  const R vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0)
  {
    const R t = d1 / (d1 - d3);
    la = 1 - t, lb = t, lc = 0;
    return a + t * ab;
  }
  const R d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= 0 && d5 <= d6)
  {
    la = 0, lb = 0, lc = 1;
    return c;
  }
  const R vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0)
  {
    const R t = d2 / (d2 - d6);
    la = 1 - t, lb = 0, lc = t;
    return a + t * ac;
  }
  const R va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0)
  {
    const R t = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    la = 0, lb = 1 - t, lc = t;
    return b + t * (c - b);
  }
  const R den = R(1) / (va + vb + vc);
  lb = vb * den;
  lc = vc * den;
  la = 1 - lb - lc;
  return a + lb * ab + lc * ac;
}

// Closest point to the origin of the simplex spanned by points 0..m-1 of s (m = 2..4).
// Outputs the barycentric weights aligned with the slots (0 for dropped points) and the keep mask.
// mask == 15 means the origin is inside the tetrahedron.
template <class R>
TC_HD Vec3<R> closestOnSimplex(const Simplex<R>& s, int m, R lam[4], int& mask)
{
  lam[0] = lam[1] = lam[2] = lam[3] = 0;
  if (m == 2)
  {
    const Vec3<R> ab = s.w[1] - s.w[0];
    const R den = norm2(ab);
    R t = (den > 0) ? -dot(s.w[0], ab) / den : R(0);
    t = tc_min(R(1), tc_max(R(0), t));
    lam[0] = 1 - t;
    lam[1] = t;
    mask = ((lam[0] > 0) ? 1 : 0) | ((lam[1] > 0) ? 2 : 0);
    return s.w[0] + t * ab;
  }
  // Triangle (m == 3) and tetrahedron (m == 4) share one instance of the triangle case analysis (code size): the
  // tetrahedron visits its four faces (0,1,2 | 3), (0,1,3 | 2), (0,2,3 | 1), (1,2,3 | 0) in this order and keeps the
  // closest point among the faces whose plane separates the origin from the opposite vertex; the triangle is the
  // single face (0,1,2) taken unconditionally.
  R best = -lowestReal<R>();
  Vec3<R> bestp;
  int bf = -1;
  R bl0 = 0, bl1 = 0, bl2 = 0;
  const int nfaces = (m == 4) ? 4 : 1;
#pragma unroll 1
  for (int f = 0; f < nfaces; ++f)
  {
    const Vec3<R> a = (f < 3) ? s.w[0] : s.w[1];
    const Vec3<R> b = (f < 2) ? s.w[1] : s.w[2];
    const Vec3<R> c = (f < 1) ? s.w[2] : s.w[3];
    bool take = true;
    if (m == 4)
    {
      const Vec3<R> d = (f == 0) ? s.w[3] : ((f == 1) ? s.w[2] : ((f == 2) ? s.w[1] : s.w[0]));
      const Vec3<R> n = cross(b - a, c - a);
      const R sp = -dot(a, n);
      const R sd = dot(d - a, n);
      const bool degenerate = tc_abs(sd) <= Tol<R>::degRel() * tc_sqrt(norm2(n) * norm2(d - a) + Tol<R>::tiny());
      take = sp * sd < 0 || degenerate;
    }
    if (take)
    {
      R la, lb, lc;
      const Vec3<R> p = closestTriangle(a, b, c, la, lb, lc);
      const R d2 = norm2(p);
      if (d2 < best)
      {
        best = d2;
        bestp = p;
        bf = f;
        bl0 = la, bl1 = lb, bl2 = lc;
      }
    }
  }
  if (bf < 0)
  {
    lam[0] = lam[1] = lam[2] = lam[3] = R(0.25);
    mask = 15;
    return Vec3<R>();
  }
  // scatter the three face weights back to their slots
  if (bf == 0)
    lam[0] = bl0, lam[1] = bl1, lam[2] = bl2;
  else if (bf == 1)
    lam[0] = bl0, lam[1] = bl1, lam[3] = bl2;
  else if (bf == 2)
    lam[0] = bl0, lam[2] = bl1, lam[3] = bl2;
  else
    lam[1] = bl0, lam[2] = bl1, lam[3] = bl2;
  mask = ((lam[0] > 0) ? 1 : 0) | ((lam[1] > 0) ? 2 : 0) | ((lam[2] > 0) ? 4 : 0) | ((lam[3] > 0) ? 8 : 0);
  return bestp;
}

// keep the points selected by `mask` (order preserved), select-only so the simplex stays in registers
template <class R>
TC_HD void compactSimplex(Simplex<R>& s, const R lam[4], int mask)
{
  const bool m0 = mask & 1, m1 = mask & 2, m2 = mask & 4, m3 = mask & 8;
  // source slot of output slot k = k-th set bit
  const int c0 = m0 ? 1 : 0, c1 = c0 + (m1 ? 1 : 0), c2 = c1 + (m2 ? 1 : 0);
  // out0: first kept
  const int s0 = m0 ? 0 : (m1 ? 1 : (m2 ? 2 : 3));
  // out1: second kept
  const int s1 = (m1 && c0 == 1) ? 1 : ((m2 && c1 == 1) ? 2 : 3);
  // out2: third kept
  const int s2 = (m2 && c1 == 2) ? 2 : 3;
  (void)c2;
  (void)m3;
#define TC_SEL(arr, idx) ((idx) == 0 ? arr[0] : ((idx) == 1 ? arr[1] : ((idx) == 2 ? arr[2] : arr[3])))
#define TC_SELV(arr, idx)                                                                                            \
  Vec3<R>(TC_SEL(arr, idx).x, TC_SEL(arr, idx).y, TC_SEL(arr, idx).z)
  const Vec3<R> w0 = TC_SELV(s.w, s0), w1 = TC_SELV(s.w, s1), w2 = TC_SELV(s.w, s2);
  const Vec3<R> a0 = TC_SELV(s.a, s0), a1 = TC_SELV(s.a, s1), a2 = TC_SELV(s.a, s2);
  const R l0 = TC_SEL(lam, s0), l1 = TC_SEL(lam, s1), l2 = TC_SEL(lam, s2);
#undef TC_SELV
#undef TC_SEL
  s.w[0] = w0, s.w[1] = w1, s.w[2] = w2;
  s.a[0] = a0, s.a[1] = a1, s.a[2] = a2;
  s.lam[0] = l0, s.lam[1] = l1, s.lam[2] = l2;
  s.lam[3] = lam[3];
  s.n = (m0 ? 1 : 0) + (m1 ? 1 : 0) + (m2 ? 1 : 0) + (m3 ? 1 : 0);
}

enum
{
  GJK_MAX_ITERS = 128,
  EPA_MAX_ITERS = 255
};

template <class R>
struct GjkOut
{
  bool separated_beyond, intersect, converged;
  bool certified;  // the loop ended on a test that bounds the error (gap, separation, enclosure); see gjk()
  R dist;
  Vec3<R> pa, pb;
  int iters;
};

template <class R>
TC_HD void putPoint(Simplex<R>& s, int k, Vec3<R> w, Vec3<R> a)
{
  // predicated writes keep the arrays in registers
  if (k == 0)
    s.w[0] = w, s.a[0] = a;
  if (k == 1)
    s.w[1] = w, s.a[1] = a;
  if (k == 2)
    s.w[2] = w, s.a[2] = a;
  if (k == 3)
    s.w[3] = w, s.a[3] = a;
}

template <class R, class PLA, class PLB>
TC_HD GjkOut<R> gjk(const PLA& A, const PLB& B, R threshold_core, Simplex<R>& S, ScanCount& scanned)
{
  GjkOut<R> r;
  r.separated_beyond = false;
  r.intersect = false;
  r.converged = true;
  r.certified = true;
  r.dist = 0;
  Vec3<R> v = A.center() - B.center();
  if (norm2(v) < Tol<R>::centerTiny())
    v = Vec3<R>(1, 0, 0);
  S.n = 0;
  S.lam[0] = 1, S.lam[1] = 0, S.lam[2] = 0, S.lam[3] = 0;
  R vv = 0;
  const R touch2 = Tol<R>::gjkTouch() * Tol<R>::gjkTouch();
  // trip 0 places the first simplex point (support along the centre difference); it shares the one support site
  // of the loop so that the support code exists once in the kernel
  for (r.iters = 0; r.iters <= GJK_MAX_ITERS; ++r.iters)
  {
    if (r.iters > 0 && vv <= touch2)
    {
      r.intersect = true;
      break;
    }
    const Vec3<R> a = A.support(-v, scanned), b = B.support(v, scanned);
    const Vec3<R> w = a - b;
    if (r.iters == 0)
    {
      S.w[0] = w;
      S.a[0] = a;
      S.n = 1;
      v = w;
      vv = norm2(v);
      continue;
    }
    const R vw = dot(v, w);
    if (vw > 0 && vw * vw > threshold_core * threshold_core * vv)
    {
      r.separated_beyond = true;
      return r;
    }
    if (vv - vw <= Tol<R>::gjkTol() * tc_sqrt(vv))
      break;
    bool dup = false;
    const R dupt = Tol<R>::dupRel() * tc_max(R(1), norm2(w));
    if (norm2(w - S.w[0]) <= dupt)
      dup = true;
    if (S.n > 1 && norm2(w - S.w[1]) <= dupt)
      dup = true;
    if (S.n > 2 && norm2(w - S.w[2]) <= dupt)
      dup = true;
    // The two exits below (support point seen before; no progress within rounding) end GJK on the closest point found
    // so far WITHOUT a bound on its error.  In FP64 they fire within ~1e-10 m of the gap test; in FP32 a sliver simplex
    // can trigger them centimetres early, so a float run that leaves through them with a gap above 5e-5 m is not
    // reported: the item is re-run in FP64 (distanceGjk)
    const bool open_gap = sizeof(R) == sizeof(float) && vv - vw > R(5e-5) * tc_sqrt(vv);
    if (dup)
    {
      r.certified = !open_gap;
      break;
    }
    putPoint(S, S.n, w, a);  // slots >= n are free: S itself is untouched until the commit below
    R lam[4];
    int mask;
    const Vec3<R> vn = closestOnSimplex(S, S.n + 1, lam, mask);
    if (mask == 15)
    {
      S.n = 4;
      S.lam[0] = S.lam[1] = S.lam[2] = S.lam[3] = R(0.25);
      r.intersect = true;
      break;
    }
    const R vvn = norm2(vn);
    if (vvn >= vv)
    {
      r.certified = !open_gap;
      break;  // no progress within rounding: S already holds the closest point
    }
    compactSimplex(S, lam, mask);
    v = vn;
    vv = vvn;
    if (r.iters == GJK_MAX_ITERS)
      r.converged = false, r.certified = false;
  }
  if (!r.intersect)
  {
    r.dist = tc_sqrt(vv);
    Vec3<R> pa, pw;
    pa = S.lam[0] * S.a[0];
    pw = S.lam[0] * S.w[0];
    if (S.n > 1)
      pa = pa + S.lam[1] * S.a[1], pw = pw + S.lam[1] * S.w[1];
    if (S.n > 2)
      pa = pa + S.lam[2] * S.a[2], pw = pw + S.lam[2] * S.w[2];
    r.pa = pa;
    r.pb = pa - pw;
  }
  return r;
}

// ---------------------------------------------------------------------------------------------
// EPA on a bounded polytope store (thread-local on the fast path, global scratch on the rare
// deep-expansion path; same code).  Dead face slots are reused first-fit.
// ---------------------------------------------------------------------------------------------
template <class R>
struct EpaStore
{
  Vec3<R>* pw;  // [max_verts]
  Vec3<R>* pa;
  Vec3<R>* fn;  // [max_faces]
  R* fd;
  int* fv;      // packed v0 | v1 << 10 | v2 << 20 ; -1 = dead slot
  int* edges;   // [max_faces] packed a | b << 16
  int max_verts, max_faces;
  unsigned long long* erow;  // [max_verts] directed-edge bit matrix of the register-mask path (all zero between trips)
};

template <class R>
struct EpaOut
{
  bool ok;         // false: the store overflowed, nothing else valid
  bool converged;
  R depth;
  Vec3<R> normal, pa, pb;
  int iters;
};

template <class R>
TC_HD Vec3<R> leastAxisCross(Vec3<R> e)
{
  Vec3<R> ax(1, 0, 0);
  if (tc_abs(e.y) <= tc_abs(e.x) && tc_abs(e.y) <= tc_abs(e.z))
    ax = Vec3<R>(0, 1, 0);
  else if (tc_abs(e.z) <= tc_abs(e.x) && tc_abs(e.z) <= tc_abs(e.y))
    ax = Vec3<R>(0, 0, 1);
  return cross(e, ax);
}

// resumable EPA: epaInit (tetrahedron + first faces), epaTrip (one support query + one polytope expansion),
// epaFinish (depth, normal, witness points)
template <class R>
struct EpaState
{
  int np, nf, best, iters;
  unsigned long long alive;  // live-face mask (MASK mode)
  R bd, eps_len, vis_eps;
  Vec3<R> centroid;
  // stall exit (curved cores, and every pair in FP32 mode): running minimum of the support values, the narrowest
  // bracket [closest face, that minimum] seen so far and the trips since it last narrowed by 10 %
  R ub_best, width_best;
  int stall;
  bool stall_exit;
};

// Stall exit.  EPA brackets the penetration depth between the closest face of the polytope (bd, never decreasing) and
// the smallest support value found along a face normal (never increasing); it is done when the bracket closes to
// epaTol.  On a polytope that happens after finitely many trips.  On a CURVED core (cylinder, cone) every new vertex
// lies on the surface next to the previous ones, the faces become slivers whose plane distance carries ~1e-10 m of
// rounding, and a few items then dither at the tolerance for a hundred trips without narrowing the bracket; in FP32 the
// same happens on every shape at ~1e-6.  EPA_STALL_TRIPS trips without narrowing the bracket by 10 % therefore end the
// expansion: in FP64 on the current closest face (the CPU checker applies the same rule to curved cores), in FP32 by
// handing the item to the FP64 re-run kernel like one whose polytope overflowed — float EPA only ever reports results it
// has closed to its tolerance.  Hull / sphere / capsule pairs in FP64 never take this exit: the gap test alone decides.
enum
{
  EPA_STALL_TRIPS = 6,       // FP64, curved cores
  EPA_STALL_TRIPS_F32 = 16   // FP32, every pair: polytope pairs legitimately spend a few trips on coplanar faces
};
template <class R>
TC_HD bool epaStalled(EpaState<R>& E, R support_value)
{
  if (!E.stall_exit)
    return false;
  const int patience = sizeof(R) == sizeof(float) ? EPA_STALL_TRIPS_F32 : EPA_STALL_TRIPS;
  if (support_value < E.ub_best)
    E.ub_best = support_value;
  const R width = E.ub_best - E.bd;
  if (width < R(0.9) * E.width_best)
  {
    E.width_best = width;
    E.stall = 0;
    return false;
  }
  return ++E.stall >= patience;
}

template <class R, bool MASK>
TC_HD bool epaIsAlive(const EpaState<R>& E, const EpaStore<R>& st, int f)
{
  return MASK ? ((E.alive >> f) & 1ull) != 0 : st.fv[f] >= 0;
}

// adds a face in the first dead slot (else appends); returns the slot or -1 on overflow; d_out = plane distance.
// pi, pj, pk are the face's points (callers that already hold them in registers pass them instead of re-reading the
// store: the apex of every face of an expansion is the new support point)
template <class R, bool MASK>
TC_HD int epaAddFacePts(EpaState<R>& E, EpaStore<R>& st, int i, int j, int k, Vec3<R> pi, Vec3<R> pj, Vec3<R> pk,
                        bool orient, R& d_out)
{
  Vec3<R> n = cross(pj - pi, pk - pi);
  if (orient && dot(n, pi - E.centroid) < 0)
  {
    const int t = j;
    j = k;
    k = t;
    n = -n;
  }
  int slot = -1;
  if (MASK)
  {
    const unsigned long long dead = ~E.alive & ((E.nf >= 64) ? ~0ull : ((1ull << E.nf) - 1ull));
    if (dead)
    {
#if defined(__CUDA_ARCH__)
      slot = __ffsll(static_cast<long long>(dead)) - 1;
#else
      unsigned long long x = dead;
      slot = 0;
      while (!(x & 1ull))
      {
        x >>= 1;
        ++slot;
      }
#endif
    }
  }
  else
  {
    for (int f = 0; f < E.nf; ++f)
      if (st.fv[f] < 0)
      {
        slot = f;
        break;
      }
  }
  if (slot < 0)
  {
    if (E.nf >= st.max_faces)
      return -1;
    slot = E.nf++;
  }
  const R nn = norm(n);
  st.fv[slot] = i | (j << 10) | (k << 20);
  if (MASK)
    E.alive |= 1ull << slot;
  if (nn > Tol<R>::tiny())
  {
    const Vec3<R> un = (R(1) / nn) * n;
    st.fn[slot] = un;
    d_out = dot(un, pi);
  }
  else
  {
    st.fn[slot] = Vec3<R>(0, 0, 0);
    d_out = -lowestReal<R>();  // degenerate sliver: never selected
  }
  st.fd[slot] = d_out;
  return slot;
}

template <class R, bool MASK>
TC_HD int epaAddFace(EpaState<R>& E, EpaStore<R>& st, int i, int j, int k, bool orient, R& d_out)
{
  return epaAddFacePts<R, MASK>(E, st, i, j, k, st.pw[i], st.pw[j], st.pw[k], orient, d_out);
}

// smallest plane distance, lowest slot on ties (= first minimum of a scan in slot order)
template <class R>
TC_HD void epaOffer(EpaState<R>& E, int f, R d)
{
  if (E.best < 0 || d < E.bd || (d == E.bd && f < E.best))
  {
    E.best = f;
    E.bd = d;
  }
}

// returns 0: polytope ready, run trips; 1: finished already (flat Minkowski difference), r is final
template <class R, bool MASK, class PLA, class PLB>
TC_HDN int epaInit(const PLA& A, const PLB& B, const Simplex<R>& start, R scale, EpaStore<R>& st,
                   EpaState<R>& E, EpaOut<R>& r, ScanCount& scanned)
{
  r.ok = true;
  r.converged = true;
  r.depth = 0;
  r.iters = 0;
  E.np = start.n;
  for (int i = 0; i < 4; ++i)
    if (i < start.n)
    {
      st.pw[i] = start.w[i];
      st.pa[i] = start.a[i];
    }
  E.eps_len = Tol<R>::epsLenRel() * tc_max(scale, R(1e-3));
  E.vis_eps = Tol<R>::visRel() * tc_max(scale, R(1e-3));

  // --- grow the GJK simplex to a tetrahedron with non-zero volume ---
  if (E.np == 4)
  {
    const R vol = dot(cross(st.pw[1] - st.pw[0], st.pw[2] - st.pw[0]), st.pw[3] - st.pw[0]);
    if (tc_abs(vol) <= E.eps_len * E.eps_len * E.eps_len)
      E.np = 3;
  }
  // Probing for the missing points, one support site for all three stages (code size): a point (np == 1) tries the
  // six axis directions for a second point, a segment (np == 2) four directions across it for a third, a triangle
  // (np == 3) its two normals for a fourth.  A stage that finds nothing ends the probing (flat difference).
  for (int k = 0; E.np < 4;)
  {
    const int lim = (E.np == 1) ? 6 : ((E.np == 2) ? 4 : 2);
    if (k >= lim)
      break;
    Vec3<R> d, e, n;
    R en = 0, area = 0;
    if (E.np == 1)
    {
      const R sg = (k & 1) ? R(-1) : R(1);
      d = Vec3<R>((k >> 1) == 0 ? sg : R(0), (k >> 1) == 1 ? sg : R(0), (k >> 1) == 2 ? sg : R(0));
    }
    else if (E.np == 2)
    {
      e = st.pw[1] - st.pw[0];
      const Vec3<R> d0 = leastAxisCross(e);
      const Vec3<R> d1 = cross(e, d0);
      en = norm(e);
      d = (k == 0) ? d0 : ((k == 1) ? -d0 : ((k == 2) ? d1 : -d1));
    }
    else
    {
      n = cross(st.pw[1] - st.pw[0], st.pw[2] - st.pw[0]);
      area = norm(n);
      d = (k == 0) ? n : -n;
    }
    const Vec3<R> a = A.support(d, scanned), b = B.support(-d, scanned);
    const Vec3<R> w = a - b;
    bool accept;
    if (E.np == 1)
      accept = norm(w - st.pw[0]) > E.eps_len;
    else if (E.np == 2)
      accept = norm(cross(w - st.pw[0], e)) > E.eps_len * en;
    else
      accept = tc_abs(dot(w - st.pw[0], n)) > E.eps_len * area;
    ++k;
    if (accept)
    {
      st.pw[E.np] = w;
      st.pa[E.np] = a;
      E.np += 1;
      k = 0;
    }
  }
  if (E.np < 4)
  {
    // flat Minkowski difference (point / segment / polygon cores): core depth is zero, the normal is
    // the direction in which the difference has no extent
    Vec3<R> n(0, 0, 1);
    if (E.np == 3)
      n = cross(st.pw[1] - st.pw[0], st.pw[2] - st.pw[0]);
    else if (E.np == 2)
      n = leastAxisCross(st.pw[1] - st.pw[0]);
    const R nn = norm(n);
    r.normal = (nn > 0) ? (R(1) / nn) * n : Vec3<R>(0, 0, 1);
    Vec3<R> pa, pw;
    if (start.n == 1)
    {
      pa = start.a[0];
      pw = start.w[0];
    }
    else
    {
      R lam[4];
      int mask;
      closestOnSimplex(start, start.n, lam, mask);
      pa = lam[0] * start.a[0] + lam[1] * start.a[1];
      pw = lam[0] * start.w[0] + lam[1] * start.w[1];
      if (start.n > 2)
        pa = pa + lam[2] * start.a[2], pw = pw + lam[2] * start.w[2];
      if (start.n > 3)
        pa = pa + lam[3] * start.a[3], pw = pw + lam[3] * start.w[3];
    }
    r.pa = pa;
    r.pb = pa - pw;
    return 1;
  }

  E.nf = 0;
  E.alive = 0;
  E.best = -1;
  E.bd = 0;
  E.iters = 1;
  E.ub_best = -lowestReal<R>();
  E.width_best = -lowestReal<R>();
  E.stall = 0;
  E.stall_exit = sizeof(R) == sizeof(float) || A.sh.type >= SH_CYLINDER || B.sh.type >= SH_CYLINDER;
  E.centroid = R(0.25) * (st.pw[0] + st.pw[1] + st.pw[2] + st.pw[3]);
  R d;
  int f = epaAddFace<R, MASK>(E, st, 0, 1, 2, true, d);
  epaOffer(E, f, d);
  f = epaAddFace<R, MASK>(E, st, 0, 1, 3, true, d);
  epaOffer(E, f, d);
  f = epaAddFace<R, MASK>(E, st, 0, 2, 3, true, d);
  epaOffer(E, f, d);
  f = epaAddFace<R, MASK>(E, st, 1, 2, 3, true, d);
  epaOffer(E, f, d);
  return 0;
}

enum
{
  EPA_CONTINUE = 0,
  EPA_DONE = 1,
  EPA_UNCONVERGED = 2,
  EPA_OVERFLOW = 3
};

TC_HD int lowestBit(unsigned long long x)
{
#if defined(__CUDA_ARCH__)
  return __ffsll(static_cast<long long>(x)) - 1;
#else
  return __builtin_ffsll(static_cast<long long>(x)) - 1;
#endif
}

// One expansion of the register-mask polytope (fast path, <= 64 faces).  The face loop of the general form below
// runs the horizon bookkeeping inside the per-face loop, so a warp executes it whenever ANY lane sees a visible face
// at that slot; here the work is split into passes whose trip counts are what each lane needs:
//   1. every live face: visibility bit + closest surviving face            (uniform body, no branches taken)
//   2. visible faces only: toggle their directed edges in a per-vertex bit matrix (an interior edge meets its twin
//      and cancels), 3. visible faces again, slot order: the edges still set are the horizon, in the order the
//      list-based form produces them, 4. one new face per horizon edge.
// `erow` is all-zero on entry and on exit.
#if defined(TC_NARROW_HIST) && defined(__CUDACC__)
static __device__ unsigned long long g_scan_dbg[4];  // faces scanned, faces a norm bound would have skipped, faces visible
#endif

template <class R, class PLA, class PLB>
TC_HD int epaTripMasked(const PLA& A, const PLB& B, EpaStore<R>& st, EpaState<R>& E,
                        unsigned long long* erow, ScanCount& scanned)
{
  const R DEGENERATE = -lowestReal<R>();
  const Vec3<R> bn = st.fn[E.best];
  const Vec3<R> xa = A.support(bn, scanned), xb = B.support(-bn, scanned);
  const Vec3<R> xw = xa - xb;
  const R sv = dot(bn, xw);
  if (sv - E.bd <= Tol<R>::epaTol())
    return EPA_DONE;
  if (epaStalled(E, sv))
    return sizeof(R) == sizeof(float) ? EPA_OVERFLOW : EPA_DONE;  // FP32: the item is re-run in FP64 (deep list)
  if (E.iters == EPA_MAX_ITERS)
    return EPA_UNCONVERGED;
  if (E.np >= st.max_verts)
    return EPA_OVERFLOW;
  int sbest = -1;
  R sbd = 0;
  unsigned long long vis = 0;
  // four faces per trip: their planes are fetched from the thread-local store together, so the loads overlap
  // instead of each face waiting for its own; the decisions are then applied in slot order
  for (unsigned long long m = E.alive; m;)
  {
    int f[4];
    bool ok[4];
    R fd[4];
    Vec3<R> fn[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
      ok[k] = m != 0;
      f[k] = ok[k] ? lowestBit(m) : f[0];
      m &= m - 1;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
      fd[k] = st.fd[f[k]];
      fn[k] = st.fn[f[k]];
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
      const bool visible = (fd[k] != DEGENERATE) && (dot(fn[k], xw) - fd[k] > E.vis_eps);
#if defined(TC_NARROW_HIST) && defined(__CUDA_ARCH__)
      if (ok[k])
      {
        atomicAdd(&g_scan_dbg[0], 1ull);
        if (fd[k] > 0 && norm2(xw) < fd[k] * fd[k] * (R(1) - R(1e-12)))
          atomicAdd(&g_scan_dbg[1], 1ull);
        if (visible)
          atomicAdd(&g_scan_dbg[2], 1ull);
      }
#endif
      if (ok[k])
      {
        if (visible)
          vis |= 1ull << f[k];
        else if (sbest < 0 || fd[k] < sbd)
        {
          sbest = f[k];
          sbd = fd[k];
        }
      }
    }
  }
  if (vis == 0)  // numerically x is not outside any face although the gap test said so: accept the face (FP32: re-run)
    return sizeof(R) == sizeof(float) ? EPA_OVERFLOW : EPA_DONE;
  for (unsigned long long m = vis; m; m &= m - 1)
  {
    const int fv = st.fv[lowestBit(m)];
    const int v0 = fv & 1023, v1 = (fv >> 10) & 1023, v2 = (fv >> 20) & 1023;
#pragma unroll
    for (int e = 0; e < 3; ++e)
    {
      const int ea = (e == 0) ? v0 : ((e == 1) ? v1 : v2);
      const int eb = (e == 0) ? v1 : ((e == 1) ? v2 : v0);
      const unsigned long long twin = erow[eb];
      if ((twin >> ea) & 1ull)
        erow[eb] = twin & ~(1ull << ea);
      else
        erow[ea] |= 1ull << eb;
    }
  }
  int ne = 0;
  bool overflow = false;
  for (unsigned long long m = vis; m; m &= m - 1)
  {
    const int fv = st.fv[lowestBit(m)];
    const int v0 = fv & 1023, v1 = (fv >> 10) & 1023, v2 = (fv >> 20) & 1023;
#pragma unroll
    for (int e = 0; e < 3; ++e)
    {
      const int ea = (e == 0) ? v0 : ((e == 1) ? v1 : v2);
      const int eb = (e == 0) ? v1 : ((e == 1) ? v2 : v0);
      const unsigned long long row = erow[ea];
      if ((row >> eb) & 1ull)
      {
        erow[ea] = row & ~(1ull << eb);
        if (ne < st.max_faces)
          st.edges[ne++] = ea | (eb << 16);
        else
          overflow = true;
      }
    }
  }
  if (overflow)
    return EPA_OVERFLOW;
  if (ne == 0)
    return sizeof(R) == sizeof(float) ? EPA_OVERFLOW : EPA_DONE;
  E.alive &= ~vis;
  const int xi = E.np;
  st.pw[xi] = xw;
  st.pa[xi] = xa;
  E.np += 1;
  E.best = sbest;
  E.bd = sbd;
  // the apex of every new face is the support point just found: passed in registers, not re-read from the store
  for (int q = 0; q < ne; ++q)
  {
    R d;
    const int e = st.edges[q];
    const int a = e & 65535, b = e >> 16;
    const int f = epaAddFacePts<R, true>(E, st, a, b, xi, st.pw[a], st.pw[b], xw, false, d);
    if (f < 0)
      return EPA_OVERFLOW;
    epaOffer(E, f, d);
  }
  E.iters += 1;
  return EPA_CONTINUE;
}

template <class R, bool MASK, class PLA, class PLB>
TC_HDN int epaTrip(const PLA& A, const PLB& B, EpaStore<R>& st, EpaState<R>& E, ScanCount& scanned)
{
  const R DEGENERATE = -lowestReal<R>();
  const Vec3<R> bn = st.fn[E.best];
  const Vec3<R> xa = A.support(bn, scanned), xb = B.support(-bn, scanned);
  const Vec3<R> xw = xa - xb;
  const R sv = dot(bn, xw);
  if (sv - E.bd <= Tol<R>::epaTol())
    return EPA_DONE;
  if (epaStalled(E, sv))
    return sizeof(R) == sizeof(float) ? EPA_OVERFLOW : EPA_DONE;  // FP32: the item is re-run in FP64 (deep list)
  if (E.iters == EPA_MAX_ITERS)
    return EPA_UNCONVERGED;
  if (E.np >= st.max_verts)
    return EPA_OVERFLOW;
  // one pass over the live faces: kill those that see x, collect the horizon, and track the closest
  // survivor so that no separate "find the closest face" scan is needed
  int ne = 0, sbest = -1;
  R sbd = 0;
  unsigned long long killed = 0;
  for (int f = 0; f < E.nf; ++f)
  {
    if (!epaIsAlive<R, MASK>(E, st, f))
      continue;
    const R fd = st.fd[f];
    bool visible = false;
    if (fd != DEGENERATE)
      visible = dot(st.fn[f], xw) - fd > E.vis_eps;
    if (!visible)
    {
      if (sbest < 0 || fd < sbd)
      {
        sbest = f;
        sbd = fd;
      }
      continue;
    }
    const int fv = st.fv[f];
    const int v0 = fv & 1023, v1 = (fv >> 10) & 1023, v2 = (fv >> 20) & 1023;
    if (MASK)
      killed |= 1ull << f;
    else
      st.fv[f] = -1;
    for (int e = 0; e < 3; ++e)
    {
      const int ea = (e == 0) ? v0 : ((e == 1) ? v1 : v2);
      const int eb = (e == 0) ? v1 : ((e == 1) ? v2 : v0);
      const int twin = eb | (ea << 16);
      int found = -1;
      for (int q = 0; q < ne; ++q)
        if (st.edges[q] == twin)
        {
          found = q;
          break;
        }
      if (found >= 0)
      {
        for (int q = found; q + 1 < ne; ++q)
          st.edges[q] = st.edges[q + 1];
        --ne;
      }
      else
      {
        if (ne >= st.max_faces)
          return EPA_OVERFLOW;
        st.edges[ne++] = ea | (eb << 16);
      }
    }
  }
  if (ne == 0)
    return EPA_DONE;  // numerically x is not outside any face although the gap test said so: accept the face
  E.alive &= ~killed;
  const int xi = E.np;
  st.pw[xi] = xw;
  st.pa[xi] = xa;
  E.np += 1;
  E.best = sbest;
  E.bd = sbd;
  for (int q = 0; q < ne; ++q)
  {
    R d;
    const int e = st.edges[q];
    const int f = epaAddFace<R, MASK>(E, st, e & 65535, e >> 16, xi, false, d);
    if (f < 0)
      return EPA_OVERFLOW;
    epaOffer(E, f, d);
  }
  E.iters += 1;
  return EPA_CONTINUE;
}

template <class R, bool MASK>
TC_HDN void epaFinish(const EpaStore<R>& st, const EpaState<R>& E, EpaOut<R>& r)
{
  const R DEGENERATE = -lowestReal<R>();
  r.iters = E.iters;
  r.depth = E.bd;
  r.normal = st.fn[E.best];
  // Witness points: the face with the smallest PLANE distance may be one of several coplanar triangles of
  // a larger Minkowski face; the witness is taken on the triangle that is actually closest to the origin.
  int wf = E.best;
  R wbest = -lowestReal<R>(), la = 0, lb = 0, lc = 0;
  unsigned long long cand = 0;  // fast path: candidate faces are collected first, then visited
  if (MASK)
  {
    for (unsigned long long m = E.alive; m;)
    {
      int f[4];
      bool ok[4];
      R fd[4];
#pragma unroll
      for (int k = 0; k < 4; ++k)
      {
        ok[k] = m != 0;
        f[k] = ok[k] ? lowestBit(m) : f[0];
        m &= m - 1;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        fd[k] = st.fd[f[k]];
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (ok[k] && fd[k] != DEGENERATE && !(fd[k] > r.depth + E.eps_len))
          cand |= 1ull << f[k];
    }
  }
  auto visit = [&](int f) {
    const int gv = st.fv[f];
    R a0, a1, a2;
    const Vec3<R> q = closestTriangle(st.pw[gv & 1023], st.pw[(gv >> 10) & 1023], st.pw[(gv >> 20) & 1023], a0, a1, a2);
    const R q2 = norm2(q);
    if (q2 < wbest)
    {
      wbest = q2;
      wf = f;
      la = a0, lb = a1, lc = a2;
    }
  };
  if (MASK)
  {
    for (; cand; cand &= cand - 1)
      visit(lowestBit(cand));
  }
  else
  {
    for (int f = 0; f < E.nf; ++f)
    {
      if (!epaIsAlive<R, MASK>(E, st, f))
        continue;
      const R fd = st.fd[f];
      if (fd == DEGENERATE || fd > r.depth + E.eps_len)
        continue;
      visit(f);
    }
  }
  const int gv = st.fv[wf];
  const int v0 = gv & 1023, v1 = (gv >> 10) & 1023, v2 = (gv >> 20) & 1023;
  r.pa = la * st.pa[v0] + lb * st.pa[v1] + lc * st.pa[v2];
  const Vec3<R> pw = la * st.pw[v0] + lb * st.pw[v1] + lc * st.pw[v2];
  r.pb = r.pa - pw;
}

template <class R, bool MASK, class PLA, class PLB>
TC_HDN EpaOut<R> epa(const PLA& A, const PLB& B, const Simplex<R>& start, R scale, EpaStore<R>& st,
                     ScanCount& scanned)
{
  EpaOut<R> r;
  EpaState<R> E;
  if (epaInit<R, MASK>(A, B, start, scale, st, E, r, scanned))
    return r;
  int code;
  if constexpr (MASK)
  {
    while ((code = epaTripMasked<R>(A, B, st, E, st.erow, scanned)) == EPA_CONTINUE)
    {
    }
  }
  else
  {
    while ((code = epaTrip<R, MASK>(A, B, st, E, scanned)) == EPA_CONTINUE)
    {
    }
  }
  if (code == EPA_OVERFLOW)
  {
    r.ok = false;
    r.iters = E.iters;
    return r;
  }
  r.converged = code != EPA_UNCONVERGED;
  epaFinish<R, MASK>(st, E, r);
  return r;
}

// length scale of the Minkowski difference used by EPA's degeneracy thresholds
template <class R, class PLA, class PLB>
TC_HD R epaScale(const PLA& A, R brad_a, const PLB& B, R brad_b)
{
  R s = 2 * (brad_a + brad_b);
  if (A.cast)
    s += norm(A.p1.apply(A.sh.center) - A.p0.apply(A.sh.center));
  if (B.cast)
    s += norm(B.p1.apply(B.sh.center) - B.p0.apply(B.sh.center));
  return s;
}

enum
{
  ND_NONE = 0,         // farther apart than the threshold
  ND_CONTACT = 1,
  ND_EPA_OVERFLOW = 2, // polytope store too small: item must be re-run on the deep-expansion path
  ND_NEED_EPA = 3      // cores touch / overlap: the simplex goes to the EPA stage
};

template <class R>
struct DistOut
{
  int status;
  R distance;
  Vec3<R> normal, pa, pb;  // normal from A to B; world witness points on the surfaces
  bool used_epa, converged;
  int gjk_iters, epa_iters;
  ScanCount scanned;
};

template <class R>
TC_HD void clearDist(DistOut<R>& out)
{
  out.status = ND_NONE;
  out.distance = 0;
  out.used_epa = false;
  out.converged = true;
  out.gjk_iters = 0;
  out.epa_iters = 0;
  out.scanned = 0;
}

template <class R, class PLA, class PLB>
TC_HD void finishDist(const PLA& A, const PLB& B, R threshold, R d_core, Vec3<R> n, Vec3<R> pa, Vec3<R> pb,
                      DistOut<R>& out)
{
  const R ra = A.sh.coreRadius(), rb = B.sh.coreRadius();
  out.distance = d_core - ra - rb;
  if (out.distance > threshold)
  {
    out.status = ND_NONE;
    return;
  }
  out.status = ND_CONTACT;
  out.normal = n;
  out.pa = pa + ra * n;
  out.pb = pb - rb * n;
}

// stage 1: GJK.  ND_NONE / ND_CONTACT are final; ND_NEED_EPA leaves the simplex in S for stage 2
template <class R, class PLA, class PLB>
TC_HD void distanceGjk(const PLA& A, const PLB& B, R threshold, Simplex<R>& S, DistOut<R>& out)
{
  clearDist(out);
  const R ra = A.sh.coreRadius(), rb = B.sh.coreRadius();
  const GjkOut<R> g = gjk(A, B, threshold + ra + rb, S, out.scanned);
  out.gjk_iters = g.iters;
  out.converged = g.converged;
  if (g.separated_beyond)
    return;
  if (g.intersect)
  {
    out.status = ND_NEED_EPA;
    return;
  }
  if (sizeof(R) == sizeof(float) && !g.certified)
  {
    out.status = ND_EPA_OVERFLOW;  // float run without an error bound: the FP64 re-run kernel takes the item
    return;
  }
  finishDist(A, B, threshold, g.dist, (R(1) / g.dist) * (g.pb - g.pa), g.pa, g.pb, out);
}

// stage 2: EPA from the GJK simplex (out carries the stage-1 counters and convergence flag)
template <class R, bool MASK, class PLA, class PLB>
TC_HD void distanceEpa(const PLA& A, R brad_a, const PLB& B, R brad_b, R threshold, const Simplex<R>& S,
                       EpaStore<R>& store, DistOut<R>& out)
{
  const EpaOut<R> e = epa<R, MASK>(A, B, S, epaScale(A, brad_a, B, brad_b), store, out.scanned);
  out.used_epa = true;
  out.epa_iters = e.iters;
  if (!e.ok)
  {
    out.status = ND_EPA_OVERFLOW;
    return;
  }
  out.converged = out.converged && e.converged;
  finishDist(A, B, threshold, -e.depth, e.normal, e.pa, e.pb, out);
}

template <class R, bool MASK, class PLA, class PLB>
TC_HDN DistOut<R> signedDistance(const PLA& A, R brad_a, const PLB& B, R brad_b, R threshold,
                                 EpaStore<R>& store)
{
  DistOut<R> out;
  Simplex<R> S;
  distanceGjk(A, B, threshold, S, out);
  if (out.status == ND_NEED_EPA)
    distanceEpa<R, MASK>(A, brad_a, B, brad_b, threshold, S, store, out);
  return out;
}
}  // namespace tc

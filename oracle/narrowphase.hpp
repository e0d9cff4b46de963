// TEST INFRASTRUCTURE ONLY — part of the CPU oracle (see oracle/README.md).
//
// Convex-convex signed distance: GJK for separated shapes, EPA for penetration depth.  This is the
// arithmetic that, in the reference, lives in an un-vendored dependency: tesseract_collision's Bullet
// managers (btGjkPairDetector + btGjkEpaPenetrationDepthSolver, double precision), selected by
// trajopt_common/data/contact_manager_plugins.yaml:5,8,25 and reached from
// trajopt/src/collision_terms.cpp:667,889,1134,1157 and
// trajopt_ifopt/src/constraints/collision/{discrete,continuous}_collision_evaluators.cpp:166,228,244,449.
// Restated here from the published algorithms (Gilbert-Johnson-Keerthi 1988; van den Bergen's EPA
// 2001), on exact geometry: no Bullet collision-margin shell; spheres / capsules are a core point /
// segment plus a radius exactly as Bullet treats a sphere's margin.
#pragma once
#include <algorithm>
#include <cfloat>
#include <vector>

#include "math3.hpp"

namespace orc
{
enum ShapeType
{
  SPHERE = 0,
  HULL = 1,
  CAPSULE = 2,
  CYLINDER = 3,
  CONE = 4
};

// A convex shape expressed in its LINK frame (shape_pose already folded in).
struct Shape
{
  int type{ HULL };
  int link{ 0 };
  int shape_id{ 0 };
  double radius{ 0 };   // sphere / capsule: inflation radius of the core; cylinder / cone: radius of the solid
  double half_len{ 0 };
  V3 center;            // sphere centre, capsule/cylinder/cone centre
  V3 axis{ 0, 0, 1 };   // unit axis of capsule/cylinder/cone in the link frame
  std::vector<V3> verts;  // hull vertices in the link frame
  V3 aabb_c, aabb_h;    // link-frame AABB (radius included)
  double brad{ 0 };     // bounding radius of the core about `center`
  double coreRadius() const { return (type == SPHERE || type == CAPSULE) ? radius : 0.0; }
};

// support point of the CORE of `s` in link-frame direction d (not normalised).
// Tie band (SUPPORT_TIE): on the side of a capsule / cylinder the contact normal is perpendicular to the axis and on
// the side of a cone it makes exactly the half-angle with it: the end-cap decision is a tie in exact arithmetic and a
// coin flip on the last bit of d in floating point (Bullet's own support mappings have the same sign test).  A value
// within 1e-12 |d| of the tie is taken as the tie and resolved one fixed way (+axis end; cone: base rim), so that two
// implementations whose d differ in the last bits agree.  The support value loses at most 1e-12 |d| x length.
constexpr double SUPPORT_TIE = 1e-12;
inline V3 supportLocal(const Shape& s, V3 d, long* scanned = nullptr)
{
  switch (s.type)
  {
    case SPHERE:
      return s.center;
    case CAPSULE:
    {
      const double al = dot(d, s.axis);
      const double tie = SUPPORT_TIE * norm(d);
      return s.center + ((al >= -tie) ? s.half_len : -s.half_len) * s.axis;
    }
    case CYLINDER:
    {
      const double al = dot(d, s.axis);
      const double tie = SUPPORT_TIE * norm(d);
      V3 rad = d - al * s.axis;
      const double rn = norm(rad);
      V3 p = s.center + ((al >= -tie) ? s.half_len : -s.half_len) * s.axis;
      if (rn > 1e-300)
        p += (s.radius / rn) * rad;
      return p;
    }
    case CONE:
    {
      // apex at +half_len, base disc of `radius` at -half_len
      const double al = dot(d, s.axis);
      const double dn = norm(d);
      const double tie = SUPPORT_TIE * dn;
      V3 rad = d - al * s.axis;
      const double rn = norm(rad);
      const double h = 2 * s.half_len;
      const double sin_a = s.radius / std::sqrt(s.radius * s.radius + h * h);
      if (al > dn * sin_a + tie)
        return s.center + s.half_len * s.axis;
      V3 p = s.center - s.half_len * s.axis;
      if (rn > 1e-300)
        p += (s.radius / rn) * rad;
      return p;
    }
    default:
    {
      double best = -DBL_MAX;
      std::size_t bi = 0;
      for (std::size_t i = 0; i < s.verts.size(); ++i)
      {
        const double v = dot(d, s.verts[i]);
        if (v > best)
        {
          best = v;
          bi = i;
        }
      }
      if (scanned)
        *scanned += static_cast<long>(s.verts.size());
      return s.verts[bi];
    }
  }
}

// A shape placed in the world; `cast` = swept between pose0 and pose1 (convex hull of both placements).
// Restates tesseract_collision's CastHullShape support rule [upstream, recalled]:
//   sv0 = s(d), sv1 = T01 * s(T01^-1 d); return dot(d, sv0) > dot(d, sv1) ? sv0 : sv1.
struct Placed
{
  const Shape* shape{ nullptr };
  Pose pose0, pose1;
  bool cast{ false };
  V3 support(V3 d, long* scanned = nullptr) const
  {
    const V3 p0 = pose0.apply(supportLocal(*shape, pose0.rotT(d), scanned));
    if (!cast)
      return p0;
    const V3 p1 = pose1.apply(supportLocal(*shape, pose1.rotT(d), scanned));
    return (dot(d, p0) > dot(d, p1)) ? p0 : p1;
  }
  V3 center() const
  {
    const V3 c0 = pose0.apply(shape->center);
    return cast ? 0.5 * (c0 + pose1.apply(shape->center)) : c0;
  }
};

struct Simplex
{
  int n{ 0 };
  V3 w[4], a[4], b[4];
  double lam[4]{ 0, 0, 0, 0 };
};

namespace detail
{
// closest point of triangle (a,b,c) to the origin; barycentric weights out (Ericson, RTCD 5.1.5)
inline V3 closestTriangle(V3 a, V3 b, V3 c, double& la, double& lb, double& lc)
{
  const V3 ab = b - a, ac = c - a;
  const double d1 = -dot(ab, a), d2 = -dot(ac, a);
  if (d1 <= 0 && d2 <= 0)
  {
    la = 1, lb = 0, lc = 0;
    return a;
  }
  const double d3 = -dot(ab, b), d4 = -dot(ac, b);
  if (d3 >= 0 && d4 <= d3)
  {
    la = 0, lb = 1, lc = 0;
    return b;
  }
  const double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0)
  {
    const double t = d1 / (d1 - d3);
    la = 1 - t, lb = t, lc = 0;
    return a + t * ab;
  }
  const double d5 = -dot(ab, c), d6 = -dot(ac, c);
  if (d6 >= 0 && d5 <= d6)
  {
    la = 0, lb = 0, lc = 1;
    return c;
  }
  const double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0)
  {
    const double t = d2 / (d2 - d6);
    la = 1 - t, lb = 0, lc = t;
    return a + t * ac;
  }
  const double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0)
  {
    const double t = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    la = 0, lb = 1 - t, lc = t;
    return b + t * (c - b);
  }
  const double den = 1.0 / (va + vb + vc);
  lb = vb * den;
  lc = vc * den;
  la = 1 - lb - lc;
  return a + lb * ab + lc * ac;
}
}  // namespace detail

// Reduce `s` to the smallest sub-simplex containing the point closest to the origin, fill lam, return
// that point.  Returns with s.n == 4 only when the origin is inside the tetrahedron.
inline V3 closestOnSimplex(Simplex& s)
{
  auto keep = [&s](std::initializer_list<int> idx, std::initializer_list<double> lam) {
    Simplex o;
    auto li = lam.begin();
    for (int i : idx)
    {
      if (*li > 0)
      {
        o.w[o.n] = s.w[i];
        o.a[o.n] = s.a[i];
        o.b[o.n] = s.b[i];
        o.lam[o.n] = *li;
        ++o.n;
      }
      ++li;
    }
    s = o;
  };
  if (s.n == 1)
  {
    s.lam[0] = 1;
    return s.w[0];
  }
  if (s.n == 2)
  {
    const V3 ab = s.w[1] - s.w[0];
    const double den = norm2(ab);
    double t = (den > 0) ? -dot(s.w[0], ab) / den : 0.0;
    t = std::min(1.0, std::max(0.0, t));
    const V3 p = s.w[0] + t * ab;
    keep({ 0, 1 }, { 1 - t, t });
    return p;
  }
  if (s.n == 3)
  {
    double la, lb, lc;
    const V3 p = detail::closestTriangle(s.w[0], s.w[1], s.w[2], la, lb, lc);
    keep({ 0, 1, 2 }, { la, lb, lc });
    return p;
  }
  // tetrahedron (Ericson, RTCD 5.1.6): test the faces whose plane separates the origin from the 4th vertex
  static const int F[4][4] = { { 0, 1, 2, 3 }, { 0, 1, 3, 2 }, { 0, 2, 3, 1 }, { 1, 2, 3, 0 } };
  double best = DBL_MAX;
  V3 bestp;
  int bf = -1;
  double bl[3] = { 0, 0, 0 };
  for (int f = 0; f < 4; ++f)
  {
    const V3 a = s.w[F[f][0]], b = s.w[F[f][1]], c = s.w[F[f][2]], d = s.w[F[f][3]];
    const V3 n = cross(b - a, c - a);
    const double sp = -dot(a, n);     // origin side
    const double sd = dot(d - a, n);  // opposite vertex side
    const bool degenerate = std::fabs(sd) <= 1e-14 * std::sqrt(norm2(n) * norm2(d - a) + 1e-300);
    if (sp * sd < 0 || degenerate)
    {
      double la, lb, lc;
      const V3 p = detail::closestTriangle(a, b, c, la, lb, lc);
      const double d2 = norm2(p);
      if (d2 < best)
      {
        best = d2;
        bestp = p;
        bf = f;
        bl[0] = la, bl[1] = lb, bl[2] = lc;
      }
    }
  }
  if (bf < 0)
  {
    // origin inside: weights are not needed by the caller (EPA takes over)
    s.lam[0] = s.lam[1] = s.lam[2] = s.lam[3] = 0.25;
    return V3();
  }
  keep({ F[bf][0], F[bf][1], F[bf][2] }, { bl[0], bl[1], bl[2] });
  return bestp;
}

struct GjkResult
{
  bool separated_beyond{ false };  // proven farther apart than the threshold: no contact, nothing else valid
  bool intersect{ false };         // cores touch / overlap: run EPA
  bool converged{ true };
  double dist{ 0 };                // core distance (>= 0)
  V3 pa, pb;                       // closest points on the cores (world)
  Simplex simplex;
  int iters{ 0 };
  long scanned{ 0 };
};

constexpr double GJK_TOL = 1e-10;        // absolute gap |v| - v.w/|v| at which GJK stops [m]
constexpr double GJK_TOUCH = 1e-7;       // core distance below which the pair is handed to EPA [m]
constexpr int GJK_MAX_ITERS = 128;

inline GjkResult gjk(const Placed& A, const Placed& B, double threshold_core)
{
  GjkResult r;
  auto sup = [&](V3 d, Simplex& s, int k) {
    s.a[k] = A.support(d, &r.scanned);
    s.b[k] = B.support(-d, &r.scanned);
    s.w[k] = s.a[k] - s.b[k];
  };
  V3 v = A.center() - B.center();
  if (norm2(v) < 1e-24)
    v = V3(1, 0, 0);
  Simplex& S = r.simplex;
  sup(-v, S, 0);
  S.n = 1;
  S.lam[0] = 1;
  v = S.w[0];
  double vv = norm2(v);
  for (r.iters = 1; r.iters <= GJK_MAX_ITERS; ++r.iters)
  {
    if (vv <= GJK_TOUCH * GJK_TOUCH)
    {
      r.intersect = true;
      break;
    }
    Simplex T = S;
    sup(-v, T, T.n);
    const V3 w = T.w[T.n];
    const double vw = dot(v, w);
    if (vw > 0 && vw * vw > threshold_core * threshold_core * vv)
    {
      r.separated_beyond = true;
      return r;
    }
    if (vv - vw <= GJK_TOL * std::sqrt(vv))
      break;
    bool dup = false;
    for (int i = 0; i < S.n; ++i)
      if (norm2(w - S.w[i]) <= 1e-28 * std::max(1.0, norm2(w)))
        dup = true;
    if (dup)
      break;
    T.n += 1;
    const V3 vn = closestOnSimplex(T);
    if (T.n == 4)
    {
      S = T;
      r.intersect = true;
      break;
    }
    const double vvn = norm2(vn);
    if (vvn >= vv)
      break;  // no progress within rounding: S already holds the closest point
    S = T;
    v = vn;
    vv = vvn;
    if (r.iters == GJK_MAX_ITERS)
      r.converged = false;
  }
  if (!r.intersect)
  {
    r.dist = std::sqrt(vv);
    for (int i = 0; i < S.n; ++i)
    {
      r.pa += S.lam[i] * S.a[i];
      r.pb += S.lam[i] * S.b[i];
    }
  }
  return r;
}

struct EpaResult
{
  bool ok{ false };
  bool converged{ true };
  double depth{ 0 };  // signed plane distance of the origin to the closest face (>= 0 when inside)
  V3 normal;          // outward normal of that face = direction from A to B
  V3 pa, pb;
  int iters{ 0 };
  long scanned{ 0 };
};

constexpr double EPA_TOL = 1e-10;
constexpr int EPA_MAX_ITERS = 255;
constexpr int EPA_STALL_TRIPS = 6;

inline EpaResult epa(const Placed& A, const Placed& B, const Simplex& start)
{
  struct Vtx
  {
    V3 w, a, b;
  };
  struct Face
  {
    int v[3];
    V3 n;
    double d;
    bool alive;
  };
  EpaResult r;
  std::vector<Vtx> P;
  auto sup = [&](V3 d) {
    Vtx x;
    x.a = A.support(d, &r.scanned);
    x.b = B.support(-d, &r.scanned);
    x.w = x.a - x.b;
    return x;
  };
  for (int i = 0; i < start.n; ++i)
    P.push_back({ start.w[i], start.a[i], start.b[i] });

  // length scale of the Minkowski difference for the "is this point new / is this volume non-zero"
  // decisions: bounding diameters of both cores plus the sweep lengths
  double scale = 2 * (A.shape->brad + B.shape->brad);
  if (A.cast)
    scale += norm(A.pose1.apply(A.shape->center) - A.pose0.apply(A.shape->center));
  if (B.cast)
    scale += norm(B.pose1.apply(B.shape->center) - B.pose0.apply(B.shape->center));
  const double eps_len = 1e-9 * std::max(scale, 1e-3);

  // --- grow the GJK simplex to a tetrahedron with non-zero volume ---
  if (P.size() == 4)
  {
    const double vol = dot(cross(P[1].w - P[0].w, P[2].w - P[0].w), P[3].w - P[0].w);
    if (std::fabs(vol) <= eps_len * eps_len * eps_len)
      P.pop_back();
  }
  if (P.size() == 1)
  {
    const V3 dirs[6] = { { 1, 0, 0 }, { -1, 0, 0 }, { 0, 1, 0 }, { 0, -1, 0 }, { 0, 0, 1 }, { 0, 0, -1 } };
    for (const V3& d : dirs)
    {
      Vtx x = sup(d);
      if (norm(x.w - P[0].w) > eps_len)
      {
        P.push_back(x);
        break;
      }
    }
  }
  if (P.size() == 2)
  {
    const V3 e = P[1].w - P[0].w;
    V3 ax(1, 0, 0);
    if (std::fabs(e.y) <= std::fabs(e.x) && std::fabs(e.y) <= std::fabs(e.z))
      ax = V3(0, 1, 0);
    else if (std::fabs(e.z) <= std::fabs(e.x) && std::fabs(e.z) <= std::fabs(e.y))
      ax = V3(0, 0, 1);
    V3 d0 = cross(e, ax);
    const V3 d1 = cross(e, d0);
    const double en = norm(e);
    const V3 cand[4] = { d0, -d0, d1, -d1 };
    for (const V3& d : cand)
    {
      Vtx x = sup(d);
      if (norm(cross(x.w - P[0].w, e)) > eps_len * en)
      {
        P.push_back(x);
        break;
      }
    }
  }
  if (P.size() == 3)
  {
    const V3 n = cross(P[1].w - P[0].w, P[2].w - P[0].w);
    const double area = norm(n);
    for (int sgn = 0; sgn < 2 && P.size() == 3; ++sgn)
    {
      Vtx x = sup(sgn == 0 ? n : -n);
      if (std::fabs(dot(x.w - P[0].w, n)) > eps_len * area)
        P.push_back(x);
    }
  }
  if (P.size() < 4)
  {
    // The Minkowski difference is flat (point / segment / polygon cores): depth of the cores is zero.
    // The normal is the direction in which the difference has no extent.
    r.ok = true;
    r.depth = 0;
    V3 n(0, 0, 1);
    if (P.size() == 3)
      n = cross(P[1].w - P[0].w, P[2].w - P[0].w);
    else if (P.size() == 2)
    {
      const V3 e = P[1].w - P[0].w;
      V3 ax(1, 0, 0);
      if (std::fabs(e.y) <= std::fabs(e.x) && std::fabs(e.y) <= std::fabs(e.z))
        ax = V3(0, 1, 0);
      else if (std::fabs(e.z) <= std::fabs(e.x) && std::fabs(e.z) <= std::fabs(e.y))
        ax = V3(0, 0, 1);
      n = cross(e, ax);
    }
    const double nn = norm(n);
    r.normal = (nn > 0) ? (1.0 / nn) * n : V3(0, 0, 1);
    Simplex s = start;
    closestOnSimplex(s);
    for (int i = 0; i < s.n; ++i)
    {
      r.pa += s.lam[i] * s.a[i];
      r.pb += s.lam[i] * s.b[i];
    }
    return r;
  }

  std::vector<Face> faces;
  const V3 centroid = 0.25 * (P[0].w + P[1].w + P[2].w + P[3].w);
  auto addFace = [&](int i, int j, int k, bool orient_by_centroid) {
    Face f;
    f.v[0] = i, f.v[1] = j, f.v[2] = k;
    V3 n = cross(P[j].w - P[i].w, P[k].w - P[i].w);
    if (orient_by_centroid && dot(n, P[i].w - centroid) < 0)
    {
      std::swap(f.v[1], f.v[2]);
      n = -n;
    }
    const double nn = norm(n);
    f.alive = true;
    if (nn > 1e-300)
    {
      f.n = (1.0 / nn) * n;
      f.d = dot(f.n, P[f.v[0]].w);
    }
    else
    {
      f.n = V3(0, 0, 0);
      f.d = DBL_MAX;  // degenerate sliver: never selected
    }
    // dead slots are reused first-fit so the face array stays bounded by the number of live faces
    for (Face& slot : faces)
      if (!slot.alive)
      {
        slot = f;
        return;
      }
    faces.push_back(f);
  };
  addFace(0, 1, 2, true);
  addFace(0, 1, 3, true);
  addFace(0, 2, 3, true);
  addFace(1, 2, 3, true);

  // On a curved core (cylinder, cone) the new vertices crowd on the surface, the faces turn into slivers whose plane
  // distance carries ~1e-10 m of rounding, and a few items dither at the tolerance for a hundred trips.  Stall exit:
  // EPA_STALL_TRIPS trips without narrowing the bracket between the closest face and the smallest support value by 10 %
  // end the expansion on the current closest face.  Polytope pairs never take it (finite termination).
  const bool stall_exit = A.shape->type >= CYLINDER || B.shape->type >= CYLINDER;
  double ub_best = DBL_MAX, width_best = DBL_MAX;
  int stall = 0;
  int best = -1;
  for (r.iters = 1; r.iters <= EPA_MAX_ITERS; ++r.iters)
  {
    best = -1;
    for (std::size_t f = 0; f < faces.size(); ++f)
      if (faces[f].alive && (best < 0 || faces[f].d < faces[static_cast<std::size_t>(best)].d))
        best = static_cast<int>(f);
    const Face bf = faces[static_cast<std::size_t>(best)];
    const Vtx x = sup(bf.n);
    const double sv = dot(bf.n, x.w);
    if (sv - bf.d <= EPA_TOL)
      break;
    if (stall_exit)
    {
      // stall exit on curved cores (see EPA_STALL_TRIPS): the bracket [closest face, smallest support value] has not
      // narrowed by 10 % for EPA_STALL_TRIPS trips -> the sliver faces are dithering at their rounding level
      if (sv < ub_best)
        ub_best = sv;
      const double width = ub_best - bf.d;
      if (width < 0.9 * width_best)
      {
        width_best = width;
        stall = 0;
      }
      else if (++stall >= EPA_STALL_TRIPS)
        break;
    }
    if (r.iters == EPA_MAX_ITERS)
    {
      r.converged = false;
      break;
    }
    // remove the faces that see x, stitch the horizon to x
    const int xi = static_cast<int>(P.size());
    P.push_back(x);
    std::vector<std::pair<int, int>> edges;
    for (Face& f : faces)
    {
      if (!f.alive || f.d == DBL_MAX)
        continue;
      if (dot(f.n, x.w) - f.d > 1e-14 * std::max(scale, 1e-3))
      {
        f.alive = false;
        for (int e = 0; e < 3; ++e)
        {
          const std::pair<int, int> ed(f.v[e], f.v[(e + 1) % 3]);
          auto twin = std::find(edges.begin(), edges.end(), std::make_pair(ed.second, ed.first));
          if (twin != edges.end())
            edges.erase(twin);
          else
            edges.push_back(ed);
        }
      }
    }
    if (edges.empty())
    {
      // numerically x is not outside any face although the gap test said so: accept current face
      P.pop_back();
      break;
    }
    for (const auto& ed : edges)
      addFace(ed.first, ed.second, xi, false);
  }
  const Face bf = faces[static_cast<std::size_t>(best)];
  r.ok = true;
  r.depth = bf.d;
  r.normal = bf.n;
  // Witness points: the face with the smallest PLANE distance may be one of several coplanar triangles of
  // a larger Minkowski face; the witness is taken on the triangle that is actually closest to the origin.
  std::size_t wf = static_cast<std::size_t>(best);
  double wbest = DBL_MAX, la = 0, lb = 0, lc = 0;
  for (std::size_t f = 0; f < faces.size(); ++f)
  {
    const Face& fc = faces[f];
    if (!fc.alive || fc.d == DBL_MAX || fc.d > bf.d + eps_len)
      continue;
    double a0, a1, a2;
    const V3 q = detail::closestTriangle(P[static_cast<std::size_t>(fc.v[0])].w, P[static_cast<std::size_t>(fc.v[1])].w,
                                         P[static_cast<std::size_t>(fc.v[2])].w, a0, a1, a2);
    const double q2 = norm2(q);
    if (q2 < wbest)
    {
      wbest = q2;
      wf = f;
      la = a0, lb = a1, lc = a2;
    }
  }
  const Face& wfc = faces[wf];
  const Vtx &a = P[static_cast<std::size_t>(wfc.v[0])], &b = P[static_cast<std::size_t>(wfc.v[1])],
            &c = P[static_cast<std::size_t>(wfc.v[2])];
  r.pa = la * a.a + lb * b.a + lc * c.a;
  r.pb = la * a.b + lb * b.b + lc * c.b;
  return r;
}

struct DistanceResult
{
  bool contact{ false };  // distance <= threshold
  double distance{ 0 };
  V3 normal;              // from A to B
  V3 pa, pb;              // world witness points on the surfaces
  bool used_epa{ false };
  bool converged{ true };
  int gjk_iters{ 0 };
  int epa_iters{ 0 };
  long scanned{ 0 };
};

// Signed distance between two placed convex shapes; reports only when distance <= threshold
// (tesseract's contactTest drops results beyond the contact margin; the evaluators then filter on
// margin + buffer again, trajopt/src/collision_terms.cpp:680-684).
inline DistanceResult signedDistance(const Placed& A, const Placed& B, double threshold)
{
  DistanceResult out;
  const double ra = A.shape->coreRadius(), rb = B.shape->coreRadius();
  GjkResult g = gjk(A, B, threshold + ra + rb);
  out.gjk_iters = g.iters;
  out.scanned = g.scanned;
  out.converged = g.converged;
  if (g.separated_beyond)
    return out;
  V3 n, pa, pb;
  double d_core;
  if (!g.intersect)
  {
    d_core = g.dist;
    n = (1.0 / g.dist) * (g.pb - g.pa);
    pa = g.pa;
    pb = g.pb;
  }
  else
  {
    EpaResult e = epa(A, B, g.simplex);
    out.used_epa = true;
    out.epa_iters = e.iters;
    out.scanned += e.scanned;
    out.converged = out.converged && e.converged;
    d_core = -e.depth;
    n = e.normal;
    pa = e.pa;
    pb = e.pb;
  }
  out.distance = d_core - ra - rb;
  if (out.distance > threshold)
    return out;
  out.contact = true;
  out.normal = n;
  out.pa = pa + ra * n;
  out.pb = pb - rb * n;
  return out;
}
}  // namespace orc

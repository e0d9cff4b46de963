// tc_contact.cuh — per-contact post-processing and gradient rows (one thread per contact).
//
//   finalizeContact : ContactResult fields, cast bookkeeping (cc_time / cc_type / cc_transform),
//                     LVS merge rules and the evaluators' result filters
//       trajopt/src/collision_terms.cpp:669-687,851-897,1082-1159
//       trajopt_common/src/collision_utils.cpp:73-114 (removeInvalidContactResults)
//       trajopt_ifopt/.../continuous_collision_evaluators.cpp:204-247,423-457
//       [upstream of the reference tree: tesseract_collision calculateContinuousData /
//        GetAverageSupport and ContactResultMap::addInterpolatedCollisionResults]
//   contactGradient : trajopt_common::calcGradient / getGradient (collision_utils.cpp:116-221) and the
//                     legacy CollisionEvaluator::GetGradient (trajopt/src/collision_terms.cpp:195-332):
//                     KDL-style translational Jacobian re-referenced to the contact point, grad = +-n^T J
#pragma once
#include "tc_scene.hpp"

namespace tc
{
// one geometry pass: which evaluator / stack conventions apply
struct EvalParams
{
  int evaluator, stack;
  int B, T, terms_per_traj, n_terms;
  double lvs;
};

// raw narrow-phase hit appended by the GJK/EPA kernel
struct RawContact
{
  int32_t unit, sp, flags, pad;
  double distance;
  double normal[3], pa[3], pb[3];
};

TC_HD double linspaced(int cnt, double lo, double hi, int i)
{
  // Eigen::VectorXd::LinSpaced(cnt, lo, hi)(i)  (trajopt/src/collision_terms.cpp:846-848)
  const int size1 = (cnt == 1) ? 1 : cnt - 1;
  const double step = (cnt == 1) ? 0.0 : (hi - lo) / static_cast<double>(cnt - 1);
  const bool flip = fabs(hi) < fabs(lo);
  if (flip)
    return (i == 0) ? lo : (hi - static_cast<double>(size1 - i) * step);
  return (i == size1) ? hi : (lo + static_cast<double>(i) * step);
}

// joint values behind an accessor, so that the chain walks below take a stored state (QArray) or a state between two
// stored ones (QLerp) without a thread-local copy of it
struct QArray
{
  const double* q;
  TC_HD double operator()(int j) const { return q[j]; }
};
struct QLerp
{
  const double *q0, *q1;
  double t;
  TC_HD double operator()(int j) const { return q0[j] + (q1[j] - q0[j]) * t; }
};

// the joint values getGradient(dofvals0, dofvals1, ...) evaluates a link at: the first state for a contact at TIME0, the
// second for TIME1, the interpolated state otherwise (collision_utils.cpp:176-221) — one accessor, so that a warp whose
// contacts are of different kinds still runs ONE copy of the chain walk
struct QByKind
{
  const double *q0, *q1;
  double t;
  int kind;  // TC_CC_TIME0 / TC_CC_TIME1 / anything else = interpolate
  TC_HD double operator()(int j) const
  {
    if (kind == TC_CC_TIME0)
      return q0[j];
    if (kind == TC_CC_TIME1)
      return q1[j];
    return q0[j] + (q1[j] - q0[j]) * t;
  }
};

// pose of frame f given its parent's pose; axis_w = world joint axis
template <class QA>
TC_HD Pose<double> frameStepA(const SceneView& v, const Pose<double>& parent, int f, const QA& q, Vec3<double>& axis_w)
{
  const Pose<double> origin = Pose<double>::from12(v.frame_origin + 12 * f);
  Pose<double> T = compose(parent, origin);
  const Vec3<double> ax(v.frame_axis[3 * f], v.frame_axis[3 * f + 1], v.frame_axis[3 * f + 2]);
  axis_w = T.rot(ax);
  const int type = v.frame_type[f];
  if (type == TC_JOINT_REVOLUTE)
    T = rotateAbout(T, ax, q(v.frame_dof[f]));
  else if (type == TC_JOINT_PRISMATIC)
    T.t = T.t + q(v.frame_dof[f]) * axis_w;
  return T;
}

TC_HD Pose<double> frameStep(const SceneView& v, const Pose<double>& parent, int f, const double* q,
                             Vec3<double>& axis_w)
{
  return frameStepA(v, parent, f, QArray{ q }, axis_w);
}

// same with the joint value passed in (the FK kernel interpolates it on the fly)
TC_HD Pose<double> frameStepQ(const SceneView& v, const Pose<double>& parent, int f, double qf)
{
  const Pose<double> origin = Pose<double>::from12(v.frame_origin + 12 * f);
  Pose<double> T = compose(parent, origin);
  const Vec3<double> ax(v.frame_axis[3 * f], v.frame_axis[3 * f + 1], v.frame_axis[3 * f + 2]);
  const int type = v.frame_type[f];
  if (type == TC_JOINT_REVOLUTE)
    T = rotateAbout(T, ax, qf);
  else if (type == TC_JOINT_PRISMATIC)
    T.t = T.t + qf * T.rot(ax);
  return T;
}

// world pose of frame lf at joint values q (walks the ancestor path)
template <class QA>
TC_HDN inline Pose<double> framePoseA(const SceneView& v, int lf, const QA& q)
{
  Pose<double> T = Pose<double>::from12(v.base_pose);
  for (int k = v.path_ptr[lf]; k < v.path_ptr[lf + 1]; ++k)
  {
    Vec3<double> z;
    T = frameStepA(v, T, v.path_frames[k], q, z);
  }
  return T;
}

TC_HDN inline Pose<double> framePose(const SceneView& v, int lf, const double* q)
{
  return framePoseA(v, lf, QArray{ q });
}

// tv . J_trans[:, j] for every joint j on the path of `link`, for two reference points on the link (origin + ref0,
// origin + ref1, world frame) at the same joint values in one walk of the chain: emit(j, g0, g1).  These are the T0 and
// T1 blocks of trajopt_common::getGradient(q0, q1, ...) (collision_utils.cpp:215-216); joints off the path have zero
// columns and are not emitted.
// dot and cross product with every product and sum placed by hand (one fused multiply-add per term after the first):
// the same bits from every instantiation of the walk below, whatever the compiler would have contracted around it
TC_HD double placedDot(Vec3<double> a, Vec3<double> b) { return fma(a.z, b.z, fma(a.y, b.y, TC_MUL(a.x, b.x))); }
TC_HD Vec3<double> placedCross(Vec3<double> a, Vec3<double> b)
{
  return Vec3<double>(fma(a.y, b.z, -TC_MUL(a.z, b.y)), fma(a.z, b.x, -TC_MUL(a.x, b.z)), fma(a.x, b.y, -TC_MUL(a.y, b.x)));
}

template <class QA, class EMIT>
TC_HDN inline void pointGradientWalk(const SceneView& v, int link, const QA& q, Vec3<double> ref0, Vec3<double> ref1,
                                     Vec3<double> tv, EMIT&& emit)
{
  const int lf = v.link_frame[link];
  const Vec3<double> pl = framePoseA(v, lf, q).t;
  const Vec3<double> p0 = pl + ref0, p1 = pl + ref1;
  Pose<double> T = Pose<double>::from12(v.base_pose);
  for (int k = v.path_ptr[lf]; k < v.path_ptr[lf + 1]; ++k)
  {
    const int f = v.path_frames[k];
    Vec3<double> z;
    T = frameStepA(v, T, f, q, z);
    const int type = v.frame_type[f];
    if (type == TC_JOINT_FIXED)
      continue;
    const Vec3<double> c0 = (type == TC_JOINT_REVOLUTE) ? placedCross(z, p0 - T.t) : z;
    const Vec3<double> c1 = (type == TC_JOINT_REVOLUTE) ? placedCross(z, p1 - T.t) : z;
    emit(v.frame_dof[f], placedDot(c0, tv), placedDot(c1, tv));
  }
}

// g[j] = tv . J_trans[:, j]  with the reference point = link origin + ref_offset (world frame)
TC_HDN inline void pointGradient(const SceneView& v, int link, const double* q, Vec3<double> ref_offset,
                                 Vec3<double> tv, double* g)
{
  for (int j = 0; j < v.n_dof; ++j)
    g[j] = 0.0;
  pointGradientWalk(v, link, QArray{ q }, ref_offset, ref_offset, tv, [&](int j, double g0, double) { g[j] = g0; });
}

// two reference points on the same link at the same joint values
TC_HDN inline void pointGradient2(const SceneView& v, int link, const double* q, Vec3<double> ref0, Vec3<double> ref1,
                                  Vec3<double> tv, double* g0, double* g1)
{
  for (int j = 0; j < v.n_dof; ++j)
    g0[j] = g1[j] = 0.0;
  pointGradientWalk(v, link, QArray{ q }, ref0, ref1, tv, [&](int j, double a, double b) { g0[j] = a, g1[j] = b; });
}

// GetAverageSupport: mean of all hull vertices within BULLET_EPSILON of the running maximum
TC_HDN inline void averageSupport(const SceneView& v, const ShapeD& sh, Vec3<double> nl, double& outsupport,
                                  Vec3<double>& outpt)
{
  const double BULLET_EPSILON = 1e-3;
  if (sh.type == SH_HULL)
  {
    Vec3<double> sum;
    double cnt = 0, max_support = -1000;
    for (int i = 0; i < sh.vert_cnt; ++i)
    {
      const Vtx<double> q = v.verts[sh.vert_off + i];
      const Vec3<double> pt(q.x, q.y, q.z);
      const double sup = dot(pt, nl);
      if (sup > max_support + BULLET_EPSILON)
      {
        cnt = 1;
        sum = pt;
        max_support = sup;
      }
      else if (sup < max_support - BULLET_EPSILON)
      {
      }
      else
      {
        cnt += 1;
        sum = sum + pt;
      }
    }
    outsupport = max_support;
    outpt = (1.0 / cnt) * sum;
  }
  else
  {
    const ShapeCore<double> core = makeCore<double>(sh, nullptr);
    const double nn = norm(nl);
    outpt = primitiveSupport(core, nl);
    if (nn > 0)
      outpt = outpt + (core.coreRadius() / nn) * nl;
    outsupport = dot(nl, outpt);
  }
}

// the two GetAverageSupport scans of calculateContinuousData (directions in the T0 and the T1 frame) in ONE pass over
// the hull's vertices; each direction keeps its own running maximum and sum, updated in vertex order exactly as in
// averageSupport
TC_HDN inline void averageSupportHull2(const SceneView& v, const ShapeD& sh, Vec3<double> n0, Vec3<double> n1, double& out0,
                                       Vec3<double>& pt0, double& out1, Vec3<double>& pt1)
{
  const double BULLET_EPSILON = 1e-3;
  Vec3<double> sum0, sum1;
  double cnt0 = 0, cnt1 = 0, max0 = -1000, max1 = -1000;
  for (int i = 0; i < sh.vert_cnt; ++i)
  {
    const Vtx<double> q = v.verts[sh.vert_off + i];
    const Vec3<double> pt(q.x, q.y, q.z);
    const double a = dot(pt, n0), b = dot(pt, n1);
    if (a > max0 + BULLET_EPSILON)
    {
      cnt0 = 1;
      sum0 = pt;
      max0 = a;
    }
    else if (!(a < max0 - BULLET_EPSILON))
    {
      cnt0 += 1;
      sum0 = sum0 + pt;
    }
    if (b > max1 + BULLET_EPSILON)
    {
      cnt1 = 1;
      sum1 = pt;
      max1 = b;
    }
    else if (!(b < max1 - BULLET_EPSILON))
    {
      cnt1 += 1;
      sum1 = sum1 + pt;
    }
  }
  out0 = max0;
  pt0 = (1.0 / cnt0) * sum0;
  out1 = max1;
  pt1 = (1.0 / cnt1) * sum1;
}

// calculateContinuousData for one link of a cast contact
TC_HDN inline void continuousData(const SceneView& v, const ShapeD& sh, const Pose<double>& tf0, const Pose<double>& tf1,
                                  Vec3<double> pt_world, Vec3<double> normal_world, double& cc_time, int& cc_type)
{
  const double SUPPORT_TOL = 0.01, LENGTH_TOL = 0.001;
  double s0, s1;
  Vec3<double> pl0, pl1;
  if (sh.type == SH_HULL)
    averageSupportHull2(v, sh, tf0.rotT(normal_world), tf1.rotT(normal_world), s0, pl0, s1, pl1);
  else
  {
    averageSupport(v, sh, tf0.rotT(normal_world), s0, pl0);
    averageSupport(v, sh, tf1.rotT(normal_world), s1, pl1);
  }
  const Vec3<double> pw0 = tf0.apply(pl0), pw1 = tf1.apply(pl1);
  const double sup0 = dot(normal_world, pw0), sup1 = dot(normal_world, pw1);
  if (sup0 - sup1 > SUPPORT_TOL)
  {
    cc_time = 0;
    cc_type = TC_CC_TIME0;
  }
  else if (sup1 - sup0 > SUPPORT_TOL)
  {
    cc_time = 1;
    cc_type = TC_CC_TIME1;
  }
  else
  {
    const double l0c = norm(pt_world - pw0), l1c = norm(pt_world - pw1);
    cc_type = TC_CC_BETWEEN;
    cc_time = (l0c + l1c < LENGTH_TOL) ? 0.5 : l0c / (l0c + l1c);
  }
}

// trajopt_common::removeInvalidContactResults predicate (true = drop)
TC_HD bool invalidContact(double distance, int cc_type0, int cc_type1, double margin, double buffer, bool f0, bool f1)
{
  if (distance > margin + buffer)
    return true;
  if (!f0 && !f1)
    return false;
  if (f0)
  {
    if (cc_type0 != TC_CC_NONE && cc_type0 != TC_CC_TIME0)
      return false;
    if (cc_type1 != TC_CC_NONE && cc_type1 != TC_CC_TIME0)
      return false;
  }
  if (f1)
  {
    if (cc_type0 != TC_CC_NONE && cc_type0 != TC_CC_TIME1)
      return false;
    if (cc_type1 != TC_CC_NONE && cc_type1 != TC_CC_TIME1)
      return false;
  }
  return true;
}

// where a unit (tested state / swept sub-segment) sits in the batch
struct UnitInfo
{
  int term;      // global term index b * terms_per_traj + ti
  int traj, t;   // trajectory, timestep (DISCRETE) or segment start
  int substep;   // i
  int last;      // cnt - 1 of the term
  int state0;    // global sub-state index of the T0 pose
  bool interp;   // results go through the LVS merge rules
  bool f0, f1;   // vars0_fixed, vars1_fixed
};

// Builds the ContactResult of a raw hit in place (c may be the record in global memory: every field is written once,
// nothing is read back).  pose0[i] / pose1[i]: 3x4 row-major pose of link i at the unit's T0 / T1 state (pose1 == pose0
// for a link that does not move); they are loaded link by link, so that only one link's poses are live at a time.
// transform[i] / cc_transform[i] of the result are not stored: they are pose0[i] and, for a cast or interpolated
// contact, pose1[i] resp. pose0[i] (contactTransforms).  Returns false when the evaluators' filters drop the contact.
TC_HDN inline bool finalizeContact(const SceneView& v, const EvalParams& ep, const RawContact& raw, const UnitInfo& u,
                                   const double* const pose0[2], const double* const pose1[2], tc_contact& c)
{
  const ShapePairD sp = v.spairs[raw.sp];
  const PairD pr = v.pairs[sp.pair];
  const bool cast = ep.evaluator == TC_EVAL_CONTINUOUS || ep.evaluator == TC_EVAL_LVS_CONTINUOUS;
  const Vec3<double> n(raw.normal[0], raw.normal[1], raw.normal[2]);
  c.distance = raw.distance;
  c.normal[0] = n.x, c.normal[1] = n.y, c.normal[2] = n.z;
  c.link_ids[0] = pr.a, c.link_ids[1] = pr.b;
  c.traj = u.traj;
  c.term = u.t;
  c.substep = u.substep;
  c.pair = sp.pair;
  c.state = u.state0;
  c.flags = raw.flags;
  int cc_type[2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
  {
    const ShapeD& sh = v.shapes[i ? sp.sb : sp.sa];
    const Vec3<double> p = i ? Vec3<double>(raw.pb[0], raw.pb[1], raw.pb[2]) : Vec3<double>(raw.pa[0], raw.pa[1], raw.pa[2]);
    const bool active = (i ? pr.b : pr.a) < v.n_active;
    const Pose<double> tf0 = Pose<double>::from12(pose0[i]);
    const Vec3<double> pl = tf0.applyInv(p);
    c.nearest_points[i][0] = p.x, c.nearest_points[i][1] = p.y, c.nearest_points[i][2] = p.z;
    c.nearest_points_local[i][0] = pl.x, c.nearest_points_local[i][1] = pl.y, c.nearest_points_local[i][2] = pl.z;
    c.shape_id[i] = sh.shape_id;
    c.subshape_id[i] = -1;
    double time = -1;
    int type = TC_CC_NONE;
    // the cast shape's normal points away from itself
    if (cast && active)
      continuousData(v, sh, tf0, Pose<double>::from12(pose1[i]), p, i ? -n : n, time, type);
    if (u.interp && active)
    {
      const double dt = 1.0 / static_cast<double>(u.last);
      const double i_dt = static_cast<double>(u.substep) * dt;
      time = (time < 0) ? i_dt : i_dt + (time * dt);
      if (u.substep == 0 && (type == TC_CC_TIME0 || type == TC_CC_NONE))
        type = TC_CC_TIME0;
      else if (u.substep == u.last && (type == TC_CC_TIME1 || type == TC_CC_NONE))
        type = TC_CC_TIME1;
      else
        type = TC_CC_BETWEEN;
    }
    c.cc_time[i] = time;
    c.cc_type[i] = type;
    cc_type[i] = type;
  }
  if (pr.zero_coeff)
    return false;
  if (ep.evaluator == TC_EVAL_DISCRETE)
    return !(raw.distance > pr.margin + v.margin_buffer);
  return !invalidContact(raw.distance, cc_type[0], cc_type[1], pr.margin, v.margin_buffer, u.f0, u.f1);
}

// bits of GradOut::has: link l, block k (0 = gradients[], 1 = cc_gradients[]) -> bit 2*l + k
struct GradOut
{
  tc_gradient rec;
  int has;
};

// One LinkGradientResults (calcGradient): writes g[n_dof], returns the scale
TC_HDN inline double linkGradient(const SceneView& v, const tc_contact& c, const double* xf, int i, const double* q,
                                  bool isTimestep1, int& cc_type_out, double* g)
{
  double scale = 1;
  const double* link_tf = xf + 24 * i;  // transform[i]
  if (c.cc_type[i] != TC_CC_NONE)
  {
    scale = isTimestep1 ? c.cc_time[i] : (1 - c.cc_time[i]);
    cc_type_out = c.cc_type[i];
    if (isTimestep1)
      link_tf = xf + 24 * i + 12;  // cc_transform[i]
  }
  const Pose<double> T = Pose<double>::from12(link_tf);
  const Vec3<double> pl(c.nearest_points_local[i][0], c.nearest_points_local[i][1], c.nearest_points_local[i][2]);
  Vec3<double> tv(c.normal[0], c.normal[1], c.normal[2]);
  if (i == 0)
    tv = -tv;
  pointGradient(v, c.link_ids[i], q, T.rot(pl), tv, g);
  return scale;
}

// Gradients of one contact.  grad = [2 links][2 blocks][n_dof]; q0/q1 = the term's two timestep states
// (q1 unused for DISCRETE).  qt = scratch [n_dof].
TC_HDN inline void contactGradient(const SceneView& v, const EvalParams& ep, const tc_contact& c, const double* xf,
                                   const double* q0, const double* q1, bool f0, bool f1, double* qt, double* grad,
                                   GradOut& out)
{
  const int n = v.n_dof;
  const PairD pr = v.pairs[c.pair];
  out.rec.error = pr.margin - c.distance;
  out.rec.error_with_buffer = pr.margin + v.margin_buffer - c.distance;
  out.has = 0;
  for (int k = 0; k < 4 * n; ++k)
    grad[k] = 0.0;
  const bool single = ep.evaluator == TC_EVAL_DISCRETE;
  for (int i = 0; i < 2; ++i)
  {
    out.rec.scale0[i] = 1;
    out.rec.scale1[i] = 1;
    out.rec.has_gradient[i] = 0;
    out.rec.cc_type[i] = TC_CC_NONE;
    if (c.link_ids[i] >= v.n_active)
      continue;
    int cc0 = TC_CC_NONE, cc1 = TC_CC_NONE;
    bool h0 = false, h1 = false;
    double* g0 = grad + (2 * i + 0) * n;
    double* g1 = grad + (2 * i + 1) * n;
    if (ep.stack == TC_STACK_LEGACY)
    {
      // one GetGradient per free timestep block, each at that block's own joint values
      if (single || !f0)
      {
        out.rec.scale0[i] = linkGradient(v, c, xf, i, q0, false, cc0, g0);
        h0 = true;
      }
      if (!single && !f1)
      {
        out.rec.scale1[i] = linkGradient(v, c, xf, i, q1, true, cc1, g1);
        h1 = true;
      }
    }
    else if (single)
    {
      out.rec.scale0[i] = linkGradient(v, c, xf, i, q0, false, cc0, g0);
      h0 = true;
    }
    else
    {
      const double* qe = qt;
      if (c.cc_type[i] == TC_CC_TIME0)
        qe = q0;
      else if (c.cc_type[i] == TC_CC_TIME1)
        qe = q1;
      else
        for (int j = 0; j < n; ++j)
          qt[j] = q0[j] + (q1[j] - q0[j]) * c.cc_time[i];
      // both blocks share the joint values: one chain walk, two reference points
      const bool cc = c.cc_type[i] != TC_CC_NONE;
      out.rec.scale0[i] = cc ? (1 - c.cc_time[i]) : 1.0;
      out.rec.scale1[i] = cc ? c.cc_time[i] : 1.0;
      cc0 = cc1 = cc ? c.cc_type[i] : TC_CC_NONE;
      const Pose<double> T0 = Pose<double>::from12(xf + 24 * i);
      const Pose<double> T1 = cc ? Pose<double>::from12(xf + 24 * i + 12) : T0;
      const Vec3<double> pl(c.nearest_points_local[i][0], c.nearest_points_local[i][1], c.nearest_points_local[i][2]);
      Vec3<double> tv(c.normal[0], c.normal[1], c.normal[2]);
      if (i == 0)
        tv = -tv;
      pointGradient2(v, c.link_ids[i], qe, T0.rot(pl), T1.rot(pl), tv, g0, g1);
      h0 = h1 = true;
    }
    out.rec.has_gradient[i] = (h0 || h1) ? 1 : 0;
    out.rec.cc_type[i] = h0 ? cc0 : cc1;
    out.has |= (h0 ? 1 : 0) << (2 * i);
    out.has |= (h1 ? 1 : 0) << (2 * i + 1);
  }
}
}  // namespace tc

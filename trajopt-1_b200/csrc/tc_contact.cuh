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

// pose of frame f given its parent's pose; axis_w = world joint axis
TC_HD Pose<double> frameStep(const SceneView& v, const Pose<double>& parent, int f, const double* q,
                             Vec3<double>& axis_w)
{
  const Pose<double> origin = Pose<double>::from12(v.frame_origin + 12 * f);
  Pose<double> T = compose(parent, origin);
  const Vec3<double> ax(v.frame_axis[3 * f], v.frame_axis[3 * f + 1], v.frame_axis[3 * f + 2]);
  axis_w = T.rot(ax);
  const int type = v.frame_type[f];
  if (type == TC_JOINT_REVOLUTE)
    T = rotateAbout(T, ax, q[v.frame_dof[f]]);
  else if (type == TC_JOINT_PRISMATIC)
    T.t = T.t + q[v.frame_dof[f]] * axis_w;
  return T;
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
TC_HDN inline Pose<double> framePose(const SceneView& v, int lf, const double* q)
{
  Pose<double> T = Pose<double>::from12(v.base_pose);
  for (int k = v.path_ptr[lf]; k < v.path_ptr[lf + 1]; ++k)
  {
    Vec3<double> z;
    T = frameStep(v, T, v.path_frames[k], q, z);
  }
  return T;
}

// g[j] = tv . J_trans[:, j]  with the reference point = link origin + ref_offset (world frame)
TC_HDN inline void pointGradient(const SceneView& v, int link, const double* q, Vec3<double> ref_offset,
                                 Vec3<double> tv, double* g)
{
  const int lf = v.link_frame[link];
  for (int j = 0; j < v.n_dof; ++j)
    g[j] = 0.0;
  const Vec3<double> p = framePose(v, lf, q).t + ref_offset;
  Pose<double> T = Pose<double>::from12(v.base_pose);
  for (int k = v.path_ptr[lf]; k < v.path_ptr[lf + 1]; ++k)
  {
    const int f = v.path_frames[k];
    Vec3<double> z;
    T = frameStep(v, T, f, q, z);
    const int type = v.frame_type[f];
    if (type == TC_JOINT_FIXED)
      continue;
    const Vec3<double> col = (type == TC_JOINT_REVOLUTE) ? cross(z, p - T.t) : z;
    g[v.frame_dof[f]] = col.x * tv.x + col.y * tv.y + col.z * tv.z;
  }
}

// two reference points on the same link at the same joint values in one walk of the chain (the T0 and T1
// blocks of trajopt_common::getGradient(q0, q1, ...), collision_utils.cpp:215-216)
TC_HDN inline void pointGradient2(const SceneView& v, int link, const double* q, Vec3<double> ref0, Vec3<double> ref1,
                                  Vec3<double> tv, double* g0, double* g1)
{
  const int lf = v.link_frame[link];
  for (int j = 0; j < v.n_dof; ++j)
    g0[j] = g1[j] = 0.0;
  const Vec3<double> pl = framePose(v, lf, q).t;
  const Vec3<double> p0 = pl + ref0, p1 = pl + ref1;
  Pose<double> T = Pose<double>::from12(v.base_pose);
  for (int k = v.path_ptr[lf]; k < v.path_ptr[lf + 1]; ++k)
  {
    const int f = v.path_frames[k];
    Vec3<double> z;
    T = frameStep(v, T, f, q, z);
    const int type = v.frame_type[f];
    if (type == TC_JOINT_FIXED)
      continue;
    const Vec3<double> c0 = (type == TC_JOINT_REVOLUTE) ? cross(z, p0 - T.t) : z;
    const Vec3<double> c1 = (type == TC_JOINT_REVOLUTE) ? cross(z, p1 - T.t) : z;
    g0[v.frame_dof[f]] = c0.x * tv.x + c0.y * tv.y + c0.z * tv.z;
    g1[v.frame_dof[f]] = c1.x * tv.x + c1.y * tv.y + c1.z * tv.z;
  }
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

// calculateContinuousData for link slot i of a cast contact
TC_HDN inline void continuousData(const SceneView& v, tc_contact& c, const ShapeD& sh, const Pose<double>& tf0,
                                  const Pose<double>& tf1, Vec3<double> pt_world, Vec3<double> normal_world, int i)
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
    c.cc_time[i] = 0;
    c.cc_type[i] = TC_CC_TIME0;
  }
  else if (sup1 - sup0 > SUPPORT_TOL)
  {
    c.cc_time[i] = 1;
    c.cc_type[i] = TC_CC_TIME1;
  }
  else
  {
    const double l0c = norm(pt_world - pw0), l1c = norm(pt_world - pw1);
    c.cc_type[i] = TC_CC_BETWEEN;
    c.cc_time[i] = (l0c + l1c < LENGTH_TOL) ? 0.5 : l0c / (l0c + l1c);
  }
}

// trajopt_common::removeInvalidContactResults predicate (true = drop)
TC_HD bool invalidContact(const tc_contact& r, double margin, double buffer, bool f0, bool f1)
{
  if (r.distance > margin + buffer)
    return true;
  if (!f0 && !f1)
    return false;
  if (f0)
  {
    if (r.cc_type[0] != TC_CC_NONE && r.cc_type[0] != TC_CC_TIME0)
      return false;
    if (r.cc_type[1] != TC_CC_NONE && r.cc_type[1] != TC_CC_TIME0)
      return false;
  }
  if (f1)
  {
    if (r.cc_type[0] != TC_CC_NONE && r.cc_type[0] != TC_CC_TIME1)
      return false;
    if (r.cc_type[1] != TC_CC_NONE && r.cc_type[1] != TC_CC_TIME1)
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

// Builds the ContactResult of a raw hit.  xf (optional) = transform[0], cc_transform[0], transform[1],
// cc_transform[1] (4 x 12 doubles).  pose0/pose1: link poses of both links at the unit's T0 / T1 state.
// Returns false when the evaluators' filters drop the contact.
TC_HDN inline bool finalizeContact(const SceneView& v, const EvalParams& ep, const RawContact& raw, const UnitInfo& u,
                                   const Pose<double> pose0[2], const Pose<double> pose1[2], tc_contact& c, double* xf)
{
  const ShapePairD sp = v.spairs[raw.sp];
  const PairD pr = v.pairs[sp.pair];
  const ShapeD& A = v.shapes[sp.sa];
  const ShapeD& B = v.shapes[sp.sb];
  const bool cast = ep.evaluator == TC_EVAL_CONTINUOUS || ep.evaluator == TC_EVAL_LVS_CONTINUOUS;
  const Vec3<double> n(raw.normal[0], raw.normal[1], raw.normal[2]);
  const Vec3<double> p[2] = { Vec3<double>(raw.pa[0], raw.pa[1], raw.pa[2]),
                              Vec3<double>(raw.pb[0], raw.pb[1], raw.pb[2]) };
  c.distance = raw.distance;
  c.normal[0] = n.x, c.normal[1] = n.y, c.normal[2] = n.z;
  Pose<double> cc_tf[2];
  for (int i = 0; i < 2; ++i)
  {
    const Vec3<double> pl = pose0[i].applyInv(p[i]);
    c.nearest_points[i][0] = p[i].x, c.nearest_points[i][1] = p[i].y, c.nearest_points[i][2] = p[i].z;
    c.nearest_points_local[i][0] = pl.x, c.nearest_points_local[i][1] = pl.y, c.nearest_points_local[i][2] = pl.z;
    c.cc_time[i] = -1;
    c.cc_type[i] = TC_CC_NONE;
    c.subshape_id[i] = -1;
  }
  c.link_ids[0] = pr.a, c.link_ids[1] = pr.b;
  c.shape_id[0] = A.shape_id, c.shape_id[1] = B.shape_id;
  c.traj = u.traj;
  c.term = u.t;
  c.substep = u.substep;
  c.pair = sp.pair;
  c.state = u.state0;
  c.flags = raw.flags;
  const bool active[2] = { pr.a < v.n_active, pr.b < v.n_active };
  if (cast)
  {
    // the cast shape's normal points away from itself
    if (active[0])
    {
      continuousData(v, c, A, pose0[0], pose1[0], p[0], n, 0);
      cc_tf[0] = pose1[0];
    }
    if (active[1])
    {
      continuousData(v, c, B, pose0[1], pose1[1], p[1], -n, 1);
      cc_tf[1] = pose1[1];
    }
  }
  if (u.interp)
  {
    const bool discrete = !cast;
    const double dt = 1.0 / static_cast<double>(u.last);
    const double i_dt = static_cast<double>(u.substep) * dt;
    for (int j = 0; j < 2; ++j)
    {
      if (!active[j])
        continue;
      c.cc_time[j] = (c.cc_time[j] < 0) ? i_dt : i_dt + (c.cc_time[j] * dt);
      if (u.substep == 0 && (c.cc_type[j] == TC_CC_TIME0 || c.cc_type[j] == TC_CC_NONE))
        c.cc_type[j] = TC_CC_TIME0;
      else if (u.substep == u.last && (c.cc_type[j] == TC_CC_TIME1 || c.cc_type[j] == TC_CC_NONE))
        c.cc_type[j] = TC_CC_TIME1;
      else
        c.cc_type[j] = TC_CC_BETWEEN;
      if (discrete)
      {
        cc_tf[0] = pose0[0];
        cc_tf[1] = pose0[1];
      }
    }
  }
  if (xf)
  {
    pose0[0].to12(xf + 0);
    cc_tf[0].to12(xf + 12);
    pose0[1].to12(xf + 24);
    cc_tf[1].to12(xf + 36);
  }
  if (pr.zero_coeff)
    return false;
  if (ep.evaluator == TC_EVAL_DISCRETE)
    return !(c.distance > pr.margin + v.margin_buffer);
  return !invalidContact(c, pr.margin, v.margin_buffer, u.f0, u.f1);
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

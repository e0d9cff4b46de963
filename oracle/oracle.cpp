// TEST INFRASTRUCTURE ONLY.
//
// CPU oracle: a dependency-free FP64 restatement of trajopt 0.35's collision-term evaluation, used
// (a) as the parity checker for the CUDA path in tests/ and __graft_entry__.smoke(), and
// (b) as the timed host-CPU baseline in bench.py (cpu_baseline / --impl reference).
// Nothing under trajopt-1_b200/ may link, import or call this file.
//
// Parity status: PINNED only by the reference's four boxbot known-answer vectors
// (trajopt_ifopt/test/collision_unit.cpp:103-158) and by closed-form sphere/box/capsule cases; the
// reference itself (tesseract + Bullet + KDL) cannot be built here, so everything tesseract computes
// is restated from its published behaviour and marked [upstream, recalled].
//
// Structure mirrors the reference: one evaluator invocation per collision term, a contact manager
// with setCollisionObjectsTransform + contactTest, ContactResultMap keyed by link pair, then the
// gradient / expression / constraint-row code of trajopt, trajopt_common and trajopt_ifopt.
#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/trajopt_cuda.h"
#include "narrowphase.hpp"
#include "oracle.h"

#include <atomic>
#include <thread>

namespace orc
{
// ---------------------------------------------------------------------------------------------
// Scene
// ---------------------------------------------------------------------------------------------
struct Frame
{
  int parent;
  int type;
  int dof;
  V3 axis;
  Pose origin;
};

struct LinkPair
{
  int a, b;
  double margin, coeff;
  bool zero_coeff;
  std::vector<std::pair<int, int>> shape_pairs;
};

struct Scene
{
  int n_dof{ 0 };
  std::vector<Frame> frames;
  std::vector<std::vector<int>> frame_dofs;  // dof-carrying ancestor frames (incl. self) of each frame
  Pose base;
  int n_links{ 0 }, n_active{ 0 };
  int n_lead{ 0 };  // manager-active links: the group's + the environment's other moving links (dynamic environment)
  std::vector<int> link_frame;
  std::vector<Pose> link_pose;  // static world poses
  std::vector<Shape> shapes;
  std::vector<std::vector<int>> link_shapes;
  std::vector<LinkPair> pairs;
  double margin_buffer{ 0 };
  bool isActive(int link) const { return link < n_active; }
};

static std::string g_err;

static Scene* buildScene(const tc_scene_desc& d)
{
  auto s = std::make_unique<Scene>();
  s->n_dof = d.n_dof;
  s->frames.resize(static_cast<std::size_t>(d.n_frames));
  s->frame_dofs.resize(static_cast<std::size_t>(d.n_frames));
  for (int f = 0; f < d.n_frames; ++f)
  {
    Frame& fr = s->frames[static_cast<std::size_t>(f)];
    fr.parent = d.frame_parent[f];
    fr.type = d.frame_joint_type[f];
    fr.dof = d.frame_dof[f];
    fr.axis = V3(d.frame_axis[3 * f], d.frame_axis[3 * f + 1], d.frame_axis[3 * f + 2]);
    fr.origin = Pose::from12(d.frame_origin + 12 * f);
    if (fr.parent >= f)
      throw std::runtime_error("frames are not topologically ordered");
    if (fr.parent >= 0)
      s->frame_dofs[static_cast<std::size_t>(f)] = s->frame_dofs[static_cast<std::size_t>(fr.parent)];
    if (fr.type != TC_JOINT_FIXED)
    {
      if (fr.dof < 0 || fr.dof >= d.n_dof)
        throw std::runtime_error("movable frame without a dof index");
      s->frame_dofs[static_cast<std::size_t>(f)].push_back(f);
    }
  }
  s->base = Pose::from12(d.base_pose);
  s->n_links = d.n_links;
  s->n_active = d.n_active_links;
  if (d.n_env_active_links < 0 || d.n_active_links + d.n_env_active_links > d.n_links)
    throw std::runtime_error("n_env_active_links out of range");
  s->n_lead = d.n_active_links + d.n_env_active_links;
  s->link_frame.assign(d.link_frame, d.link_frame + d.n_links);
  for (int l = 0; l < d.n_links; ++l)
  {
    s->link_pose.push_back(Pose::from12(d.link_pose + 12 * l));
    if ((l < d.n_active_links) != (d.link_frame[l] >= 0))
      throw std::runtime_error("active links must come first and carry a frame");
  }
  s->link_shapes.resize(static_cast<std::size_t>(d.n_links));
  for (int i = 0; i < d.n_shapes; ++i)
  {
    Shape sh;
    sh.type = d.shape_type[i];
    sh.link = d.shape_link[i];
    sh.shape_id = d.shape_id[i];
    const Pose sp = Pose::from12(d.shape_pose + 12 * i);
    const double* prm = d.shape_params + 4 * i;
    sh.center = sp.t;
    sh.axis = sp.rot(V3(0, 0, 1));
    V3 lo(DBL_MAX, DBL_MAX, DBL_MAX), hi(-DBL_MAX, -DBL_MAX, -DBL_MAX);
    switch (sh.type)
    {
      case TC_SHAPE_SPHERE:
        sh.radius = prm[0];
        lo = sh.center - V3(prm[0], prm[0], prm[0]);
        hi = sh.center + V3(prm[0], prm[0], prm[0]);
        break;
      case TC_SHAPE_CAPSULE:
      case TC_SHAPE_CYLINDER:
      case TC_SHAPE_CONE:
      {
        sh.radius = prm[0];
        sh.half_len = prm[1];
        // conservative box: segment extent + radius in every direction
        for (int k = 0; k < 3; ++k)
        {
          const double e = std::fabs(sh.axis[k]) * sh.half_len + sh.radius;
          lo[k] = sh.center[k] - e;
          hi[k] = sh.center[k] + e;
        }
        break;
      }
      case TC_SHAPE_HULL:
      {
        if (d.shape_vert_count[i] <= 0)
          throw std::runtime_error("hull without vertices");
        V3 c;
        for (int k = 0; k < d.shape_vert_count[i]; ++k)
        {
          const double* p = d.verts + 3 * (d.shape_vert_offset[i] + k);
          const V3 v = sp.apply(V3(p[0], p[1], p[2]));
          sh.verts.push_back(v);
          c += v;
          for (int a = 0; a < 3; ++a)
          {
            lo[a] = std::min(lo[a], v[a]);
            hi[a] = std::max(hi[a], v[a]);
          }
        }
        sh.center = (1.0 / d.shape_vert_count[i]) * c;
        break;
      }
      default:
        throw std::runtime_error("unsupported shape type");
    }
    switch (sh.type)
    {
      case TC_SHAPE_SPHERE:
        sh.brad = 0;
        break;
      case TC_SHAPE_CAPSULE:
        sh.brad = sh.half_len;
        break;
      case TC_SHAPE_CYLINDER:
      case TC_SHAPE_CONE:
        sh.brad = std::sqrt(sh.half_len * sh.half_len + sh.radius * sh.radius);
        break;
      default:
        for (const V3& v : sh.verts)
          sh.brad = std::max(sh.brad, norm(v - sh.center));
    }
    sh.aabb_c = 0.5 * (lo + hi);
    sh.aabb_h = 0.5 * (hi - lo);
    s->link_shapes[static_cast<std::size_t>(sh.link)].push_back(i);
    s->shapes.push_back(sh);
  }
  s->margin_buffer = d.margin_buffer;
  for (int a = 0; a < d.n_links; ++a)
    for (int b = a + 1; b < d.n_links; ++b)
    {
      if (!(a < s->n_lead || b < s->n_lead))  // contactTest reports pairs with at least one manager-active link
        continue;
      if (d.acm && (d.acm[a * d.n_links + b] || d.acm[b * d.n_links + a]))
        continue;
      if (s->link_shapes[static_cast<std::size_t>(a)].empty() || s->link_shapes[static_cast<std::size_t>(b)].empty())
        continue;
      LinkPair p;
      p.a = a;
      p.b = b;
      p.margin = d.pair_margin ? d.pair_margin[a * d.n_links + b] : d.default_margin;
      p.coeff = d.pair_coeff ? d.pair_coeff[a * d.n_links + b] : d.default_coeff;
      // CollisionCoeffData::setCollisionCoeff: almostEqualRelativeAndAbs(coeff, 0)  (collision_types.cpp:56)
      p.zero_coeff = std::fabs(p.coeff) <= 1e-6;
      for (int sa : s->link_shapes[static_cast<std::size_t>(a)])
        for (int sb : s->link_shapes[static_cast<std::size_t>(b)])
          p.shape_pairs.emplace_back(sa, sb);
      s->pairs.push_back(p);
    }
  return s.release();
}

// ---------------------------------------------------------------------------------------------
// Kinematics — stands in for tesseract::kinematics::JointGroup::{calcFwdKin, calcJacobian} and
// tesseract::common::jacobianChangeRefPoint [upstream, recalled]; call contract at
// trajopt/src/collision_terms.cpp:209-223 and trajopt_common/src/collision_utils.cpp:128-151.
// ---------------------------------------------------------------------------------------------
struct FkState
{
  std::vector<Pose> frame;  // world pose of every frame after its joint motion
  std::vector<V3> axis_w;   // world joint axis per frame
};

static void fk(const Scene& s, const double* q, FkState& out)
{
  const std::size_t F = s.frames.size();
  out.frame.resize(F);
  out.axis_w.resize(F);
  for (std::size_t f = 0; f < F; ++f)
  {
    const Frame& fr = s.frames[f];
    const Pose& parent = (fr.parent < 0) ? s.base : out.frame[static_cast<std::size_t>(fr.parent)];
    Pose T = parent * fr.origin;
    out.axis_w[f] = T.rot(fr.axis);
    if (fr.type == TC_JOINT_REVOLUTE)
      T = T * axisAngle(fr.axis, q[fr.dof]);
    else if (fr.type == TC_JOINT_PRISMATIC)
      T.t = T.t + q[fr.dof] * out.axis_w[f];
    out.frame[f] = T;
  }
}

static Pose linkPose(const Scene& s, const FkState& st, int link)
{
  return s.isActive(link) ? st.frame[static_cast<std::size_t>(s.link_frame[static_cast<std::size_t>(link)])] :
                            s.link_pose[static_cast<std::size_t>(link)];
}

// Translational rows of the geometric Jacobian of `link`, reference point = link origin + ref_offset
// (world frame): column j = z_j x (p - p_j) for revolute, z_j for prismatic, 0 if joint j does not move
// the link.  J is 3 x n_dof row-major.
static void jacobianTrans(const Scene& s, const FkState& st, int link, V3 ref_offset, double* J)
{
  const int n = s.n_dof;
  std::fill(J, J + 3 * n, 0.0);
  const int lf = s.link_frame[static_cast<std::size_t>(link)];
  const V3 p = st.frame[static_cast<std::size_t>(lf)].t + ref_offset;
  for (int f : s.frame_dofs[static_cast<std::size_t>(lf)])
  {
    const Frame& fr = s.frames[static_cast<std::size_t>(f)];
    const V3 z = st.axis_w[static_cast<std::size_t>(f)];
    const V3 col = (fr.type == TC_JOINT_REVOLUTE) ? cross(z, p - st.frame[static_cast<std::size_t>(f)].t) : z;
    J[0 * n + fr.dof] = col.x;
    J[1 * n + fr.dof] = col.y;
    J[2 * n + fr.dof] = col.z;
  }
}

// ---------------------------------------------------------------------------------------------
// Contact results — the fields of tesseract::collision::ContactResult the path reads
// ---------------------------------------------------------------------------------------------
struct Contact
{
  tc_contact c{};
  Pose transform[2];
  Pose cc_transform[2];
};
using ContactResultMap = std::map<int, std::vector<Contact>>;  // key = link-pair index (sorted LinkIdPair order)

struct Stats
{
  long candidates{ 0 }, narrow{ 0 }, gjk_iters{ 0 }, scanned{ 0 }, epa{ 0 }, unconverged{ 0 }, epa_iters{ 0 };
};

// tesseract_collision bullet_utils GetAverageSupport [upstream, recalled]: for polyhedral shapes the
// support point is the mean of all vertices within BULLET_EPSILON of the running maximum.
constexpr double BULLET_EPSILON = 1e-3;
constexpr double BULLET_SUPPORT_FUNC_TOLERANCE = 0.01;
constexpr double BULLET_LENGTH_TOLERANCE = 0.001;

static void averageSupport(const Shape& sh, V3 nl, double& outsupport, V3& outpt)
{
  if (sh.type == HULL)
  {
    V3 sum;
    double cnt = 0, max_support = -1000;
    for (const V3& pt : sh.verts)
    {
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
        sum += pt;
      }
    }
    outsupport = max_support;
    outpt = (1.0 / cnt) * sum;
  }
  else
  {
    // localGetSupportingVertex: core support plus the radius shell along the direction
    const double nn = norm(nl);
    outpt = supportLocal(sh, nl);
    if (nn > 0)
      outpt += (sh.coreRadius() / nn) * nl;
    outsupport = dot(nl, outpt);
  }
}

// tesseract_collision bullet_utils calculateContinuousData [upstream, recalled]
static void calculateContinuousData(Contact& col, const Shape& sh, const Pose& tf0, const Pose& tf1, V3 pt_world,
                                    V3 normal_world, int i)
{
  col.cc_transform[i] = tf1;
  const V3 nl0 = tf0.rotT(normal_world), nl1 = tf1.rotT(normal_world);
  double s0, s1;
  V3 pl0, pl1;
  averageSupport(sh, nl0, s0, pl0);
  averageSupport(sh, nl1, s1, pl1);
  const V3 pw0 = tf0.apply(pl0), pw1 = tf1.apply(pl1);
  const double sup0 = dot(normal_world, pw0), sup1 = dot(normal_world, pw1);
  if (sup0 - sup1 > BULLET_SUPPORT_FUNC_TOLERANCE)
  {
    col.c.cc_time[i] = 0;
    col.c.cc_type[i] = TC_CC_TIME0;
  }
  else if (sup1 - sup0 > BULLET_SUPPORT_FUNC_TOLERANCE)
  {
    col.c.cc_time[i] = 1;
    col.c.cc_type[i] = TC_CC_TIME1;
  }
  else
  {
    const double l0c = norm(pt_world - pw0), l1c = norm(pt_world - pw1);
    col.c.cc_type[i] = TC_CC_BETWEEN;
    col.c.cc_time[i] = (l0c + l1c < BULLET_LENGTH_TOLERANCE) ? 0.5 : l0c / (l0c + l1c);
  }
}

// The contact manager: setCollisionObjectsTransform + contactTest (ContactTestType::ALL).
// poses1 == nullptr -> discrete manager; otherwise cast manager (active links swept pose0 -> pose1).
// `type`: tesseract::collision::ContactTestType as include/trajopt_cuda.h numbers it (TC_CONTACT_*): CLOSEST keeps,
// per link pair, the result with the smallest distance (processResult replaces the stored one only when the new one is
// strictly closer [upstream, recalled]; ties inside CLOSEST_TIE go to the first, see below); FIRST ends the whole test at the first result (cdata.done) — tesseract's
// visiting order is Bullet's broadphase order, here the (link pair, shape pair) order of the scene.
constexpr double CLOSEST_TIE = 1e-9;  // [m]
static void contactTest(const Scene& s, const std::vector<Pose>& poses0, const std::vector<Pose>* poses1,
                        ContactResultMap& out, Stats& st, int type = TC_CONTACT_ALL)
{
  const bool cast = poses1 != nullptr;
  auto worldAabb = [&](const Shape& sh, V3& lo, V3& hi) {
    auto one = [&](const Pose& T, V3& l, V3& h) {
      const V3 c = T.apply(sh.aabb_c);
      V3 e;
      for (int k = 0; k < 3; ++k)
        e[k] = std::fabs(T.R[k][0]) * sh.aabb_h.x + std::fabs(T.R[k][1]) * sh.aabb_h.y +
               std::fabs(T.R[k][2]) * sh.aabb_h.z;
      l = c - e;
      h = c + e;
    };
    one(poses0[static_cast<std::size_t>(sh.link)], lo, hi);
    if (cast && s.isActive(sh.link))
    {
      V3 l1, h1;
      one((*poses1)[static_cast<std::size_t>(sh.link)], l1, h1);
      for (int k = 0; k < 3; ++k)
      {
        lo[k] = std::min(lo[k], l1[k]);
        hi[k] = std::max(hi[k], h1[k]);
      }
    }
  };
  for (std::size_t pi = 0; pi < s.pairs.size(); ++pi)
  {
    const LinkPair& lp = s.pairs[pi];
    const double thr = lp.margin + s.margin_buffer;  // manager margin after incrementCollisionMargin(buffer)
    for (const auto& sp : lp.shape_pairs)
    {
      const Shape& A = s.shapes[static_cast<std::size_t>(sp.first)];
      const Shape& B = s.shapes[static_cast<std::size_t>(sp.second)];
      ++st.candidates;
      V3 alo, ahi, blo, bhi;
      worldAabb(A, alo, ahi);
      worldAabb(B, blo, bhi);
      bool overlap = true;
      for (int k = 0; k < 3; ++k)
        if (alo[k] - bhi[k] > thr || blo[k] - ahi[k] > thr)
          overlap = false;
      if (!overlap)
        continue;
      ++st.narrow;
      Placed PA, PB;
      PA.shape = &A;
      PB.shape = &B;
      PA.pose0 = poses0[static_cast<std::size_t>(A.link)];
      PB.pose0 = poses0[static_cast<std::size_t>(B.link)];
      if (cast && s.isActive(A.link))
      {
        PA.cast = true;
        PA.pose1 = (*poses1)[static_cast<std::size_t>(A.link)];
      }
      if (cast && s.isActive(B.link))
      {
        PB.cast = true;
        PB.pose1 = (*poses1)[static_cast<std::size_t>(B.link)];
      }
      const DistanceResult d = signedDistance(PA, PB, thr);
      st.gjk_iters += d.gjk_iters;
      st.scanned += d.scanned;
      st.epa += d.used_epa ? 1 : 0;
      st.epa_iters += d.epa_iters;
      st.unconverged += d.converged ? 0 : 1;
      if (!d.contact)
        continue;
      Contact c;
      c.c.distance = d.distance;
      const V3 p[2] = { d.pa, d.pb };
      const Shape* sh[2] = { &A, &B };
      for (int k = 0; k < 3; ++k)
        c.c.normal[k] = d.normal[k];
      for (int i = 0; i < 2; ++i)
      {
        c.transform[i] = poses0[static_cast<std::size_t>(sh[i]->link)];
        c.cc_transform[i] = Pose();  // ContactResult default (identity)
        const V3 pl = c.transform[i].applyInv(p[i]);
        for (int k = 0; k < 3; ++k)
        {
          c.c.nearest_points[i][k] = p[i][k];
          c.c.nearest_points_local[i][k] = pl[k];
        }
        c.c.cc_time[i] = -1;
        c.c.cc_type[i] = TC_CC_NONE;
        c.c.link_ids[i] = sh[i]->link;
        c.c.shape_id[i] = sh[i]->shape_id;
        c.c.subshape_id[i] = -1;
      }
      c.c.pair = static_cast<int>(pi);
      c.c.flags = (d.used_epa ? TC_FLAG_EPA : 0) | (d.converged ? 0 : TC_FLAG_UNCONVERGED);
      if (cast)
      {
        // addCastSingleResult: the cast shape's normal points away from itself
        if (PA.cast)
          calculateContinuousData(c, A, PA.pose0, PA.pose1, d.pa, d.normal, 0);
        if (PB.cast)
          calculateContinuousData(c, B, PB.pose0, PB.pose1, d.pb, -d.normal, 1);
      }
      out[static_cast<int>(pi)].push_back(c);
      if (type == TC_CONTACT_FIRST)
        return;
    }
    if (type == TC_CONTACT_CLOSEST)
    {
      // Sub-shapes of a compound link often share faces, so two shape pairs report the SAME distance and the strict
      // "replace when closer" of the manager is decided by the last bit.  Stated rule: the first result, in shape-pair
      // order, within CLOSEST_TIE of the smallest distance (identical to the manager's outside that band).
      auto it = out.find(static_cast<int>(pi));
      if (it != out.end() && it->second.size() > 1)
      {
        double dmin = it->second[0].c.distance;
        for (const Contact& c : it->second)
          dmin = std::min(dmin, c.c.distance);
        for (const Contact& c : it->second)
          if (c.c.distance <= dmin + CLOSEST_TIE)
          {
            const Contact keep = c;
            it->second.assign(1, keep);
            break;
          }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// Evaluators
// ---------------------------------------------------------------------------------------------
// Eigen::VectorXd::LinSpaced(cnt, lo, hi)(i) for floating point scalars [Eigen 3.4 linspaced_op_impl]
static double linspaced(long cnt, double lo, double hi, long i)
{
  const long size1 = (cnt == 1) ? 1 : cnt - 1;
  const double step = (cnt == 1) ? 0.0 : (hi - lo) / static_cast<double>(cnt - 1);
  const bool flip = std::fabs(hi) < std::fabs(lo);
  if (flip)
    return (i == 0) ? lo : (hi - static_cast<double>(size1 - i) * step);
  return (i == size1) ? hi : (lo + static_cast<double>(i) * step);
}

// trajopt_common::removeInvalidContactResults  (trajopt_common/src/collision_utils.cpp:73-114)
static void removeInvalidContactResults(std::vector<Contact>& v, double margin, double buffer, bool f0, bool f1)
{
  auto end = std::remove_if(v.begin(), v.end(), [=](const Contact& r) {
    if (r.c.distance > margin + buffer)
      return true;
    if (!f0 && !f1)
      return false;
    if (f0)
    {
      if (r.c.cc_type[0] != TC_CC_NONE && r.c.cc_type[0] != TC_CC_TIME0)
        return false;
      if (r.c.cc_type[1] != TC_CC_NONE && r.c.cc_type[1] != TC_CC_TIME0)
        return false;
    }
    if (f1)
    {
      if (r.c.cc_type[0] != TC_CC_NONE && r.c.cc_type[0] != TC_CC_TIME1)
        return false;
      if (r.c.cc_type[1] != TC_CC_NONE && r.c.cc_type[1] != TC_CC_TIME1)
        return false;
    }
    return true;
  });
  v.erase(end, v.end());
}

struct TermEval
{
  const Scene& s;
  double lvs;
  Stats stats;
  long units{ 0 }, states{ 0 };
  FkState st0, st1;
  std::vector<Pose> poses0, poses1;
  int type{ TC_CONTACT_ALL };
  explicit TermEval(const Scene& scene, double lvs_) : s(scene), lvs(lvs_)
  {
    poses0 = s.link_pose;
    poses1 = s.link_pose;
  }
  void setPoses(const double* q, FkState& st, std::vector<Pose>& poses)
  {
    fk(s, q, st);  // get_state_fn_ -> manip_->calcFwdKin
    for (int l = 0; l < s.n_active; ++l)
      poses[static_cast<std::size_t>(l)] = linkPose(s, st, l);
  }
  // filter lambda of the evaluators: zero coeff -> clear; then removeInvalidContactResults
  void filter(ContactResultMap& m, bool f0, bool f1, bool single) const
  {
    for (auto it = m.begin(); it != m.end();)
    {
      const LinkPair& lp = s.pairs[static_cast<std::size_t>(it->first)];
      if (lp.zero_coeff)
        it->second.clear();
      else if (single)
      {
        auto& v = it->second;
        v.erase(std::remove_if(v.begin(), v.end(),
                               [&](const Contact& r) { return r.c.distance > lp.margin + s.margin_buffer; }),
                v.end());
      }
      else
        removeInvalidContactResults(it->second, lp.margin, s.margin_buffer, f0, f1);
      it = it->second.empty() ? m.erase(it) : std::next(it);
    }
  }
  // tesseract ContactResultMap::addInterpolatedCollisionResults [upstream, recalled]
  void addInterpolated(ContactResultMap& dst, ContactResultMap& sub, long i, long last, double dt, bool discrete,
                       bool f0, bool f1, int state_index) const
  {
    for (auto& kv : sub)
      for (Contact& r : kv.second)
      {
        r.c.substep = static_cast<int>(i);
        r.c.state = state_index;
        for (int j = 0; j < 2; ++j)
        {
          if (!s.isActive(r.c.link_ids[j]))
            continue;
          r.c.cc_time[j] = (r.c.cc_time[j] < 0) ? (static_cast<double>(i) * dt) :
                                                   (static_cast<double>(i) * dt) + (r.c.cc_time[j] * dt);
          if (i == 0 && (r.c.cc_type[j] == TC_CC_TIME0 || r.c.cc_type[j] == TC_CC_NONE))
            r.c.cc_type[j] = TC_CC_TIME0;
          else if (i == last && (r.c.cc_type[j] == TC_CC_TIME1 || r.c.cc_type[j] == TC_CC_NONE))
            r.c.cc_type[j] = TC_CC_TIME1;
          else
            r.c.cc_type[j] = TC_CC_BETWEEN;
          if (discrete)
          {
            r.cc_transform[0] = r.transform[0];
            r.cc_transform[1] = r.transform[1];
          }
        }
      }
    filter(sub, f0, f1, false);
    for (auto& kv : sub)
    {
      auto& v = dst[kv.first];
      v.insert(v.end(), kv.second.begin(), kv.second.end());
    }
  }

  // SingleTimestepCollisionEvaluator::CalcCollisions (trajopt/src/collision_terms.cpp:653-688) ==
  // trajopt_ifopt SingleTimestepCollisionEvaluator::calcCollisionsHelper (discrete_collision_evaluators.cpp:151-189)
  void single(const double* q, ContactResultMap& out, int state_base)
  {
    setPoses(q, st0, poses0);
    contactTest(s, poses0, nullptr, out, stats, type);
    filter(out, false, false, true);
    for (auto& kv : out)
      for (Contact& c : kv.second)
        c.c.state = state_base;
    units += 1;
    states += 1;
  }

  // `longer` = dist > longest_valid_segment_length: the reference branches on THIS, not on cnt
  // (collision_terms.cpp:1107, continuous_collision_evaluators.cpp:204); when dist barely exceeds the length,
  // dist / length rounds to 1.0, cnt is 2 and the single sweep still goes through addInterpolatedCollisionResults
  long subCount(const double* q0, const double* q1, bool* longer = nullptr) const
  {
    double d2 = 0;
    for (int j = 0; j < s.n_dof; ++j)
      d2 += (q1[j] - q0[j]) * (q1[j] - q0[j]);
    const double dist = std::sqrt(d2);
    long cnt = 2;
    if (dist > lvs)
      cnt = static_cast<long>(std::ceil(dist / lvs)) + 1;
    if (longer)
      *longer = dist > lvs;
    return cnt;
  }
  void subState(const double* q0, const double* q1, long cnt, long i, std::vector<double>& q) const
  {
    q.resize(static_cast<std::size_t>(s.n_dof));
    for (int j = 0; j < s.n_dof; ++j)
      q[static_cast<std::size_t>(j)] = linspaced(cnt, q0[j], q1[j], i);
  }

  // DiscreteCollisionEvaluator::CalcCollisions (collision_terms.cpp:817-898) ==
  // LVSDiscreteCollisionEvaluator::calcCollisionsHelper (continuous_collision_evaluators.cpp:386-458)
  void lvsDiscrete(const double* q0, const double* q1, bool f0, bool f1, ContactResultMap& out, int state_base)
  {
    const long cnt = subCount(q0, q1);
    const long last = cnt - 1;
    const double dt = 1.0 / static_cast<double>(last);
    std::vector<double> q;
    for (long i = 0; i < cnt; ++i)
    {
      subState(q0, q1, cnt, i, q);
      setPoses(q.data(), st0, poses0);
      ContactResultMap contacts;
      contactTest(s, poses0, nullptr, contacts, stats, type);
      if (!contacts.empty())
        addInterpolated(out, contacts, i, last, dt, true, f0, f1, state_base + static_cast<int>(i));
    }
    units += cnt;
    states += cnt;
  }

  // CastCollisionEvaluator::CalcCollisions (collision_terms.cpp:1060-1161) /
  // LVSContinuousCollisionEvaluator::calcCollisionsHelper (continuous_collision_evaluators.cpp:166-248)
  void cast(const double* q0, const double* q1, bool f0, bool f1, bool single_sweep, ContactResultMap& out,
            int state_base)
  {
    bool longer = false;
    const long cnt = subCount(q0, q1, &longer);
    if (!single_sweep && longer)
    {
      const long last = cnt - 1;
      const double dt = 1.0 / static_cast<double>(last);
      std::vector<double> qa, qb;
      for (long i = 0; i < cnt - 1; ++i)
      {
        subState(q0, q1, cnt, i, qa);
        subState(q0, q1, cnt, i + 1, qb);
        setPoses(qa.data(), st0, poses0);
        setPoses(qb.data(), st1, poses1);
        ContactResultMap contacts;
        contactTest(s, poses0, &poses1, contacts, stats, type);
        if (!contacts.empty())
          addInterpolated(out, contacts, i, last, dt, false, f0, f1, state_base + static_cast<int>(i));
      }
      units += cnt - 1;
      states += cnt;
    }
    else
    {
      setPoses(q0, st0, poses0);
      setPoses(q1, st1, poses1);
      contactTest(s, poses0, &poses1, out, stats, type);
      filter(out, f0, f1, false);
      for (auto& kv : out)
        for (Contact& c : kv.second)
          c.c.state = state_base;
      units += 1;
      states += 2;
    }
  }
};

// ---------------------------------------------------------------------------------------------
// Gradients — trajopt_common::calcGradient / getGradient (collision_utils.cpp:116-221) and the legacy
// CollisionEvaluator::GetGradient single-state form used by CollisionsToDistanceExpressions
// (collision_terms.cpp:195-242,341-386)
// ---------------------------------------------------------------------------------------------
struct LinkGradient
{
  bool has{ false };
  double scale{ 1 };
  int cc_type{ TC_CC_NONE };
  std::vector<double> g;
};
struct GradientResults
{
  LinkGradient gradients[2], cc_gradients[2];
  double error{ 0 }, error_with_buffer{ 0 };
};

static void calcGradient(const Scene& s, GradientResults& res, int i, const double* dofvals, const Contact& c,
                         bool isTimestep1)
{
  LinkGradient& lg = isTimestep1 ? res.cc_gradients[i] : res.gradients[i];
  lg.has = true;
  FkState st;
  fk(s, dofvals, st);
  lg.scale = 1;
  Pose link_transform = c.transform[i];
  if (c.c.cc_type[i] != TC_CC_NONE)
  {
    lg.scale = isTimestep1 ? c.c.cc_time[i] : (1 - c.c.cc_time[i]);
    lg.cc_type = c.c.cc_type[i];
    link_transform = isTimestep1 ? c.cc_transform[i] : c.transform[i];
  }
  const V3 pl(c.c.nearest_points_local[i][0], c.c.nearest_points_local[i][1], c.c.nearest_points_local[i][2]);
  std::vector<double> J(static_cast<std::size_t>(3 * s.n_dof));
  jacobianTrans(s, st, c.c.link_ids[i], link_transform.rot(pl), J.data());
  V3 tv(c.c.normal[0], c.c.normal[1], c.c.normal[2]);
  if (i == 0)
    tv = -tv;
  lg.g.resize(static_cast<std::size_t>(s.n_dof));
  for (int j = 0; j < s.n_dof; ++j)
    lg.g[static_cast<std::size_t>(j)] = J[static_cast<std::size_t>(j)] * tv.x +
                                        J[static_cast<std::size_t>(s.n_dof + j)] * tv.y +
                                        J[static_cast<std::size_t>(2 * s.n_dof + j)] * tv.z;
}

// getGradient(results, dofvals, contact, margin, buffer, manip)  (collision_utils.cpp:175-190)
static void getGradient1(const Scene& s, GradientResults& res, const double* q, const Contact& c, double margin,
                         double buffer)
{
  res.error = margin - c.c.distance;
  res.error_with_buffer = margin + buffer - c.c.distance;
  for (int i = 0; i < 2; ++i)
    if (s.isActive(c.c.link_ids[i]))
      calcGradient(s, res, i, q, c, false);
}

// getGradient(results, dofvals0, dofvals1, contact, margin, buffer, manip)  (collision_utils.cpp:192-221)
static void getGradient2(const Scene& s, GradientResults& res, const double* q0, const double* q1, const Contact& c,
                         double margin, double buffer)
{
  res.error = margin - c.c.distance;
  res.error_with_buffer = margin + buffer - c.c.distance;
  std::vector<double> qt(static_cast<std::size_t>(s.n_dof));
  for (int i = 0; i < 2; ++i)
  {
    if (!s.isActive(c.c.link_ids[i]))
      continue;
    for (int j = 0; j < s.n_dof; ++j)
    {
      if (c.c.cc_type[i] == TC_CC_TIME0)
        qt[static_cast<std::size_t>(j)] = q0[j];
      else if (c.c.cc_type[i] == TC_CC_TIME1)
        qt[static_cast<std::size_t>(j)] = q1[j];
      else
        qt[static_cast<std::size_t>(j)] = q0[j] + (q1[j] - q0[j]) * c.c.cc_time[i];
    }
    calcGradient(s, res, i, qt.data(), c, false);
    calcGradient(s, res, i, qt.data(), c, true);
  }
}

// legacy: GetGradient(dofvals, contact, margin, coeff, isTimestep1) evaluated once per free timestep
// block with that block's own joint values (collision_terms.cpp:349,369)
static void getGradientLegacy(const Scene& s, GradientResults& res, const double* q0, const double* q1,
                              const Contact& c, double margin, double buffer, bool want0, bool want1)
{
  res.error = margin - c.c.distance;
  res.error_with_buffer = margin + buffer - c.c.distance;
  for (int i = 0; i < 2; ++i)
  {
    if (!s.isActive(c.c.link_ids[i]))
      continue;
    if (want0)
      calcGradient(s, res, i, q0, c, false);
    if (want1)
      calcGradient(s, res, i, q1, c, true);
  }
}

// ---------------------------------------------------------------------------------------------
// Result of one batched evaluation
// ---------------------------------------------------------------------------------------------
struct Result
{
  const Scene* scene{ nullptr };
  tc_eval_desc desc{};
  int B{ 0 }, T{ 0 }, terms_per_traj{ 0 };
  std::vector<uint8_t> fixed;
  std::vector<double> q;
  std::vector<Contact> contacts;
  std::vector<GradientResults> grads;
  std::vector<int64_t> term_ptr;
  std::vector<int> term_t;                    // timestep / segment index of each term (per trajectory)
  std::vector<std::array<bool, 2>> term_fixed;  // vars0_fixed, vars1_fixed
  tc_counts counts{};
};

static Result* evaluate(const Scene& s, const tc_eval_desc& d, const double* q, int B, int T, int n_threads)
{
  auto r = std::make_unique<Result>();
  r->scene = &s;
  r->desc = d;
  r->B = B;
  r->T = T;
  r->fixed.assign(static_cast<std::size_t>(T), 0);
  if (d.fixed_steps)
    r->fixed.assign(d.fixed_steps, d.fixed_steps + T);
  r->desc.fixed_steps = nullptr;
  r->q.assign(q, q + static_cast<std::size_t>(B) * static_cast<std::size_t>(T) * static_cast<std::size_t>(s.n_dof));
  const int n = s.n_dof;
  // term list per trajectory: CollisionTermInfo::hatch (trajopt/src/problem_description.cpp:1735-1858)
  if (d.evaluator == TC_EVAL_DISCRETE)
  {
    for (int t = 0; t < T; ++t)
      if (!r->fixed[static_cast<std::size_t>(t)])
      {
        r->term_t.push_back(t);
        r->term_fixed.push_back({ false, false });
      }
  }
  else
  {
    for (int t = 0; t + 1 < T; ++t)
    {
      const bool cur = r->fixed[static_cast<std::size_t>(t)] != 0, nxt = r->fixed[static_cast<std::size_t>(t + 1)] != 0;
      if (cur && nxt)
        continue;  // both ends fixed: nothing to optimise (ifopt constraints throw, legacy would hatch START_FIXED)
      r->term_t.push_back(t);
      r->term_fixed.push_back({ cur, !cur && nxt });
    }
  }
  r->terms_per_traj = static_cast<int>(r->term_t.size());
  const long n_terms = static_cast<long>(B) * r->terms_per_traj;
  std::vector<std::vector<Contact>> per_term(static_cast<std::size_t>(n_terms));
  std::vector<std::vector<GradientResults>> per_term_g(static_cast<std::size_t>(n_terms));
  std::vector<Stats> stats(static_cast<std::size_t>(n_terms));
  std::vector<long> units(static_cast<std::size_t>(n_terms)), states(static_cast<std::size_t>(n_terms));
  const bool want_grad = (d.want & TC_WANT_GRADIENTS) != 0;
  const bool single_sweep = (d.evaluator == TC_EVAL_CONTINUOUS) && (d.stack == TC_STACK_IFOPT);
  // One evaluator per term; terms are distributed over host threads with dynamic scheduling, the
  // structure of BasicTrustRegionSQPMultiThreaded (trajopt_sco/src/optimizers.cpp:286-298:
  // "#pragma omp parallel for schedule(dynamic)" over cost terms).  std::thread instead of OpenMP so the
  // library has no runtime dependency.
  auto run_term = [&](long k) {
    const int b = static_cast<int>(k / r->terms_per_traj), ti = static_cast<int>(k % r->terms_per_traj);
    const int t = r->term_t[static_cast<std::size_t>(ti)];
    const bool f0 = r->term_fixed[static_cast<std::size_t>(ti)][0], f1 = r->term_fixed[static_cast<std::size_t>(ti)][1];
    const double* q0 = q + (static_cast<std::size_t>(b) * static_cast<std::size_t>(T) + static_cast<std::size_t>(t)) *
                               static_cast<std::size_t>(n);
    const double* q1 = q0 + n;
    TermEval ev(s, d.lvs_length);
    ev.type = d.contact_test_type;
    ContactResultMap m;
    if (d.evaluator == TC_EVAL_DISCRETE)
      ev.single(q0, m, 0);
    else if (d.evaluator == TC_EVAL_LVS_DISCRETE)
      ev.lvsDiscrete(q0, q1, f0, f1, m, 0);
    else
      ev.cast(q0, q1, f0, f1, single_sweep, m, 0);
    // flattenWrapperResults: map order, then append order
    auto& out = per_term[static_cast<std::size_t>(k)];
    for (auto& kv : m)
      for (Contact& c : kv.second)
      {
        c.c.traj = b;
        c.c.term = t;
        out.push_back(c);
      }
    if (want_grad)
    {
      auto& og = per_term_g[static_cast<std::size_t>(k)];
      og.resize(out.size());
      for (std::size_t i = 0; i < out.size(); ++i)
      {
        const LinkPair& lp = s.pairs[static_cast<std::size_t>(out[i].c.pair)];
        if (d.stack == TC_STACK_LEGACY)
        {
          const bool single = d.evaluator == TC_EVAL_DISCRETE;
          getGradientLegacy(s, og[i], q0, single ? q0 : q1, out[i], lp.margin, s.margin_buffer, single || !f0,
                            !single && !f1);
        }
        else if (d.evaluator == TC_EVAL_DISCRETE)
          getGradient1(s, og[i], q0, out[i], lp.margin, s.margin_buffer);
        else
          getGradient2(s, og[i], q0, q1, out[i], lp.margin, s.margin_buffer);
      }
    }
    stats[static_cast<std::size_t>(k)] = ev.stats;
    units[static_cast<std::size_t>(k)] = ev.units;
    states[static_cast<std::size_t>(k)] = ev.states;
  };
  const int nt = std::max(1, std::min<int>(n_threads, static_cast<int>(std::min<long>(n_terms, 1024))));
  if (nt <= 1)
  {
    for (long k = 0; k < n_terms; ++k)
      run_term(k);
  }
  else
  {
    std::atomic<long> next{ 0 };
    std::vector<std::thread> pool;
    for (int w = 0; w < nt; ++w)
      pool.emplace_back([&]() {
        for (long k = next.fetch_add(1); k < n_terms; k = next.fetch_add(1))
          run_term(k);
      });
    for (auto& th : pool)
      th.join();
  }
  r->term_ptr.assign(static_cast<std::size_t>(n_terms) + 1, 0);
  long shape_pairs = 0;
  for (const LinkPair& lp : s.pairs)
    shape_pairs += static_cast<long>(lp.shape_pairs.size());
  (void)shape_pairs;
  for (long k = 0; k < n_terms; ++k)
  {
    // contacts carry the term-local sub-state index; make it global (states are numbered term by term)
    for (Contact& c : per_term[static_cast<std::size_t>(k)])
      c.c.state += static_cast<int>(r->counts.n_states);
    r->term_ptr[static_cast<std::size_t>(k) + 1] =
        r->term_ptr[static_cast<std::size_t>(k)] + static_cast<int64_t>(per_term[static_cast<std::size_t>(k)].size());
    r->contacts.insert(r->contacts.end(), per_term[static_cast<std::size_t>(k)].begin(),
                       per_term[static_cast<std::size_t>(k)].end());
    if (want_grad)
      r->grads.insert(r->grads.end(), std::make_move_iterator(per_term_g[static_cast<std::size_t>(k)].begin()),
                      std::make_move_iterator(per_term_g[static_cast<std::size_t>(k)].end()));
    const Stats& st = stats[static_cast<std::size_t>(k)];
    r->counts.n_states += states[static_cast<std::size_t>(k)];
    r->counts.n_units += units[static_cast<std::size_t>(k)];
    r->counts.n_candidates += st.candidates;
    r->counts.n_narrowphase += st.narrow;
    r->counts.gjk_iters += st.gjk_iters;
    r->counts.support_verts += st.scanned;
    r->counts.support_compared += st.scanned;  // the plain scan compares every vertex
    r->counts.n_epa += st.epa;
    r->counts.epa_iters += st.epa_iters;
    r->counts.n_unconverged += st.unconverged;
  }
  r->counts.n_dense = r->counts.n_units * static_cast<int64_t>(s.pairs.size());
  r->counts.n_contacts = static_cast<int64_t>(r->contacts.size());
  r->counts.n_rows = r->counts.n_contacts;
  r->counts.n_terms = n_terms;
  return r.release();
}

// The contact manager driven directly (seam B3): poses in, contacts out, no kinematics and no evaluator filter beyond
// the manager's own contact threshold.  pose0 / pose1: [n_states][n_active][12]; pose1 == nullptr -> discrete manager.
static Result* contactTestStates(const Scene& s, const double* pose0, const double* pose1, int n_states, int type)
{
  auto r = std::make_unique<Result>();
  r->scene = &s;
  r->desc.evaluator = pose1 ? TC_EVAL_CONTINUOUS : TC_EVAL_DISCRETE;
  r->desc.stack = TC_STACK_IFOPT;
  r->desc.want = TC_WANT_CONTACTS | TC_WANT_COST;
  r->desc.contact_test_type = type;
  r->B = n_states;
  r->T = pose1 ? 2 : 1;
  r->terms_per_traj = 1;
  r->term_t.push_back(0);
  r->term_fixed.push_back({ false, false });
  r->term_ptr.assign(static_cast<std::size_t>(n_states) + 1, 0);
  std::vector<Pose> p0 = s.link_pose, p1 = s.link_pose;
  const int spt = pose1 ? 2 : 1;
  for (int k = 0; k < n_states; ++k)
  {
    for (int l = 0; l < s.n_active; ++l)
    {
      p0[static_cast<std::size_t>(l)] = Pose::from12(pose0 + (static_cast<std::size_t>(k) * s.n_active + l) * 12);
      if (pose1)
        p1[static_cast<std::size_t>(l)] = Pose::from12(pose1 + (static_cast<std::size_t>(k) * s.n_active + l) * 12);
    }
    ContactResultMap m;
    Stats st;
    contactTest(s, p0, pose1 ? &p1 : nullptr, m, st, type);
    for (auto& kv : m)
    {
      if (s.pairs[static_cast<std::size_t>(kv.first)].zero_coeff)
        continue;
      for (Contact& c : kv.second)
      {
        c.c.traj = k;
        c.c.term = 0;
        c.c.substep = 0;
        c.c.state = spt * k;
        if (pose1)
          for (int i = 0; i < 2; ++i)
            if (s.isActive(c.c.link_ids[i]))
              c.cc_transform[i] = p1[static_cast<std::size_t>(c.c.link_ids[i])];
        r->contacts.push_back(c);
      }
    }
    r->term_ptr[static_cast<std::size_t>(k) + 1] = static_cast<int64_t>(r->contacts.size());
    r->counts.n_candidates += st.candidates;
    r->counts.n_narrowphase += st.narrow;
    r->counts.gjk_iters += st.gjk_iters;
    r->counts.support_verts += st.scanned;
    r->counts.support_compared += st.scanned;
    r->counts.n_epa += st.epa;
    r->counts.epa_iters += st.epa_iters;
    r->counts.n_unconverged += st.unconverged;
  }
  r->counts.n_states = static_cast<int64_t>(spt) * n_states;
  r->counts.n_units = n_states;
  r->counts.n_dense = r->counts.n_units * static_cast<int64_t>(s.pairs.size());
  r->counts.n_contacts = static_cast<int64_t>(r->contacts.size());
  r->counts.n_rows = r->counts.n_contacts;
  r->counts.n_terms = n_states;
  return r.release();
}

// ---------------------------------------------------------------------------------------------
// GradientResultsSet and the fixed-size constraint rows
// (trajopt_common/src/collision_types.cpp:173-333; weighted_average_methods.cpp:31-107;
//  discrete_collision_constraint.cpp:68-141; continuous_collision_constraint.cpp:80-236)
// ---------------------------------------------------------------------------------------------
constexpr double LOWEST = std::numeric_limits<double>::lowest();
struct LinkMaxError
{
  bool has_error[2]{ false, false };
  double error[2]{ LOWEST, LOWEST };
  double error_with_buffer[2]{ LOWEST, LOWEST };
};
struct GradientResultsSet
{
  int pair{ 0 };
  std::pair<std::size_t, std::size_t> shape_key;
  double coeff{ 1 };
  bool is_continuous{ false };
  LinkMaxError max_error[2];
  std::vector<const GradientResults*> results;
  void add(const GradientResults& g)
  {
    for (int i = 0; i < 2; ++i)
    {
      const LinkGradient& lg = g.gradients[i];
      if (!lg.has)
        continue;
      if (lg.cc_type != TC_CC_TIME1)
      {
        max_error[i].has_error[0] = true;
        max_error[i].error[0] = std::max(max_error[i].error[0], g.error);
        max_error[i].error_with_buffer[0] = std::max(max_error[i].error_with_buffer[0], g.error_with_buffer);
      }
      if (lg.cc_type != TC_CC_TIME0)
      {
        max_error[i].has_error[1] = true;
        max_error[i].error[1] = std::max(max_error[i].error[1], g.error);
        max_error[i].error_with_buffer[1] = std::max(max_error[i].error_with_buffer[1], g.error_with_buffer);
      }
    }
    results.push_back(&g);
  }
  // which: 0 = T0 only, 1 = T1 only, 2 = both; buffer selects error_with_buffer
  bool maxError(int which, bool buffer, double& e) const
  {
    e = LOWEST;
    bool found = false;
    for (int i = 0; i < 2; ++i)
      for (int k = 0; k < 2; ++k)
      {
        if (which != 2 && which != k)
          continue;
        const double v = buffer ? max_error[i].error_with_buffer[k] : max_error[i].error[k];
        if (max_error[i].has_error[k] && v > e)
        {
          e = v;
          found = true;
        }
      }
    return found;
  }
  double get(int which, bool buffer) const
  {
    double e;
    if (!maxError(which, buffer, e))
      throw std::runtime_error("Max error does not exist.");
    return e;
  }
};

static std::size_t cantorHash(int shape_id, int subshape_id)
{
  if (subshape_id < 0)
    return static_cast<std::size_t>(shape_id);
  return static_cast<std::size_t>(1 / 2.0 * (shape_id + subshape_id) * (shape_id + subshape_id + 1) + subshape_id);
}

static std::vector<double> weightedAvg(const GradientResultsSet& set, double max_error_with_buffer, int n, bool t1)
{
  std::vector<double> grad(static_cast<std::size_t>(n), 0.0);
  if (set.results.empty())
    return grad;
  double total_weight = 0;
  long cnt = 0;
  for (const GradientResults* g : set.results)
    for (int i = 0; i < 2; ++i)
    {
      const LinkGradient& lg = t1 ? g->cc_gradients[i] : g->gradients[i];
      if (lg.has && lg.cc_type != (t1 ? TC_CC_TIME0 : TC_CC_TIME1))
      {
        if (set.max_error[i].error_with_buffer[t1 ? 1 : 0] > 0)
        {
          const double w = std::max(g->error_with_buffer, 0.0) / max_error_with_buffer;
          total_weight += w;
          for (int j = 0; j < n; ++j)
            grad[static_cast<std::size_t>(j)] += w * (lg.scale * lg.g[static_cast<std::size_t>(j)]);
          ++cnt;
        }
      }
    }
  if (cnt == 0)
    return grad;
  const double scale = 1 / total_weight;
  for (double& v : grad)
    v *= scale;
  return grad;
}

static std::vector<GradientResultsSet> buildSets(const Result& r, long term, std::size_t max_allowed)
{
  const Scene& s = *r.scene;
  const bool continuous = r.desc.evaluator != TC_EVAL_DISCRETE;
  const std::array<bool, 2> fx = r.term_fixed[static_cast<std::size_t>(term % r.terms_per_traj)];
  std::vector<GradientResultsSet> sets;
  const int64_t lo = r.term_ptr[static_cast<std::size_t>(term)], hi = r.term_ptr[static_cast<std::size_t>(term) + 1];
  int64_t i = lo;
  while (i < hi)
  {
    const int pair = r.contacts[static_cast<std::size_t>(i)].c.pair;
    std::map<std::pair<std::size_t, std::size_t>, GradientResultsSet> shape_grs;
    for (; i < hi && r.contacts[static_cast<std::size_t>(i)].c.pair == pair; ++i)
    {
      const tc_contact& c = r.contacts[static_cast<std::size_t>(i)].c;
      const auto key = std::make_pair(cantorHash(c.shape_id[0], c.subshape_id[0]), cantorHash(c.shape_id[1], c.subshape_id[1]));
      auto ins = shape_grs.try_emplace(key);
      GradientResultsSet& grs = ins.first->second;
      if (ins.second)
      {
        grs.pair = pair;
        grs.shape_key = key;
        grs.coeff = s.pairs[static_cast<std::size_t>(pair)].coeff;
        grs.is_continuous = continuous;
      }
      grs.add(r.grads[static_cast<std::size_t>(i)]);
    }
    for (auto& kv : shape_grs)
      sets.push_back(std::move(kv.second));
  }
  if (sets.size() > max_allowed)
  {
    if (!continuous)
    {
      // comparator of discrete_collision_evaluators.cpp:143-147: std::array<double,2> operator> on max_error[0].error
      std::stable_sort(sets.begin(), sets.end(), [](const GradientResultsSet& a, const GradientResultsSet& b) {
        return std::lexicographical_compare(b.max_error[0].error, b.max_error[0].error + 2, a.max_error[0].error,
                                            a.max_error[0].error + 2);
      });
    }
    else
    {
      const int which = (!fx[0] && !fx[1]) ? 2 : (!fx[0] ? 0 : 1);
      std::stable_sort(sets.begin(), sets.end(), [which](const GradientResultsSet& a, const GradientResultsSet& b) {
        return a.get(which, true) > b.get(which, true);
      });
    }
  }
  return sets;
}

}  // namespace orc

// =================================================================================================
// C API
// =================================================================================================
using namespace orc;

#define ORC_TRY try {
#define ORC_CATCH                                                                                                    \
  }                                                                                                                  \
  catch (const std::exception& e)                                                                                    \
  {                                                                                                                  \
    g_err = e.what();                                                                                                \
    return TC_ERR_INVALID;                                                                                           \
  }                                                                                                                  \
  return TC_OK;

extern "C" {
const char* orc_last_error(void) { return g_err.c_str(); }

int orc_scene_create(const tc_scene_desc* desc, void** out)
{
  ORC_TRY
  *out = buildScene(*desc);
  ORC_CATCH
}
void orc_scene_destroy(void* scene) { delete static_cast<Scene*>(scene); }
int orc_scene_num_pairs(const void* scene) { return static_cast<int>(static_cast<const Scene*>(scene)->pairs.size()); }
int orc_scene_pairs(const void* scene, int32_t* l0, int32_t* l1, double* margin, double* coeff)
{
  const Scene& s = *static_cast<const Scene*>(scene);
  for (std::size_t i = 0; i < s.pairs.size(); ++i)
  {
    l0[i] = s.pairs[i].a;
    l1[i] = s.pairs[i].b;
    margin[i] = s.pairs[i].margin;
    coeff[i] = s.pairs[i].coeff;
  }
  return TC_OK;
}

void orc_sincos(double x, double* sn, double* cs) { orc::sincosShared(x, *sn, *cs); }

int orc_fk(const void* scene, const double* q, double* link_poses /*[n_active][12]*/)
{
  const Scene& s = *static_cast<const Scene*>(scene);
  FkState st;
  fk(s, q, st);
  for (int l = 0; l < s.n_active; ++l)
    linkPose(s, st, l).to12(link_poses + 12 * l);
  return TC_OK;
}

int orc_jacobian(const void* scene, const double* q, int link, const double* ref_offset, double* J /*[3][n_dof]*/)
{
  const Scene& s = *static_cast<const Scene*>(scene);
  if (!s.isActive(link))
    return TC_ERR_INVALID;
  FkState st;
  fk(s, q, st);
  jacobianTrans(s, st, link, V3(ref_offset[0], ref_offset[1], ref_offset[2]), J);
  return TC_OK;
}

// signed distance of two scene shapes at given LINK poses; pose*1 may be NULL (not swept)
int orc_shape_distance(const void* scene, int shape_a, const double* pose_a0, const double* pose_a1, int shape_b,
                       const double* pose_b0, const double* pose_b1, double threshold, int* contact, double* distance,
                       double* normal, double* pa, double* pb, int* used_epa, int* iters)
{
  const Scene& s = *static_cast<const Scene*>(scene);
  Placed A, B;
  A.shape = &s.shapes[static_cast<std::size_t>(shape_a)];
  B.shape = &s.shapes[static_cast<std::size_t>(shape_b)];
  A.pose0 = Pose::from12(pose_a0);
  B.pose0 = Pose::from12(pose_b0);
  if (pose_a1)
  {
    A.cast = true;
    A.pose1 = Pose::from12(pose_a1);
  }
  if (pose_b1)
  {
    B.cast = true;
    B.pose1 = Pose::from12(pose_b1);
  }
  const DistanceResult d = signedDistance(A, B, threshold);
  *contact = d.contact;
  *distance = d.distance;
  for (int k = 0; k < 3; ++k)
  {
    normal[k] = d.normal[k];
    pa[k] = d.pa[k];
    pb[k] = d.pb[k];
  }
  if (used_epa)
    *used_epa = d.used_epa;
  if (iters)
    *iters = d.gjk_iters;
  return TC_OK;
}

// largest support value of a scene shape along world direction `dir` at link pose(s): max_x dir.x over
// the (swept) solid.  Used by tests to validate witness points independently of GJK.
int orc_shape_support(const void* scene, int shape, const double* pose0, const double* pose1, const double* dir,
                      double* value)
{
  const Scene& s = *static_cast<const Scene*>(scene);
  Placed P;
  P.shape = &s.shapes[static_cast<std::size_t>(shape)];
  P.pose0 = Pose::from12(pose0);
  if (pose1)
  {
    P.cast = true;
    P.pose1 = Pose::from12(pose1);
  }
  const V3 d(dir[0], dir[1], dir[2]);
  *value = dot(d, P.support(d)) + P.shape->coreRadius() * norm(d);
  return TC_OK;
}

int orc_eval(void* scene, const tc_eval_desc* desc, const double* q, int32_t B, int32_t T, void** out, int n_threads)
{
  ORC_TRY
  if (desc->evaluator < TC_EVAL_DISCRETE || desc->evaluator > TC_EVAL_LVS_CONTINUOUS)
    throw std::runtime_error("invalid evaluator type");
  if (desc->evaluator != TC_EVAL_DISCRETE && T < 2)
    throw std::runtime_error("segment evaluators need at least two timesteps");
  *out = evaluate(*static_cast<Scene*>(scene), *desc, q, B, T, n_threads);
  ORC_CATCH
}
int orc_contact_test(void* scene, const double* pose0, const double* pose1, int32_t n_states, int32_t type, void** out)
{
  if (!scene || !pose0 || !out || n_states <= 0)
  {
    g_err = "bad argument";
    return TC_ERR_INVALID;
  }
  *out = contactTestStates(*static_cast<Scene*>(scene), pose0, pose1, n_states, type);
  return TC_OK;
}

void orc_result_destroy(void* res) { delete static_cast<Result*>(res); }

int orc_result_counts(const void* res, tc_counts* c)
{
  const Result& r = *static_cast<const Result*>(res);
  *c = r.counts;
  const bool two = r.desc.evaluator != TC_EVAL_DISCRETE && r.desc.stack == TC_STACK_IFOPT;
  (void)two;
  return TC_OK;
}
int orc_result_contacts(const void* res, tc_contact* dst, size_t cap)
{
  const Result& r = *static_cast<const Result*>(res);
  for (std::size_t i = 0; i < std::min(cap, r.contacts.size()); ++i)
    dst[i] = r.contacts[i].c;
  return TC_OK;
}
int orc_result_contact_transforms(const void* res, double* dst, size_t cap)
{
  const Result& r = *static_cast<const Result*>(res);
  for (std::size_t i = 0; i < std::min(cap, r.contacts.size()); ++i)
    for (int l = 0; l < 2; ++l)
    {
      r.contacts[i].transform[l].to12(dst + ((i * 2 + static_cast<std::size_t>(l)) * 2 + 0) * 12);
      r.contacts[i].cc_transform[l].to12(dst + ((i * 2 + static_cast<std::size_t>(l)) * 2 + 1) * 12);
    }
  return TC_OK;
}
int orc_result_term_ptr(const void* res, int64_t* term_ptr)
{
  const Result& r = *static_cast<const Result*>(res);
  std::copy(r.term_ptr.begin(), r.term_ptr.end(), term_ptr);
  return TC_OK;
}
int orc_result_gradients(const void* res, tc_gradient* rec, double* grad, size_t cap)
{
  const Result& r = *static_cast<const Result*>(res);
  const std::size_t n = static_cast<std::size_t>(r.scene->n_dof);
  for (std::size_t i = 0; i < std::min(cap, r.grads.size()); ++i)
  {
    const GradientResults& g = r.grads[i];
    rec[i].error = g.error;
    rec[i].error_with_buffer = g.error_with_buffer;
    for (std::size_t l = 0; l < 2; ++l)
    {
      rec[i].scale0[l] = g.gradients[l].scale;
      rec[i].scale1[l] = g.cc_gradients[l].scale;
      rec[i].has_gradient[l] = (g.gradients[l].has || g.cc_gradients[l].has) ? 1 : 0;
      rec[i].cc_type[l] = g.gradients[l].has ? g.gradients[l].cc_type : g.cc_gradients[l].cc_type;
      for (std::size_t k = 0; k < 2; ++k)
      {
        const LinkGradient& lg = k ? g.cc_gradients[l] : g.gradients[l];
        double* dst = grad + ((i * 2 + l) * 2 + k) * n;
        for (std::size_t j = 0; j < n; ++j)
          dst[j] = lg.has ? lg.g[j] : 0.0;
      }
    }
  }
  return TC_OK;
}

// rows: see include/trajopt_cuda.h tc_result_rows
int orc_result_rows(const void* res, int64_t* row_ptr, int32_t* col_idx, double* val, double* row_value,
                    double* row_coeff, double* row_margin)
{
  const Result& r = *static_cast<const Result*>(res);
  const Scene& s = *r.scene;
  const int n = s.n_dof;
  const bool single = r.desc.evaluator == TC_EVAL_DISCRETE;
  int64_t nnz = 0;
  row_ptr[0] = 0;
  if (r.grads.size() != r.contacts.size())
  {
    g_err = "rows need an evaluation with TC_WANT_GRADIENTS";
    return TC_ERR_INVALID;
  }
  for (std::size_t i = 0; i < r.contacts.size(); ++i)
  {
    const Contact& c = r.contacts[i];
    const GradientResults& g = r.grads[i];
    const LinkPair& lp = s.pairs[static_cast<std::size_t>(c.c.pair)];
    // which term -> fixed flags
    int ti = 0;
    for (; ti < r.terms_per_traj; ++ti)
      if (r.term_t[static_cast<std::size_t>(ti)] == c.c.term)
        break;
    const bool f0 = r.term_fixed[static_cast<std::size_t>(ti)][0], f1 = r.term_fixed[static_cast<std::size_t>(ti)][1];
    const double* q0 = r.q.data() + (static_cast<std::size_t>(c.c.traj) * static_cast<std::size_t>(r.T) +
                                     static_cast<std::size_t>(c.c.term)) *
                                        static_cast<std::size_t>(n);
    const double* q1 = q0 + n;
    double constant = c.c.distance;
    for (int blk = 0; blk < 2; ++blk)
    {
      if (blk == 0 && f0)
        continue;
      if (blk == 1 && (single || f1))
        continue;
      const double* qb = blk ? q1 : q0;
      for (int j = 0; j < n; ++j)
      {
        double v = 0;
        for (int l = 0; l < 2; ++l)
        {
          const LinkGradient& lg = blk ? g.cc_gradients[l] : g.gradients[l];
          if (lg.has)
            v += lg.scale * lg.g[static_cast<std::size_t>(j)];
        }
        if (r.desc.stack == TC_STACK_LEGACY)
        {
          // exprInc(dist, varDot(scale*grad, vars)); exprInc(dist, scale * -grad.dot(dofvals))
          constant -= v * qb[j];
          val[nnz] = v;
        }
        else
          val[nnz] = -1.0 * v;
        col_idx[nnz] = (c.c.term + blk) * n + j;
        ++nnz;
      }
    }
    row_ptr[i + 1] = nnz;
    row_value[i] = (r.desc.stack == TC_STACK_LEGACY) ? constant : (lp.margin - c.c.distance);
    row_coeff[i] = lp.coeff;
    row_margin[i] = lp.margin;
  }
  return TC_OK;
}

int orc_result_sets(const void* res, int32_t max_num_cnt, int64_t* set_ptr, tc_set* sets, size_t cap_sets,
                    double* values, double* coeffs, double* jac)
{
  ORC_TRY
  const Result& r = *static_cast<const Result*>(res);
  const Scene& s = *r.scene;
  const int n = s.n_dof;
  const long n_terms = static_cast<long>(r.term_ptr.size()) - 1;
  const bool single = r.desc.evaluator == TC_EVAL_DISCRETE;
  std::size_t ns = 0;
  set_ptr[0] = 0;
  for (long k = 0; k < n_terms; ++k)
  {
    std::vector<GradientResultsSet> v = buildSets(r, k, static_cast<std::size_t>(max_num_cnt));
    const std::array<bool, 2> fx = r.term_fixed[static_cast<std::size_t>(k % r.terms_per_traj)];
    for (const GradientResultsSet& g : v)
    {
      if (ns < cap_sets && sets)
      {
        tc_set& o = sets[ns];
        o.traj = static_cast<int>(k / r.terms_per_traj);
        o.term = r.term_t[static_cast<std::size_t>(k % r.terms_per_traj)];
        o.pair = g.pair;
        o.shape_key[0] = static_cast<int>(g.shape_key.first);
        o.shape_key[1] = static_cast<int>(g.shape_key.second);
        o.n_results = static_cast<int>(g.results.size());
        o.coeff = g.coeff;
        for (int i = 0; i < 2; ++i)
          for (int j = 0; j < 2; ++j)
          {
            o.has_error[i][j] = g.max_error[i].has_error[j];
            o.error[i][j] = g.max_error[i].error[j];
            o.error_with_buffer[i][j] = g.max_error[i].error_with_buffer[j];
          }
      }
      ++ns;
    }
    set_ptr[k + 1] = static_cast<int64_t>(ns);
    double* val = values + k * max_num_cnt;
    double* cf = coeffs + k * max_num_cnt;
    double* J = jac + static_cast<std::size_t>(k) * static_cast<std::size_t>(max_num_cnt) * 2 * static_cast<std::size_t>(n);
    std::fill(val, val + max_num_cnt, -s.margin_buffer);
    std::fill(cf, cf + max_num_cnt, 1.0);
    std::fill(J, J + static_cast<std::size_t>(max_num_cnt) * 2 * static_cast<std::size_t>(n), 0.0);
    const std::size_t cnt = std::min<std::size_t>(static_cast<std::size_t>(max_num_cnt), v.size());
    for (std::size_t i = 0; i < cnt; ++i)
    {
      const GradientResultsSet& g = v[i];
      cf[i] = g.coeff;
      double* J0 = J + (i * 2 + 0) * static_cast<std::size_t>(n);
      double* J1 = J + (i * 2 + 1) * static_cast<std::size_t>(n);
      if (single)
      {
        val[i] = g.get(0, false);
        const std::vector<double> w = weightedAvg(g, g.get(0, true), n, false);
        for (int j = 0; j < n; ++j)
          J0[j] = -1.0 * w[static_cast<std::size_t>(j)];
        continue;
      }
      const bool h0 = g.max_error[0].has_error[0] || g.max_error[1].has_error[0];
      const bool h1 = g.max_error[0].has_error[1] || g.max_error[1].has_error[1];
      if (!fx[0] && !fx[1])
        val[i] = g.get(2, false);
      else if (!fx[0])
      {
        if (h0)
          val[i] = g.get(0, false);
      }
      else if (h1)
        val[i] = g.get(1, false);
      if (!fx[0] && h0)
      {
        const double m = !fx[1] ? g.get(2, true) : g.get(0, true);
        const std::vector<double> w = weightedAvg(g, m, n, false);
        for (int j = 0; j < n; ++j)
          J0[j] = -1.0 * w[static_cast<std::size_t>(j)];
      }
      if (!fx[1] && h1)
      {
        const double m = !fx[0] ? g.get(2, true) : g.get(1, true);
        const std::vector<double> w = weightedAvg(g, m, n, true);
        for (int j = 0; j < n; ++j)
          J1[j] = -1.0 * w[static_cast<std::size_t>(j)];
      }
    }
  }
  ORC_CATCH
}

// CollisionCost::value (collision_terms.cpp:1287-1306): margin without the buffer
int orc_result_cost(const void* res, double* cost_per_term, double* cost_per_traj)
{
  const Result& r = *static_cast<const Result*>(res);
  const Scene& s = *r.scene;
  const long n_terms = static_cast<long>(r.term_ptr.size()) - 1;
  if (cost_per_traj)
    std::fill(cost_per_traj, cost_per_traj + r.B, 0.0);
  for (long k = 0; k < n_terms; ++k)
  {
    double out = 0;
    for (int64_t i = r.term_ptr[static_cast<std::size_t>(k)]; i < r.term_ptr[static_cast<std::size_t>(k) + 1]; ++i)
    {
      const Contact& c = r.contacts[static_cast<std::size_t>(i)];
      const LinkPair& lp = s.pairs[static_cast<std::size_t>(c.c.pair)];
      out += std::max(lp.margin - c.c.distance, 0.0) * lp.coeff;
    }
    if (cost_per_term)
      cost_per_term[k] = out;
    if (cost_per_traj)
      cost_per_traj[k / r.terms_per_traj] += out;
  }
  return TC_OK;
}

int orc_num_threads(void)
{
  const unsigned n = std::thread::hardware_concurrency();
  return n ? static_cast<int>(n) : 1;
}
}

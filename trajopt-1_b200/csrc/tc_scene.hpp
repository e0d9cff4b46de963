// tc_scene.hpp — flattened scene tables shared by the host builder and the kernels.
//
// tc_scene_create() stands where the reference's evaluator constructors fetch a contact manager,
// set the active links, snapshot the margin data and inflate the manager's margin by the buffer
// (trajopt/src/collision_terms.cpp:600-644; trajopt_ifopt/.../discrete_collision_evaluators.cpp:38-87;
// continuous_collision_evaluators.cpp:41-95,266-315).  The builder is plain host C++ so that the
// unit-test harness can reuse it without CUDA.
#pragma once
#include <cmath>
#include <cstring>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/trajopt_cuda.h"
#include "tc_narrow.cuh"

namespace tc
{
// one candidate link pair (>= 1 active link, not ACM-disabled); link a is always active, a < b
struct PairD
{
  int32_t a, b;
  int32_t sp_begin, sp_count;  // its shape pairs
  int32_t zero_coeff, pad;
  double margin, coeff, thr;   // thr = margin + buffer (manager margin after incrementCollisionMargin)
};

struct ShapePairD
{
  int32_t pair, sa, sb, k;  // k = position inside the link pair (contact order within a link pair)
};

struct SceneError : std::runtime_error
{
  int code;
  SceneError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

struct HostScene
{
  int n_dof{ 0 }, n_frames{ 0 }, n_links{ 0 }, n_active{ 0 }, n_shapes{ 0 }, n_active_shapes{ 0 };
  int n_lead{ 0 };  // links that may lead a candidate pair: the group's links + the environment's other moving links
  std::vector<int32_t> frame_parent, frame_type, frame_dof;
  std::vector<double> frame_axis, frame_origin;
  double base_pose[12];
  std::vector<int32_t> path_ptr, path_frames;    // ancestor frames (root first, incl. self) of every frame
  std::vector<int32_t> link_frame;
  std::vector<int32_t> frame_link;               // first active link riding on a frame, -1 if none
  std::vector<unsigned long long> link_dof_mask; // per link: bit j set if joint j moves the link (0 for static links)
  std::vector<double> link_pose;                 // [n_links][12] (static links)
  std::vector<ShapeD> shapes;
  std::vector<Vtx<double>> verts;                // link-frame hull vertices, 32 B each
  std::vector<int32_t> active_shapes;            // shape index of every active shape
  std::vector<int32_t> active_shape_frame;       // frame its link rides on
  std::vector<double> static_aabb;               // [n_shapes][6] world lo/hi (static shapes only)
  std::vector<PairD> pairs;
  std::vector<ShapePairD> spairs;
  std::vector<int32_t> partner_ptr, partner_sp;  // per active shape: the shape pairs it leads
  std::vector<int32_t> partner_word_ptr;         // per active shape: first 32-bit word of its broadphase mask
  std::vector<int32_t> curved_sp;                // shape pairs with a cylinder or cone (served by tc_strict.cu)
  double margin_buffer{ 0 };
  int max_hull_verts{ 0 };
  std::vector<uint8_t> lut;  // support look-up tables of the hulls (ShapeD::lut_off / lut_bytes)
  int max_lut_bytes{ 0 };
};

// raw-pointer view of the tables (host vectors in the test harness, device arrays in the kernels)
struct SceneView
{
  int n_dof, n_frames, n_links, n_active, n_shapes, n_active_shapes, n_pairs, n_spairs, max_hull_verts;
  int max_lut_bytes;
  const uint8_t* lut;
  const int32_t *frame_parent, *frame_type, *frame_dof;
  const double *frame_axis, *frame_origin;
  double base_pose[12];
  const int32_t *path_ptr, *path_frames;
  const int32_t* link_frame;
  const int32_t* frame_link;
  const unsigned long long* link_dof_mask;
  const double* link_pose;
  const ShapeD* shapes;
  const Vtx<double>* verts;
  const Vtx<float>* verts_f32;
  const int32_t* active_shapes;
  const int32_t* active_shape_frame;
  const double* static_aabb;
  const PairD* pairs;
  const ShapePairD* spairs;
  const int32_t *partner_ptr, *partner_sp, *partner_word_ptr;
  const int32_t* curved_sp;
  int n_curved_sp;
  double margin_buffer;
};

inline SceneView hostView(const HostScene& s)
{
  SceneView v{};
  v.n_dof = s.n_dof, v.n_frames = s.n_frames, v.n_links = s.n_links, v.n_active = s.n_active;
  v.n_shapes = s.n_shapes, v.n_active_shapes = s.n_active_shapes;
  v.n_pairs = static_cast<int>(s.pairs.size()), v.n_spairs = static_cast<int>(s.spairs.size());
  v.max_hull_verts = s.max_hull_verts;
  v.max_lut_bytes = s.max_lut_bytes;
  v.lut = s.lut.data();
  v.frame_parent = s.frame_parent.data(), v.frame_type = s.frame_type.data(), v.frame_dof = s.frame_dof.data();
  v.frame_axis = s.frame_axis.data(), v.frame_origin = s.frame_origin.data();
  for (int i = 0; i < 12; ++i)
    v.base_pose[i] = s.base_pose[i];
  v.path_ptr = s.path_ptr.data(), v.path_frames = s.path_frames.data();
  v.link_frame = s.link_frame.data(), v.link_pose = s.link_pose.data();
  v.frame_link = s.frame_link.data();
  v.link_dof_mask = s.link_dof_mask.data();
  v.shapes = s.shapes.data(), v.verts = s.verts.data(), v.verts_f32 = nullptr;
  v.active_shapes = s.active_shapes.data(), v.static_aabb = s.static_aabb.data();
  v.active_shape_frame = s.active_shape_frame.data();
  v.pairs = s.pairs.data(), v.spairs = s.spairs.data();
  v.partner_ptr = s.partner_ptr.data(), v.partner_sp = s.partner_sp.data();
  v.partner_word_ptr = s.partner_word_ptr.data();
  v.curved_sp = s.curved_sp.data();
  v.n_curved_sp = static_cast<int>(s.curved_sp.size());
  v.margin_buffer = s.margin_buffer;
  return v;
}

// Support look-up table of one hull (layout and use: tc_narrow.cuh).  For every cube-map cell take the cone around
// the cell's centre direction c that contains the cell (half-angle theta = largest angle to a corner, plus slack for
// the float binning on the device).  Vertex i cannot be extreme for any direction of the cone when some vertex j has
// d.(v_j - v_i) > 0 for every d of the cone, i.e. c.e > |e| sin(theta) with e = v_j - v_i; a margin keeps the
// decision clear of rounding on either side.  Everything else is listed.  Appends to `blob`, 16-byte aligned.
inline void buildSupportLut(const Vtx<double>* verts, int n, std::vector<uint8_t>& blob)
{
  std::vector<uint16_t> off(LUT_CELLS + 1, 0);
  std::vector<uint8_t> cand;
  std::vector<double> proj(static_cast<size_t>(n));
  for (int cell = 0; cell < LUT_CELLS; ++cell)
  {
    const int iu = cell % LUT_G, iv = (cell / LUT_G) % LUT_G, face = cell / (LUT_G * LUT_G);
    const int m = face / 2;
    const double sg = (face & 1) ? -1.0 : 1.0;
    auto dir = [&](double u, double v) {
      double d[3];
      d[m] = sg;
      d[(m + 1) % 3] = u;
      d[(m + 2) % 3] = v;
      const double nn = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
      return Vec3<double>(d[0] / nn, d[1] / nn, d[2] / nn);
    };
    const double u0 = -1.0 + 2.0 * iu / LUT_G, u1 = -1.0 + 2.0 * (iu + 1) / LUT_G;
    const double v0 = -1.0 + 2.0 * iv / LUT_G, v1 = -1.0 + 2.0 * (iv + 1) / LUT_G;
    const Vec3<double> c = dir(0.5 * (u0 + u1), 0.5 * (v0 + v1));
    double cmin = 1.0;
    for (int k = 0; k < 4; ++k)
      cmin = std::fmin(cmin, dot(c, dir((k & 1) ? u1 : u0, (k & 2) ? v1 : v0)));
    const double theta = std::acos(std::fmax(-1.0, std::fmin(1.0, cmin))) + 1e-4;
    const double sin_t = std::sin(theta);
    int jbest = 0;
    for (int i = 0; i < n; ++i)
    {
      proj[static_cast<size_t>(i)] = c.x * verts[i].x + c.y * verts[i].y + c.z * verts[i].z;
      if (proj[static_cast<size_t>(i)] > proj[static_cast<size_t>(jbest)])
        jbest = i;
    }
    auto beats = [&](int j, int i) {
      const double ex = verts[j].x - verts[i].x, ey = verts[j].y - verts[i].y, ez = verts[j].z - verts[i].z;
      const double len = std::sqrt(ex * ex + ey * ey + ez * ez);
      return proj[static_cast<size_t>(j)] - proj[static_cast<size_t>(i)] > len * (sin_t + 1e-9) + 1e-12;
    };
    for (int i = 0; i < n; ++i)
    {
      bool out = beats(jbest, i);
      for (int j = 0; j < n && !out; ++j)
        out = beats(j, i);
      if (!out)
        cand.push_back(static_cast<uint8_t>(i));
    }
    if (cand.size() > 65535)
      throw SceneError(TC_ERR_UNSUPPORTED, "support look-up table too large");
    off[static_cast<size_t>(cell) + 1] = static_cast<uint16_t>(cand.size());
  }
  const size_t base = blob.size();
  const size_t bytes = (off.size() * sizeof(uint16_t) + cand.size() + 15) / 16 * 16;
  blob.resize(base + bytes, 0);
  std::memcpy(blob.data() + base, off.data(), off.size() * sizeof(uint16_t));
  std::memcpy(blob.data() + base + off.size() * sizeof(uint16_t), cand.data(), cand.size());
}

// world AABB of static shape i at its link's current pose
inline void staticShapeAabb(HostScene& s, int i)
{
  const ShapeD& sh = s.shapes[static_cast<size_t>(i)];
  const Pose<double> T = Pose<double>::from12(s.link_pose.data() + 12 * sh.link);
  const Vec3<double> wc = T.apply(Vec3<double>(sh.aabb_c[0], sh.aabb_c[1], sh.aabb_c[2]));
  for (int k = 0; k < 3; ++k)
  {
    const double e = std::fabs(T.r[3 * k + 0]) * sh.aabb_h[0] + std::fabs(T.r[3 * k + 1]) * sh.aabb_h[1] +
                     std::fabs(T.r[3 * k + 2]) * sh.aabb_h[2];
    s.static_aabb[static_cast<size_t>(6 * i + k)] = wc.get(k) - e;
    s.static_aabb[static_cast<size_t>(6 * i + 3 + k)] = wc.get(k) + e;
  }
}

// Dynamic environment (trajopt/src/collision_terms.cpp:399-411,659-660,828-836): links that are not moved by the
// planning group may still move between evaluations (another robot, a positioner, an attached object).  The reference
// reads their poses from the environment state on every call; here the caller pushes the new world poses of the
// non-group links and the static boxes are refreshed.  Poses of group links in `link_pose` are ignored.
inline void updateStaticPoses(HostScene& s, const double* link_pose)
{
  for (int l = s.n_active; l < s.n_links; ++l)
    for (int k = 0; k < 12; ++k)
      s.link_pose[static_cast<size_t>(12 * l + k)] = link_pose[12 * l + k];
  for (int i = 0; i < s.n_shapes; ++i)
    if (s.shapes[static_cast<size_t>(i)].link >= s.n_active)
      staticShapeAabb(s, i);
}

inline void buildHostScene(const tc_scene_desc& d, HostScene& s)
{
  auto fail = [](const std::string& m) { throw SceneError(TC_ERR_INVALID, m); };
  if (d.n_dof <= 0 || d.n_frames <= 0 || d.n_links <= 0 || d.n_active_links <= 0 || d.n_active_links > d.n_links)
    fail("scene description has empty kinematics or links");
  if (d.n_dof > 64)
    throw SceneError(TC_ERR_UNSUPPORTED, "more than 64 degrees of freedom");
  s.n_dof = d.n_dof;
  s.n_frames = d.n_frames;
  s.frame_parent.assign(d.frame_parent, d.frame_parent + d.n_frames);
  s.frame_type.assign(d.frame_joint_type, d.frame_joint_type + d.n_frames);
  s.frame_dof.assign(d.frame_dof, d.frame_dof + d.n_frames);
  s.frame_axis.assign(d.frame_axis, d.frame_axis + 3 * d.n_frames);
  s.frame_origin.assign(d.frame_origin, d.frame_origin + 12 * d.n_frames);
  for (int i = 0; i < 12; ++i)
    s.base_pose[i] = d.base_pose[i];
  std::vector<std::vector<int32_t>> paths(static_cast<size_t>(d.n_frames));
  for (int f = 0; f < d.n_frames; ++f)
  {
    if (s.frame_parent[f] >= f)
      fail("frames are not topologically ordered");
    if (s.frame_parent[f] >= 0)
      paths[f] = paths[s.frame_parent[f]];
    paths[f].push_back(f);
    if (s.frame_type[f] != TC_JOINT_FIXED)
    {
      if (s.frame_type[f] != TC_JOINT_REVOLUTE && s.frame_type[f] != TC_JOINT_PRISMATIC)
        fail("unknown joint type");
      if (s.frame_dof[f] < 0 || s.frame_dof[f] >= d.n_dof)
        fail("movable frame without a dof index");
    }
  }
  s.path_ptr.assign(1, 0);
  for (int f = 0; f < d.n_frames; ++f)
  {
    s.path_frames.insert(s.path_frames.end(), paths[f].begin(), paths[f].end());
    s.path_ptr.push_back(static_cast<int32_t>(s.path_frames.size()));
  }
  s.n_links = d.n_links;
  s.n_active = d.n_active_links;
  if (d.n_env_active_links < 0 || d.n_active_links + d.n_env_active_links > d.n_links)
    fail("n_env_active_links out of range");
  s.n_lead = d.n_active_links + d.n_env_active_links;
  s.link_frame.assign(d.link_frame, d.link_frame + d.n_links);
  s.link_pose.assign(d.link_pose, d.link_pose + 12 * d.n_links);
  for (int l = 0; l < d.n_links; ++l)
  {
    if ((l < d.n_active_links) != (d.link_frame[l] >= 0))
      fail("active links must come first and carry a frame");
    if (d.link_frame[l] >= d.n_frames)
      fail("link frame out of range");
  }
  // joints that move a link = the movable frames on its chain path (the structural non-zeros of its Jacobian rows)
  s.link_dof_mask.assign(static_cast<size_t>(d.n_links), 0ull);
  for (int l = 0; l < d.n_links; ++l)
  {
    const int lf = d.link_frame[l];
    if (lf < 0)
      continue;
    for (int k = s.path_ptr[lf]; k < s.path_ptr[lf + 1]; ++k)
    {
      const int f = s.path_frames[k];
      if (s.frame_type[f] != TC_JOINT_FIXED && s.frame_dof[f] < 64)
        s.link_dof_mask[static_cast<size_t>(l)] |= 1ull << s.frame_dof[f];
    }
  }
  s.frame_link.assign(static_cast<size_t>(d.n_frames), -1);
  for (int l = d.n_active_links - 1; l >= 0; --l)
    s.frame_link[static_cast<size_t>(d.link_frame[l])] = l;
  s.n_shapes = d.n_shapes;
  s.margin_buffer = d.margin_buffer;
  std::vector<std::vector<int>> link_shapes(static_cast<size_t>(d.n_links));
  s.static_aabb.assign(static_cast<size_t>(6 * d.n_shapes), 0.0);
  for (int i = 0; i < d.n_shapes; ++i)
  {
    ShapeD sh{};
    sh.type = d.shape_type[i];
    sh.link = d.shape_link[i];
    sh.shape_id = d.shape_id[i];
    if (sh.link < 0 || sh.link >= d.n_links)
      fail("shape link out of range");
    const Pose<double> sp = Pose<double>::from12(d.shape_pose + 12 * i);
    const double* prm = d.shape_params + 4 * i;
    Vec3<double> c = sp.t;
    const Vec3<double> ax = sp.rot(Vec3<double>(0, 0, 1));
    double lo[3] = { 1e300, 1e300, 1e300 }, hi[3] = { -1e300, -1e300, -1e300 };
    sh.vert_off = static_cast<int32_t>(s.verts.size());
    sh.vert_cnt = 0;
    sh.lut_off = -1;
    sh.lut_bytes = 0;
    switch (sh.type)
    {
      case TC_SHAPE_SPHERE:
        sh.radius = prm[0];
        for (int k = 0; k < 3; ++k)
          lo[k] = c.get(k) - prm[0], hi[k] = c.get(k) + prm[0];
        sh.brad = 0;
        break;
      case TC_SHAPE_CAPSULE:
      case TC_SHAPE_CYLINDER:
      case TC_SHAPE_CONE:
        sh.radius = prm[0];
        sh.half_len = prm[1];
        for (int k = 0; k < 3; ++k)
        {
          const double e = std::fabs(ax.get(k)) * sh.half_len + sh.radius;
          lo[k] = c.get(k) - e, hi[k] = c.get(k) + e;
        }
        sh.brad = (sh.type == TC_SHAPE_CAPSULE) ? sh.half_len : std::sqrt(sh.half_len * sh.half_len + sh.radius * sh.radius);
        break;
      case TC_SHAPE_HULL:
      {
        const int n = d.shape_vert_count[i];
        if (n <= 0)
          fail("hull without vertices");
        if (n > 1023)
          throw SceneError(TC_ERR_UNSUPPORTED, "convex hull with more than 1023 vertices");
        Vec3<double> sum;
        for (int k = 0; k < n; ++k)
        {
          const double* p = d.verts + 3 * (d.shape_vert_offset[i] + k);
          const Vec3<double> v = sp.apply(Vec3<double>(p[0], p[1], p[2]));
          s.verts.push_back(Vtx<double>{ v.x, v.y, v.z, 0.0 });
          sum = sum + v;
          for (int a = 0; a < 3; ++a)
          {
            lo[a] = std::fmin(lo[a], v.get(a));
            hi[a] = std::fmax(hi[a], v.get(a));
          }
        }
        sh.vert_cnt = n;
        c = (1.0 / n) * sum;
        for (int k = 0; k < n; ++k)
        {
          const Vtx<double>& q = s.verts[static_cast<size_t>(sh.vert_off + k)];
          sh.brad = std::fmax(sh.brad, norm(Vec3<double>(q.x, q.y, q.z) - c));
        }
        if (n > s.max_hull_verts)
          s.max_hull_verts = n;
        if (n >= LUT_MIN_VERTS && n <= LUT_MAX_VERTS)
        {
          sh.lut_off = static_cast<int32_t>(s.lut.size());
          buildSupportLut(s.verts.data() + sh.vert_off, n, s.lut);
          sh.lut_bytes = static_cast<int32_t>(s.lut.size()) - sh.lut_off;
          if (sh.lut_bytes > s.max_lut_bytes)
            s.max_lut_bytes = sh.lut_bytes;
        }
        break;
      }
      default:
        throw SceneError(TC_ERR_UNSUPPORTED, "shape type outside the scoped geometry (convex primitives and hulls)");
    }
    for (int k = 0; k < 3; ++k)
    {
      sh.center[k] = c.get(k);
      sh.axis[k] = ax.get(k);
      sh.aabb_c[k] = 0.5 * (lo[k] + hi[k]);
      sh.aabb_h[k] = 0.5 * (hi[k] - lo[k]);
    }
    sh.active_index = -1;
    if (sh.link < s.n_lead)
    {
      // a shape that leads pairs in the broadphase owns a slot of the box arena; a moving link outside the group has no
      // frame (-1): its box is the same in every state and is written from its current world pose
      sh.active_index = static_cast<int32_t>(s.active_shapes.size());
      s.active_shapes.push_back(i);
      s.active_shape_frame.push_back(sh.link < d.n_active_links ? d.link_frame[sh.link] : -1);
    }
    if (sh.link >= d.n_active_links)
    {
      const Pose<double> T = Pose<double>::from12(d.link_pose + 12 * sh.link);
      const Vec3<double> wc = T.apply(Vec3<double>(sh.aabb_c[0], sh.aabb_c[1], sh.aabb_c[2]));
      for (int k = 0; k < 3; ++k)
      {
        const double e = std::fabs(T.r[3 * k + 0]) * sh.aabb_h[0] + std::fabs(T.r[3 * k + 1]) * sh.aabb_h[1] +
                         std::fabs(T.r[3 * k + 2]) * sh.aabb_h[2];
        s.static_aabb[static_cast<size_t>(6 * i + k)] = wc.get(k) - e;
        s.static_aabb[static_cast<size_t>(6 * i + 3 + k)] = wc.get(k) + e;
      }
    }
    link_shapes[static_cast<size_t>(sh.link)].push_back(i);
    s.shapes.push_back(sh);
  }
  s.n_active_shapes = static_cast<int>(s.active_shapes.size());
  std::vector<std::vector<int32_t>> partners(static_cast<size_t>(s.n_active_shapes));
  for (int a = 0; a < d.n_links; ++a)
    for (int b = a + 1; b < d.n_links; ++b)
    {
      if (a >= s.n_lead)
        continue;  // leading links come first, so link a of a candidate pair always is one
      if (d.acm && (d.acm[a * d.n_links + b] || d.acm[b * d.n_links + a]))
        continue;
      if (link_shapes[static_cast<size_t>(a)].empty() || link_shapes[static_cast<size_t>(b)].empty())
        continue;
      PairD p{};
      p.a = a;
      p.b = b;
      p.margin = d.pair_margin ? d.pair_margin[a * d.n_links + b] : d.default_margin;
      p.coeff = d.pair_coeff ? d.pair_coeff[a * d.n_links + b] : d.default_coeff;
      // CollisionCoeffData::setCollisionCoeff treats |coeff| ~ 0 as "pair disabled"
      // (trajopt_common/src/collision_types.cpp:56,69-72)
      p.zero_coeff = std::fabs(p.coeff) <= 1e-6 ? 1 : 0;
      p.thr = p.margin + d.margin_buffer;
      p.sp_begin = static_cast<int32_t>(s.spairs.size());
      int k = 0;
      for (int sa : link_shapes[static_cast<size_t>(a)])
        for (int sb : link_shapes[static_cast<size_t>(b)])
        {
          partners[static_cast<size_t>(s.shapes[static_cast<size_t>(sa)].active_index)].push_back(
              static_cast<int32_t>(s.spairs.size()));
          if (s.shapes[static_cast<size_t>(sa)].type >= TC_SHAPE_CYLINDER || s.shapes[static_cast<size_t>(sb)].type >= TC_SHAPE_CYLINDER)
            s.curved_sp.push_back(static_cast<int32_t>(s.spairs.size()));
          s.spairs.push_back(ShapePairD{ static_cast<int32_t>(s.pairs.size()), sa, sb, k++ });
        }
      p.sp_count = k;
      if (k > 256)
        throw SceneError(TC_ERR_UNSUPPORTED, "more than 256 shape pairs in one link pair");
      s.pairs.push_back(p);
    }
  if (s.pairs.size() > 16384)
    throw SceneError(TC_ERR_UNSUPPORTED, "more than 16384 candidate link pairs");
  s.partner_ptr.assign(1, 0);
  s.partner_word_ptr.assign(1, 0);
  for (const auto& v : partners)
  {
    s.partner_sp.insert(s.partner_sp.end(), v.begin(), v.end());
    s.partner_ptr.push_back(static_cast<int32_t>(s.partner_sp.size()));
    s.partner_word_ptr.push_back(s.partner_word_ptr.back() + static_cast<int32_t>((v.size() + 31) / 32));
  }
}
}  // namespace tc

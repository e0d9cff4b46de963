// tc_evaluators.hpp — host-side mirror of the reference's collision evaluators over the C ABI.
//
// Seam B1 of SURVEY.md 8(b): trajopt_ifopt's constraints hold a std::shared_ptr to
//   DiscreteCollisionEvaluator   ::calcCollisions(CollisionCacheData&, dof_vals, max_allowed)
//       trajopt_ifopt/include/trajopt_ifopt/constraints/collision/discrete_collision_evaluators.h:50-98
//   ContinuousCollisionEvaluator ::calcCollisionData(CollisionCacheData&, dof_vals0, dof_vals1,
//                                                    vars0_fixed, vars1_fixed, max_allowed)
//       trajopt_ifopt/include/trajopt_ifopt/constraints/collision/continuous_collision_evaluators.h:51-108
// The classes below have the same names, arguments and error behaviour (std::runtime_error for a
// mis-typed config, callee clears and fully overwrites the caller-owned CollisionCacheData) and do their
// work through libtrajopt_cuda.so.  There is no CPU fallback: without a B200 tc_init fails and the
// constructors throw.
//
// Two ways to run:
//   * standalone  — each calcCollisions / calcCollisionData call is one tc_eval (B = 1, T = 1 or 2).
//                   Zero changes to the calling constraint; latency-bound (one small pass per call).
//   * batched     — evaluators share a CollisionBatch.  The SQP driver calls batch->update(x) once per
//                   iterate with the whole T x n_dof variable vector; one tc_eval then covers every term
//                   and each evaluator only slices the shared result (the reference's per-x memoisation,
//                   trajopt_common/include/trajopt_common/cache.h:32-329, lifted to the whole iterate).
//                   A call whose joint values are not the batch's falls back to a standalone GPU pass.
#pragma once
#include <algorithm>
#include <cstring>
#include <limits>
#include <mutex>
#include <string>

#include "tc_types.hpp"

namespace trajopt_cuda
{
inline void check(int status, const char* what)
{
  if (status != TC_OK)
    throw std::runtime_error(std::string(what) + ": " + tc_last_error());
}

// One (Environment, JointGroup, margin/coeff data) triple uploaded to the GPU: evaluator construction,
// trajopt_ifopt/src/constraints/collision/discrete_collision_evaluators.cpp:38-87.
class Scene
{
public:
  using Ptr = std::shared_ptr<Scene>;
  struct Pair
  {
    int32_t link0, link1;
    double margin, coeff;
  };
  Scene(const tc_scene_desc& desc, int device = 0) : n_dof_(desc.n_dof), margin_buffer_(desc.margin_buffer)
  {
    check(tc_init(device), "tc_init");
    check(tc_scene_create(&desc, &scene_), "tc_scene_create");
    const int n = tc_scene_num_pairs(scene_);
    std::vector<int32_t> l0(static_cast<std::size_t>(n)), l1(static_cast<std::size_t>(n));
    std::vector<double> m(static_cast<std::size_t>(n)), c(static_cast<std::size_t>(n));
    check(tc_scene_pairs(scene_, l0.data(), l1.data(), m.data(), c.data()), "tc_scene_pairs");
    pairs_.resize(static_cast<std::size_t>(n));
    for (std::size_t i = 0; i < pairs_.size(); ++i)
      pairs_[i] = Pair{ l0[i], l1[i], m[i], c[i] };
  }
  ~Scene() { tc_scene_destroy(scene_); }
  Scene(const Scene&) = delete;
  Scene& operator=(const Scene&) = delete;
  tc_scene* handle() const { return scene_; }
  // dynamic environment (collision_terms.cpp:399-411): new world poses of the links outside the planning group,
  // link_pose[n_links][12]; must not overlap an evaluation on this scene
  void setLinkPoses(const double* link_pose) { check(tc_scene_set_link_poses(scene_, link_pose), "tc_scene_set_link_poses"); }
  int nDof() const { return n_dof_; }
  double marginBuffer() const { return margin_buffer_; }
  const std::vector<Pair>& pairs() const { return pairs_; }  // CollisionMarginData + CollisionCoeffData per pair

private:
  tc_scene* scene_{ nullptr };
  int n_dof_;
  double margin_buffer_;
  std::vector<Pair> pairs_;
};

// Host copy of one tc_eval's contacts and gradients.
struct HostResult
{
  std::vector<tc_contact> contacts;
  std::vector<double> transforms;  // [n][2 links][T0, T1][12]
  std::vector<tc_gradient> grec;
  std::vector<double> grad;        // [n][2 links][T0 block, T1 block][n_dof]
  std::vector<int64_t> term_ptr;
};

class Evaluation
{
public:
  explicit Evaluation(Scene::Ptr scene) : scene_(std::move(scene))
  {
    check(tc_result_create(scene_->handle(), &res_), "tc_result_create");
  }
  ~Evaluation() { tc_result_destroy(res_); }
  Evaluation(const Evaluation&) = delete;
  Evaluation& operator=(const Evaluation&) = delete;

  // one pass over q[B][T][n_dof]; contacts and gradients are copied to the host
  void run(const TrajOptCollisionConfig& cfg, const double* q, int B, int T, const uint8_t* fixed, HostResult& out,
           int stack = TC_STACK_IFOPT)
  {
    tc_eval_desc d{};
    d.evaluator = static_cast<int32_t>(cfg.type);
    d.stack = stack;
    d.precision = cfg.precision;
    d.want = TC_WANT_CONTACTS | TC_WANT_GRADIENTS;
    d.lvs_length = cfg.longest_valid_segment_length;
    d.fixed_steps = fixed;
    check(tc_eval(scene_->handle(), &d, q, B, T, res_, nullptr), "tc_eval");
    tc_counts c{};
    check(tc_result_counts(res_, &c), "tc_result_counts");
    if (c.overflow)
      throw std::runtime_error("tc_eval: device arena overflow");
    const std::size_t n = static_cast<std::size_t>(c.n_contacts), nd = static_cast<std::size_t>(scene_->nDof());
    out.contacts.resize(n);
    out.transforms.resize(n * 48);
    out.grec.resize(n);
    out.grad.resize(n * 4 * nd);
    out.term_ptr.resize(static_cast<std::size_t>(c.n_terms) + 1);
    check(tc_result_term_ptr(res_, out.term_ptr.data()), "tc_result_term_ptr");
    if (n == 0)
      return;
    check(tc_result_contacts(res_, out.contacts.data(), n), "tc_result_contacts");
    check(tc_result_contact_transforms(res_, out.transforms.data(), n), "tc_result_contact_transforms");
    check(tc_result_gradients(res_, out.grec.data(), out.grad.data(), n), "tc_result_gradients");
  }
  const Scene::Ptr& scene() const { return scene_; }

private:
  Scene::Ptr scene_;
  tc_result* res_{ nullptr };
};

namespace detail
{
inline ContactResult toContact(const tc_contact& c, const double* xf)
{
  ContactResult r;
  r.distance = c.distance;
  for (std::size_t i = 0; i < 2; ++i)
  {
    r.link_ids[i] = c.link_ids[i];
    r.shape_id[i] = c.shape_id[i];
    r.subshape_id[i] = c.subshape_id[i];
    r.cc_time[i] = c.cc_time[i];
    r.cc_type[i] = static_cast<ContinuousCollisionType>(c.cc_type[i]);
    for (std::size_t k = 0; k < 3; ++k)
    {
      r.nearest_points[i][k] = c.nearest_points[i][k];
      r.nearest_points_local[i][k] = c.nearest_points_local[i][k];
    }
    std::memcpy(r.transform[i].data(), xf + (2 * i) * 12, 12 * sizeof(double));
    std::memcpy(r.cc_transform[i].data(), xf + (2 * i + 1) * 12, 12 * sizeof(double));
  }
  for (std::size_t k = 0; k < 3; ++k)
    r.normal[k] = c.normal[k];
  return r;
}

inline GradientResults toGradient(const tc_contact& c, const tc_gradient& g, const double* grad, std::size_t n_dof,
                                  bool continuous)
{
  GradientResults r;
  r.error = g.error;
  r.error_with_buffer = g.error_with_buffer;
  for (std::size_t i = 0; i < 2; ++i)
  {
    const double sign = (i == 0) ? -1.0 : 1.0;  // link A moves against the normal (collision_utils.cpp:153-159)
    for (int blk = 0; blk < (continuous ? 2 : 1); ++blk)
    {
      LinkGradientResults& lg = blk ? r.cc_gradients[i] : r.gradients[i];
      lg.has_gradient = g.has_gradient[i] != 0;
      lg.cc_type = static_cast<ContinuousCollisionType>(g.cc_type[i]);
      lg.scale = blk ? g.scale1[i] : g.scale0[i];
      if (!lg.has_gradient)
        continue;
      const double* src = grad + (2 * i + static_cast<std::size_t>(blk)) * n_dof;
      lg.gradient.assign(src, src + n_dof);
      for (std::size_t k = 0; k < 3; ++k)
        lg.translation_vector[k] = sign * c.normal[k];
    }
  }
  return r;
}

// contacts [lo, hi) of one term -> CollisionCacheData, following
// SingleTimestepCollisionEvaluator::calcCollisions (discrete_collision_evaluators.cpp:89-149) and
// LVS{Continuous,Discrete}CollisionEvaluator::calcCollisionData (continuous_collision_evaluators.cpp:97-164,317-384)
inline void fillCacheData(CollisionCacheData& out, const Scene& scene, const HostResult& h, int64_t lo, int64_t hi,
                          bool continuous, bool vars0_fixed, bool vars1_fixed, std::size_t max_allowed)
{
  out.contact_results_map.clear();
  out.gradient_results_sets.clear();
  const std::size_t nd = static_cast<std::size_t>(scene.nDof());
  for (int64_t i = lo; i < hi; ++i)
  {
    const tc_contact& c = h.contacts[static_cast<std::size_t>(i)];
    out.contact_results_map[LinkIdPair(c.link_ids[0], c.link_ids[1])].push_back(
        toContact(c, h.transforms.data() + static_cast<std::size_t>(i) * 48));
  }
  if (max_allowed == 0 || lo == hi)  // contacts only (discrete_collision_evaluators.cpp:97-99)
    return;
  out.gradient_results_sets.reserve(static_cast<std::size_t>(hi - lo));
  int64_t i = lo;
  while (i < hi)
  {
    // contacts arrive sorted by link pair; inside a pair they are grouped by the shape-pair key
    const int32_t pair = h.contacts[static_cast<std::size_t>(i)].pair;
    std::map<std::pair<std::size_t, std::size_t>, GradientResultsSet> by_shape;
    for (; i < hi && h.contacts[static_cast<std::size_t>(i)].pair == pair; ++i)
    {
      const std::size_t k = static_cast<std::size_t>(i);
      const tc_contact& c = h.contacts[k];
      const auto key = std::make_pair(cantorHash(c.shape_id[0], c.subshape_id[0]),
                                      cantorHash(c.shape_id[1], c.subshape_id[1]));
      auto ins = by_shape.try_emplace(key);
      GradientResultsSet& grs = ins.first->second;
      if (ins.second)
      {
        grs.key = LinkIdPair(c.link_ids[0], c.link_ids[1]);
        grs.shape_key = key;
        grs.coeff = scene.pairs()[static_cast<std::size_t>(pair)].coeff;
        grs.is_continuous = continuous;
      }
      grs.add(toGradient(c, h.grec[k], h.grad.data() + k * 4 * nd, nd, continuous));
    }
    for (auto& kv : by_shape)
      out.gradient_results_sets.push_back(std::move(kv.second));
  }
  if (out.gradient_results_sets.size() <= max_allowed)
    return;
  auto& sets = out.gradient_results_sets;
  if (!continuous)
  {
    std::stable_sort(sets.begin(), sets.end(), [](const GradientResultsSet& a, const GradientResultsSet& b) {
      return a.max_error[0].error > b.max_error[0].error;  // std::array comparison, as the reference's
    });
  }
  else if (!vars0_fixed && !vars1_fixed)
    std::stable_sort(sets.begin(), sets.end(), [](const GradientResultsSet& a, const GradientResultsSet& b) {
      return a.getMaxErrorWithBuffer() > b.getMaxErrorWithBuffer();
    });
  else if (!vars0_fixed)
    std::stable_sort(sets.begin(), sets.end(), [](const GradientResultsSet& a, const GradientResultsSet& b) {
      return a.getMaxErrorWithBufferT0() > b.getMaxErrorWithBufferT0();
    });
  else
    std::stable_sort(sets.begin(), sets.end(), [](const GradientResultsSet& a, const GradientResultsSet& b) {
      return a.getMaxErrorWithBufferT1() > b.getMaxErrorWithBufferT1();
    });
}
}  // namespace detail

// Whole-iterate memo shared by the evaluators of one problem (see the header comment).
class CollisionBatch
{
public:
  using Ptr = std::shared_ptr<CollisionBatch>;
  CollisionBatch(Scene::Ptr scene, const TrajOptCollisionConfig& cfg, int n_steps, std::vector<uint8_t> fixed_steps)
    : eval_(std::move(scene)), cfg_(cfg), T_(n_steps), fixed_(std::move(fixed_steps))
  {
    if (static_cast<int>(fixed_.size()) != T_)
      throw std::runtime_error("CollisionBatch: fixed_steps must have one entry per timestep");
    // tc_eval skips fixed timesteps (DISCRETE) and segments with both ends fixed (two-state evaluators) when it builds
    // its term list (CollisionTermInfo::hatch, trajopt/src/problem_description.cpp:1735-1858): term k of the result is
    // not timestep / segment k.  term_of_[t] = index into term_ptr, -1 = skipped.
    const bool discrete = cfg_.type == CollisionEvaluatorType::DISCRETE;
    term_of_.assign(static_cast<std::size_t>(T_), -1);
    int k = 0;
    for (int t = 0; t < T_; ++t)
    {
      const bool cur = fixed_[static_cast<std::size_t>(t)] != 0;
      if (discrete)
      {
        if (!cur)
          term_of_[static_cast<std::size_t>(t)] = k++;
      }
      else if (t + 1 < T_ && !(cur && fixed_[static_cast<std::size_t>(t) + 1] != 0))
        term_of_[static_cast<std::size_t>(t)] = k++;
    }
  }
  // x = the iterate's joint values, row-major [n_steps][n_dof].  Re-evaluates only when x changed.
  void update(const double* x)
  {
    std::lock_guard<std::mutex> lock(mutex_);
    const std::size_t n = static_cast<std::size_t>(T_) * static_cast<std::size_t>(eval_.scene()->nDof());
    if (valid_ && std::memcmp(x_.data(), x, n * sizeof(double)) == 0)
      return;
    x_.assign(x, x + n);
    eval_.run(cfg_, x_.data(), 1, T_, fixed_.data(), host_);
    valid_ = true;
    ++passes_;
  }
  // term = timestep (discrete) or segment index (continuous).  Returns false when the joint values asked for are
  // not the ones of the memoised iterate, or when the batched pass holds no term for that timestep / segment (fixed).
  bool slice(CollisionCacheData& out, int term, const double* dof0, const double* dof1, bool continuous,
             bool vars0_fixed, bool vars1_fixed, std::size_t max_allowed)
  {
    std::lock_guard<std::mutex> lock(mutex_);
    const std::size_t nd = static_cast<std::size_t>(eval_.scene()->nDof());
    if (!valid_ || term < 0 || term + (continuous ? 1 : 0) >= T_)
      return false;
    if (std::memcmp(x_.data() + static_cast<std::size_t>(term) * nd, dof0, nd * sizeof(double)) != 0)
      return false;
    if (continuous && std::memcmp(x_.data() + static_cast<std::size_t>(term + 1) * nd, dof1, nd * sizeof(double)) != 0)
      return false;
    if (continuous && (vars0_fixed != (fixed_[static_cast<std::size_t>(term)] != 0) ||
                       vars1_fixed != (fixed_[static_cast<std::size_t>(term) + 1] != 0)))
      return false;
    if (continuous == (cfg_.type == CollisionEvaluatorType::DISCRETE))
      return false;  // a one-state evaluator on a two-state batch or the reverse
    const int k = term_of_[static_cast<std::size_t>(term)];
    if (k < 0 || static_cast<std::size_t>(k) + 1 >= host_.term_ptr.size())
      return false;
    detail::fillCacheData(out, *eval_.scene(), host_, host_.term_ptr[static_cast<std::size_t>(k)],
                          host_.term_ptr[static_cast<std::size_t>(k) + 1], continuous, vars0_fixed, vars1_fixed,
                          max_allowed);
    return true;
  }
  long passes() const { return passes_; }
  const TrajOptCollisionConfig& config() const { return cfg_; }

private:
  Evaluation eval_;
  TrajOptCollisionConfig cfg_;
  int T_;
  std::vector<uint8_t> fixed_;
  std::vector<int> term_of_;
  std::vector<double> x_;
  HostResult host_;
  bool valid_{ false };
  long passes_{ 0 };
  std::mutex mutex_;
};

// ---------------------------------------------------------------------------------------------
// seam B1, discrete
// ---------------------------------------------------------------------------------------------
class DiscreteCollisionEvaluator
{
public:
  using Ptr = std::shared_ptr<DiscreteCollisionEvaluator>;
  DiscreteCollisionEvaluator() = default;
  virtual ~DiscreteCollisionEvaluator() = default;
  DiscreteCollisionEvaluator(const DiscreteCollisionEvaluator&) = delete;
  DiscreteCollisionEvaluator& operator=(const DiscreteCollisionEvaluator&) = delete;
  virtual void calcCollisions(CollisionCacheData& collision_data, const VectorXd& dof_vals, std::size_t max_allowed) = 0;
  virtual double getCollisionMarginBuffer() const = 0;
  virtual const std::vector<Scene::Pair>& getCollisionMarginData() const = 0;  // margin per candidate pair
  virtual const std::vector<Scene::Pair>& getCollisionCoeffData() const = 0;   // coeff per candidate pair
  virtual const Scene& getJointGroup() const = 0;  // kinematics live in the scene
};

class SingleTimestepCollisionEvaluator : public DiscreteCollisionEvaluator
{
public:
  SingleTimestepCollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& collision_config,
                                   CollisionBatch::Ptr batch = nullptr, int timestep = -1)
    : eval_(scene), scene_(std::move(scene)), cfg_(collision_config), batch_(std::move(batch)), timestep_(timestep)
  {
    if (cfg_.type != CollisionEvaluatorType::DISCRETE)
      throw std::runtime_error("SingleTimestepCollisionEvaluator, should be configured with DISCRETE");
  }
  void calcCollisions(CollisionCacheData& collision_data, const VectorXd& dof_vals, std::size_t max_allowed) override final
  {
    if (static_cast<int>(dof_vals.size()) != scene_->nDof())
      throw std::runtime_error("SingleTimestepCollisionEvaluator: dof_vals has the wrong size");
    if (batch_ && batch_->slice(collision_data, timestep_, dof_vals.data(), nullptr, false, false, false, max_allowed))
      return;
    eval_.run(cfg_, dof_vals.data(), 1, 1, nullptr, host_);
    detail::fillCacheData(collision_data, *scene_, host_, host_.term_ptr[0], host_.term_ptr[1], false, false, false,
                          max_allowed);
  }
  double getCollisionMarginBuffer() const override final { return scene_->marginBuffer(); }
  const std::vector<Scene::Pair>& getCollisionMarginData() const override final { return scene_->pairs(); }
  const std::vector<Scene::Pair>& getCollisionCoeffData() const override final { return scene_->pairs(); }
  const Scene& getJointGroup() const override final { return *scene_; }

private:
  Evaluation eval_;
  Scene::Ptr scene_;
  TrajOptCollisionConfig cfg_;
  CollisionBatch::Ptr batch_;
  int timestep_;
  HostResult host_;
};

// ---------------------------------------------------------------------------------------------
// seam B1, continuous
// ---------------------------------------------------------------------------------------------
class ContinuousCollisionEvaluator
{
public:
  using Ptr = std::shared_ptr<ContinuousCollisionEvaluator>;
  ContinuousCollisionEvaluator() = default;
  virtual ~ContinuousCollisionEvaluator() = default;
  ContinuousCollisionEvaluator(const ContinuousCollisionEvaluator&) = delete;
  ContinuousCollisionEvaluator& operator=(const ContinuousCollisionEvaluator&) = delete;
  virtual void calcCollisionData(CollisionCacheData& collision_data, const VectorXd& dof_vals0, const VectorXd& dof_vals1,
                                 bool vars0_fixed, bool vars1_fixed, std::size_t max_allowed) = 0;
  virtual double getCollisionMarginBuffer() const = 0;
  virtual const std::vector<Scene::Pair>& getCollisionMarginData() const = 0;
  virtual const std::vector<Scene::Pair>& getCollisionCoeffData() const = 0;
  virtual const Scene& getJointGroup() const = 0;
};

namespace detail
{
class TwoStateEvaluator : public ContinuousCollisionEvaluator
{
public:
  TwoStateEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& cfg, CollisionBatch::Ptr batch, int segment)
    : eval_(scene), scene_(std::move(scene)), cfg_(cfg), batch_(std::move(batch)), segment_(segment)
  {
  }
  void calcCollisionData(CollisionCacheData& collision_data, const VectorXd& dof_vals0, const VectorXd& dof_vals1,
                         bool vars0_fixed, bool vars1_fixed, std::size_t max_allowed) override final
  {
    const std::size_t nd = static_cast<std::size_t>(scene_->nDof());
    if (dof_vals0.size() != nd || dof_vals1.size() != nd)
      throw std::runtime_error("ContinuousCollisionEvaluator: dof_vals have the wrong size");
    if (vars0_fixed && vars1_fixed)  // the reference's constraints refuse this case (continuous_collision_constraint.cpp:70,273)
      throw std::runtime_error("ContinuousCollisionEvaluator: both ends cannot be fixed!");
    if (batch_ && batch_->slice(collision_data, segment_, dof_vals0.data(), dof_vals1.data(), true, vars0_fixed,
                                vars1_fixed, max_allowed))
      return;
    q_.assign(dof_vals0.begin(), dof_vals0.end());
    q_.insert(q_.end(), dof_vals1.begin(), dof_vals1.end());
    const uint8_t fixed[2] = { static_cast<uint8_t>(vars0_fixed), static_cast<uint8_t>(vars1_fixed) };
    eval_.run(cfg_, q_.data(), 1, 2, fixed, host_);
    detail::fillCacheData(collision_data, *scene_, host_, host_.term_ptr[0], host_.term_ptr[1], true, vars0_fixed,
                          vars1_fixed, max_allowed);
  }
  double getCollisionMarginBuffer() const override final { return scene_->marginBuffer(); }
  const std::vector<Scene::Pair>& getCollisionMarginData() const override final { return scene_->pairs(); }
  const std::vector<Scene::Pair>& getCollisionCoeffData() const override final { return scene_->pairs(); }
  const Scene& getJointGroup() const override final { return *scene_; }

private:
  Evaluation eval_;
  Scene::Ptr scene_;
  TrajOptCollisionConfig cfg_;
  CollisionBatch::Ptr batch_;
  int segment_;
  std::vector<double> q_;
  HostResult host_;
};
}  // namespace detail

// continuous_collision_evaluators.cpp:41-95: swept (cast) test between the two states, LVS sub-division unless
// the config asks for CONTINUOUS (one sweep)
class LVSContinuousCollisionEvaluator : public detail::TwoStateEvaluator
{
public:
  LVSContinuousCollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& collision_config,
                                  CollisionBatch::Ptr batch = nullptr, int segment = -1)
    : TwoStateEvaluator(std::move(scene), checked(collision_config), std::move(batch), segment)
  {
  }

private:
  static const TrajOptCollisionConfig& checked(const TrajOptCollisionConfig& c)
  {
    if (c.type != CollisionEvaluatorType::CONTINUOUS && c.type != CollisionEvaluatorType::LVS_CONTINUOUS)
      throw std::runtime_error("LVSContinuousCollisionEvaluator, should be configured with CONTINUOUS or "
                               "LVS_CONTINUOUS");
    return c;
  }
};

// continuous_collision_evaluators.cpp:266-315: discrete tests at the LVS sub-states between the two states
class LVSDiscreteCollisionEvaluator : public detail::TwoStateEvaluator
{
public:
  LVSDiscreteCollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& collision_config,
                                CollisionBatch::Ptr batch = nullptr, int segment = -1)
    : TwoStateEvaluator(std::move(scene), checked(collision_config), std::move(batch), segment)
  {
  }

private:
  static const TrajOptCollisionConfig& checked(const TrajOptCollisionConfig& c)
  {
    if (c.type != CollisionEvaluatorType::LVS_DISCRETE)
      throw std::runtime_error("LVSDiscreteCollisionEvaluator, should be configured with LVS_DISCRETE");
    return c;
  }
};
// ---------------------------------------------------------------------------------------------
// SURVEY 8a A18 — numerical-gradient constraints, batched.
//   trajopt_ifopt/src/constraints/collision/discrete_collision_numerical_constraint.cpp:65-200
//   trajopt_ifopt/src/constraints/collision/continuous_collision_numerical_constraint.cpp:91-232
// The reference evaluates once at x and once per perturbed joint (delta = 1e-8) and matches the perturbed
// gradient-result sets to the baseline's by (key, shape_key).  The classes below run the n_dof + 1 (2 n_dof + 1)
// states as ONE tc_eval of independent "trajectories" and then follow the reference's matching loops.
// (An unchanged reference numerical constraint also works on top of the evaluators above: one small pass per call.)
// ---------------------------------------------------------------------------------------------
class DiscreteCollisionNumericalConstraint
{
public:
  DiscreteCollisionNumericalConstraint(Scene::Ptr scene, const TrajOptCollisionConfig& cfg, int max_num_cnt = 1)
    : eval_(scene), scene_(std::move(scene)), cfg_(cfg), rows_(static_cast<std::size_t>(std::max(max_num_cnt, 0)))
  {
    if (cfg_.type != CollisionEvaluatorType::DISCRETE)
      throw std::runtime_error("SingleTimestepCollisionEvaluator, should be configured with DISCRETE");
    if (max_num_cnt < 1)
      throw std::runtime_error("max_num_cnt must be greater than zero!");
  }
  // values[rows], coeffs[rows], jac[rows][n_dof] (row-major)
  void evaluate(const VectorXd& joint_vals, std::vector<double>& values, std::vector<double>& coeffs,
                std::vector<double>& jac)
  {
    const std::size_t n = static_cast<std::size_t>(scene_->nDof());
    if (joint_vals.size() != n)
      throw std::runtime_error("DiscreteCollisionNumericalConstraint: joint_vals has the wrong size");
    const double delta = 1e-8, mb = scene_->marginBuffer();
    q_.assign((n + 1) * n, 0.0);
    for (std::size_t b = 0; b <= n; ++b)
      for (std::size_t j = 0; j < n; ++j)
        q_[b * n + j] = joint_vals[j] + ((b == j + 1) ? delta : 0.0);
    eval_.run(cfg_, q_.data(), static_cast<int>(n + 1), 1, nullptr, host_);
    ++passes_;
    detail::fillCacheData(base_, *scene_, host_, host_.term_ptr[0], host_.term_ptr[1], false, false, false, rows_);
    values.assign(rows_, -mb);
    coeffs.assign(rows_, 1.0);
    jac.assign(rows_ * n, 0.0);
    const std::size_t cnt = std::min(rows_, base_.gradient_results_sets.size());
    for (std::size_t i = 0; i < cnt; ++i)
    {
      values[i] = base_.gradient_results_sets[i].getMaxErrorT0();
      coeffs[i] = base_.gradient_results_sets[i].coeff;
    }
    for (std::size_t j = 0; j < n && cnt > 0; ++j)
    {
      detail::fillCacheData(delta_, *scene_, host_, host_.term_ptr[j + 1], host_.term_ptr[j + 2], false, false, false,
                            rows_);
      for (std::size_t i = 0; i < cnt; ++i)
      {
        const GradientResultsSet& baseline = base_.gradient_results_sets[i];
        double e = -mb;
        for (const GradientResultsSet& cr : delta_.gradient_results_sets)
          if (cr.key == baseline.key && cr.shape_key == baseline.shape_key)
          {
            e = cr.getMaxErrorT0();
            break;
          }
        jac[i * n + j] = (e - baseline.getMaxErrorT0()) / delta;
      }
    }
  }
  long passes() const { return passes_; }

private:
  Evaluation eval_;
  Scene::Ptr scene_;
  TrajOptCollisionConfig cfg_;
  std::size_t rows_;
  std::vector<double> q_;
  HostResult host_;
  CollisionCacheData base_, delta_;
  long passes_{ 0 };
};

class ContinuousCollisionNumericalConstraint
{
public:
  ContinuousCollisionNumericalConstraint(Scene::Ptr scene, const TrajOptCollisionConfig& cfg, bool vars0_fixed,
                                         bool vars1_fixed, int max_num_cnt = 1)
    : eval_(scene)
    , scene_(std::move(scene))
    , cfg_(cfg)
    , f0_(vars0_fixed)
    , f1_(vars1_fixed)
    , rows_(static_cast<std::size_t>(std::max(max_num_cnt, 0)))
  {
    if (cfg_.type == CollisionEvaluatorType::DISCRETE || cfg_.type == CollisionEvaluatorType::NONE)
      throw std::runtime_error("ContinuousCollisionNumericalConstraint needs a two-state evaluator type");
    if (f0_ && f1_)
      throw std::runtime_error("position_vars are both fixed!");
    if (max_num_cnt < 1)
      throw std::runtime_error("max_num_cnt must be greater than zero!");
  }
  // values[rows], coeffs[rows], jac[rows][2][n_dof] (block p = state p)
  void evaluate(const VectorXd& joint_vals0, const VectorXd& joint_vals1, std::vector<double>& values,
                std::vector<double>& coeffs, std::vector<double>& jac)
  {
    const std::size_t n = static_cast<std::size_t>(scene_->nDof());
    if (joint_vals0.size() != n || joint_vals1.size() != n)
      throw std::runtime_error("ContinuousCollisionNumericalConstraint: joint_vals have the wrong size");
    const double delta = 1e-8, mb = scene_->marginBuffer();
    const std::size_t B = 2 * n + 1;
    q_.assign(B * 2 * n, 0.0);
    auto put = [&](std::size_t b, const VectorXd& a0, const VectorXd& a1) {
      for (std::size_t j = 0; j < n; ++j)
        q_[(b * 2) * n + j] = a0[j], q_[(b * 2 + 1) * n + j] = a1[j];
    };
    put(0, joint_vals0, joint_vals1);
    for (std::size_t p = 0; p < 2; ++p)
      for (std::size_t j = 0; j < n; ++j)
      {
        VectorXd jv = p ? joint_vals1 : joint_vals0;
        jv[j] += delta;
        put(1 + p * n + j, jv, joint_vals1);  // as the reference: (jv, joint_vals1) for both blocks (:182-183)
      }
    const uint8_t fixed[2] = { static_cast<uint8_t>(f0_), static_cast<uint8_t>(f1_) };
    eval_.run(cfg_, q_.data(), static_cast<int>(B), 2, fixed, host_);
    ++passes_;
    detail::fillCacheData(base_, *scene_, host_, host_.term_ptr[0], host_.term_ptr[1], true, f0_, f1_, rows_);
    values.assign(rows_, -mb);
    coeffs.assign(rows_, 1.0);
    jac.assign(rows_ * 2 * n, 0.0);
    const std::size_t cnt = std::min(rows_, base_.gradient_results_sets.size());
    auto err = [&](const GradientResultsSet& s) {
      return (!f0_ && !f1_) ? s.getMaxError() : (!f0_ ? s.getMaxErrorT0() : s.getMaxErrorT1());
    };
    for (std::size_t i = 0; i < cnt; ++i)
    {
      values[i] = err(base_.gradient_results_sets[i]);
      coeffs[i] = base_.gradient_results_sets[i].coeff;
    }
    for (std::size_t k = 0; k < 2 * n && cnt > 0; ++k)
    {
      detail::fillCacheData(delta_, *scene_, host_, host_.term_ptr[k + 1], host_.term_ptr[k + 2], true, f0_, f1_, rows_);
      for (std::size_t i = 0; i < cnt; ++i)
      {
        const GradientResultsSet& baseline = base_.gradient_results_sets[i];
        double e = -mb;
        for (const GradientResultsSet& cr : delta_.gradient_results_sets)
          if (cr.key == baseline.key && cr.shape_key == baseline.shape_key)
          {
            e = err(cr);
            break;
          }
        jac[(i * 2 + k / n) * n + k % n] = (e - err(baseline)) / delta;
      }
    }
  }
  long passes() const { return passes_; }

private:
  Evaluation eval_;
  Scene::Ptr scene_;
  TrajOptCollisionConfig cfg_;
  bool f0_, f1_;
  std::size_t rows_;
  std::vector<double> q_;
  HostResult host_;
  CollisionCacheData base_, delta_;
  long passes_{ 0 };
};

// SURVEY 8f N2 — the outcome tests' checkTrajectory (trajopt/test/planning_unit.cpp:99-100,145-148) as one batched
// pass: trajs[B][T][n_dof] -> minimum signed distance of every trajectory (+inf when nothing is within margin +
// buffer); a trajectory is in collision when that minimum is below `contact_distance`.
inline std::vector<bool> checkTrajectories(Evaluation& eval, const TrajOptCollisionConfig& cfg, const double* trajs, int B,
                                           int T, double contact_distance, std::vector<double>* min_distance = nullptr)
{
  HostResult h;
  eval.run(cfg, trajs, B, T, nullptr, h);
  std::vector<double> md(static_cast<std::size_t>(B), std::numeric_limits<double>::infinity());
  for (const tc_contact& c : h.contacts)
    md[static_cast<std::size_t>(c.traj)] = std::min(md[static_cast<std::size_t>(c.traj)], c.distance);
  std::vector<bool> hit(static_cast<std::size_t>(B));
  for (std::size_t b = 0; b < md.size(); ++b)
    hit[b] = md[b] < contact_distance;
  if (min_distance)
    *min_distance = md;
  return hit;
}
}  // namespace trajopt_cuda

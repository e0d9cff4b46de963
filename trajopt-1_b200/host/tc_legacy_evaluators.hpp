// tc_legacy_evaluators.hpp — host-side mirror of the legacy stack's collision evaluators over the C ABI.
//
// Seam B2 of SURVEY.md 8(b): trajopt::CollisionEvaluator and its three implementations
//   SingleTimestepCollisionEvaluator / DiscreteCollisionEvaluator / CastCollisionEvaluator
//       trajopt/include/trajopt/collision_terms.hpp:79-430, trajopt/src/collision_terms.cpp:600-1161
// with the two virtuals CollisionCost / CollisionConstraint call through m_calc:
//   CalcDistExpressions(const DblVec& x, sco::AffExprVector&, margins, coeffs)   (collision_terms.cpp:341-386,463-554)
//   CalcCollisions(const DblVec& x, ContactResultMap&)
// Same names, argument meaning and error behaviour (std::runtime_error on a mis-typed config, callee clears the
// caller-owned outputs).  sco::Var / sco::AffExpr are stood in for by the two small structs below.
// One difference is deliberate: the reference appends one block of coefficients per link to an expression
// (exprInc of varDot, :374), so a contact between two active links lists every variable twice; here the two
// blocks arrive merged per variable.  As affine functions they are equal; cleanupAff's 1e-7 pruning (:480,501,534,
// trajopt_sco/src/expr_ops.cpp:86-99) is applied to the merged coefficient.
#pragma once
#include <cmath>

#include "tc_evaluators.hpp"

namespace trajopt_cuda
{
namespace legacy
{
using DblVec = std::vector<double>;

// sco::Var: position of the variable in the optimiser's vector (var_rep->index)
struct Var
{
  int index{ -1 };
};
using VarVector = std::vector<Var>;

// sco::AffExpr
struct AffExpr
{
  double constant{ 0 };
  std::vector<double> coeffs;
  VarVector vars;
  std::size_t size() const { return coeffs.size(); }
  double value(const DblVec& x) const
  {
    double out = constant;
    for (std::size_t i = 0; i < coeffs.size(); ++i)
      out += coeffs[i] * x[static_cast<std::size_t>(vars[i].index)];
    return out;
  }
};
using AffExprVector = std::vector<AffExpr>;

// trajopt::CollisionExpressionEvaluatorType (collision_terms.hpp:28-34)
enum class CollisionExpressionEvaluatorType : std::uint8_t
{
  START_FREE_END_FREE = 0,
  START_FREE_END_FIXED = 1,
  START_FIXED_END_FREE = 2,
  SINGLE_TIME_STEP = 3
};

struct CollisionEvaluator
{
  using Ptr = std::shared_ptr<CollisionEvaluator>;
  virtual ~CollisionEvaluator() { tc_result_destroy(rows_res_); }
  virtual void CalcDistExpressions(const DblVec& x, AffExprVector& exprs, std::vector<double>& exprs_margin,
                                   std::vector<double>& exprs_coeff) = 0;
  virtual void CalcCollisions(const DblVec& x, ContactResultMap& dist_results) = 0;
  virtual VarVector GetVars() = 0;
  double getCollisionMarginBuffer() const { return scene_->marginBuffer(); }
  const std::vector<Scene::Pair>& getCollisionMarginData() const { return scene_->pairs(); }
  const std::vector<Scene::Pair>& getCollisionCoeffData() const { return scene_->pairs(); }

protected:
  CollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& cfg, VarVector vars0, VarVector vars1,
                     CollisionExpressionEvaluatorType type)
    : scene_(scene), cfg_(cfg), vars0_(std::move(vars0)), vars1_(std::move(vars1)), type_(type), eval_(scene)
  {
    const std::size_t nd = static_cast<std::size_t>(scene_->nDof());
    if (vars0_.size() != nd || (type != CollisionExpressionEvaluatorType::SINGLE_TIME_STEP && vars1_.size() != nd))
      throw std::runtime_error("CollisionEvaluator: one variable per joint expected");
    check(tc_result_create(scene_->handle(), &rows_res_), "tc_result_create");
  }

public:
  CollisionEvaluator(const CollisionEvaluator&) = delete;
  CollisionEvaluator& operator=(const CollisionEvaluator&) = delete;

protected:
  bool single() const { return type_ == CollisionExpressionEvaluatorType::SINGLE_TIME_STEP; }

  // joint values of the evaluator's own variables: q = [x(vars0), x(vars1)]
  void gather(const DblVec& x)
  {
    q_.clear();
    for (const Var& v : vars0_)
      q_.push_back(x.at(static_cast<std::size_t>(v.index)));
    if (!single())
      for (const Var& v : vars1_)
        q_.push_back(x.at(static_cast<std::size_t>(v.index)));
    fixed_[0] = type_ == CollisionExpressionEvaluatorType::START_FIXED_END_FREE;
    fixed_[1] = type_ == CollisionExpressionEvaluatorType::START_FREE_END_FIXED;
  }

  void collisions(const DblVec& x, ContactResultMap& dist_results)
  {
    gather(x);
    // the geometry pass itself does not depend on which end is free; the legacy CalcCollisions filters on
    // distance only unless an end is fixed (collision_terms.cpp:851-873), which tc_eval applies from fixed_steps
    eval_.run(legacyConfig(), q_.data(), 1, single() ? 1 : 2, single() ? nullptr : fixed_, host_, TC_STACK_LEGACY);
    CollisionCacheData tmp;
    detail::fillCacheData(tmp, *scene_, host_, host_.term_ptr.empty() ? 0 : host_.term_ptr[0],
                          host_.term_ptr.empty() ? 0 : host_.term_ptr.back(), !single(), false, false, 0);
    dist_results = tmp.contact_results_map;
  }

  void expressions(const DblVec& x, AffExprVector& exprs, std::vector<double>& exprs_margin,
                   std::vector<double>& exprs_coeff)
  {
    exprs.clear();
    exprs_margin.clear();
    exprs_coeff.clear();
    gather(x);
    const std::size_t nd = static_cast<std::size_t>(scene_->nDof());
    tc_eval_desc d{};
    d.evaluator = static_cast<int32_t>(cfg_.type);
    d.stack = TC_STACK_LEGACY;
    d.precision = cfg_.precision;
    d.want = TC_WANT_CONTACTS | TC_WANT_ROWS;
    d.lvs_length = cfg_.longest_valid_segment_length;
    d.fixed_steps = single() ? nullptr : fixed_;
    check(tc_eval(scene_->handle(), &d, q_.data(), 1, single() ? 1 : 2, rows_res_, nullptr), "tc_eval");
    tc_counts c{};
    check(tc_result_counts(rows_res_, &c), "tc_result_counts");
    const std::size_t n = static_cast<std::size_t>(c.n_contacts);
    row_ptr_.resize(n + 1);
    col_.resize(std::max<std::size_t>(1, 2 * n * nd));
    val_.resize(std::max<std::size_t>(1, 2 * n * nd));
    value_.resize(n), coeff_.resize(n), margin_.resize(n);
    check(tc_result_rows(rows_res_, row_ptr_.data(), col_.data(), val_.data(), value_.data(), coeff_.data(),
                         margin_.data()),
          "tc_result_rows");
    exprs.reserve(n);
    for (std::size_t i = 0; i < n; ++i)
    {
      AffExpr e;
      e.constant = value_[i];
      for (int64_t k = row_ptr_[i]; k < row_ptr_[i + 1]; ++k)
      {
        const double a = val_[static_cast<std::size_t>(k)];
        if (!(std::fabs(a) > 1e-7))  // cleanupAff
          continue;
        const std::size_t j = static_cast<std::size_t>(col_[static_cast<std::size_t>(k)]);
        e.coeffs.push_back(a);
        e.vars.push_back(j < nd ? vars0_[j] : vars1_[j - nd]);
      }
      exprs.push_back(std::move(e));
      exprs_margin.push_back(margin_[i]);
      exprs_coeff.push_back(coeff_[i]);
    }
  }

  TrajOptCollisionConfig legacyConfig() const { return cfg_; }

  Scene::Ptr scene_;
  TrajOptCollisionConfig cfg_;
  VarVector vars0_, vars1_;
  CollisionExpressionEvaluatorType type_;
  Evaluation eval_;
  tc_result* rows_res_{ nullptr };
  HostResult host_;
  std::vector<double> q_;
  uint8_t fixed_[2]{ 0, 0 };
  std::vector<int64_t> row_ptr_;
  std::vector<int32_t> col_;
  std::vector<double> val_, value_, coeff_, margin_;

};

namespace detail
{
inline VarVector concat(VarVector a, const VarVector& b)
{
  a.insert(a.end(), b.begin(), b.end());
  return a;
}
}  // namespace detail

// collision_terms.cpp:600-735: one state, discrete contact test
struct SingleTimestepCollisionEvaluator : public CollisionEvaluator
{
  SingleTimestepCollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& collision_config, VarVector vars,
                                   CollisionExpressionEvaluatorType type = CollisionExpressionEvaluatorType::SINGLE_TIME_STEP)
    : CollisionEvaluator(std::move(scene), checked(collision_config, type), std::move(vars), {}, type)
  {
  }
  void CalcDistExpressions(const DblVec& x, AffExprVector& exprs, std::vector<double>& exprs_margin,
                           std::vector<double>& exprs_coeff) override
  {
    expressions(x, exprs, exprs_margin, exprs_coeff);
  }
  void CalcCollisions(const DblVec& x, ContactResultMap& dist_results) override { collisions(x, dist_results); }
  VarVector GetVars() override { return vars0_; }

private:
  static const TrajOptCollisionConfig& checked(const TrajOptCollisionConfig& c, CollisionExpressionEvaluatorType t)
  {
    if (c.type != CollisionEvaluatorType::DISCRETE)
      throw std::runtime_error("SingleTimestepCollisionEvaluator, should be configured with DISCRETE");
    if (t != CollisionExpressionEvaluatorType::SINGLE_TIME_STEP)
      throw std::runtime_error("Invalid CollisionExpressionEvaluatorType for SingleTimestepCollisionEvaluator!");
    return c;
  }
};

// collision_terms.cpp:737-976: discrete tests at the LVS sub-states between two timesteps
struct DiscreteCollisionEvaluator : public CollisionEvaluator
{
  DiscreteCollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& collision_config, VarVector vars0,
                             VarVector vars1, CollisionExpressionEvaluatorType type)
    : CollisionEvaluator(std::move(scene), checked(collision_config, type), std::move(vars0), std::move(vars1), type)
  {
  }
  void CalcDistExpressions(const DblVec& x, AffExprVector& exprs, std::vector<double>& exprs_margin,
                           std::vector<double>& exprs_coeff) override
  {
    expressions(x, exprs, exprs_margin, exprs_coeff);
  }
  void CalcCollisions(const DblVec& x, ContactResultMap& dist_results) override { collisions(x, dist_results); }
  VarVector GetVars() override { return detail::concat(vars0_, vars1_); }

private:
  static const TrajOptCollisionConfig& checked(const TrajOptCollisionConfig& c, CollisionExpressionEvaluatorType t)
  {
    if (c.type != CollisionEvaluatorType::LVS_DISCRETE)
      throw std::runtime_error("DiscreteCollisionEvaluator, should be configured with LVS_DISCRETE");
    if (t == CollisionExpressionEvaluatorType::SINGLE_TIME_STEP)
      throw std::runtime_error("Invalid CollisionExpressionEvaluatorType for DiscreteCollisionEvaluator!");
    return c;
  }
};

// collision_terms.cpp:978-1229: swept (cast) tests between two timesteps
struct CastCollisionEvaluator : public CollisionEvaluator
{
  CastCollisionEvaluator(Scene::Ptr scene, const TrajOptCollisionConfig& collision_config, VarVector vars0, VarVector vars1,
                         CollisionExpressionEvaluatorType type)
    : CollisionEvaluator(std::move(scene), checked(collision_config, type), std::move(vars0), std::move(vars1), type)
  {
  }
  void CalcDistExpressions(const DblVec& x, AffExprVector& exprs, std::vector<double>& exprs_margin,
                           std::vector<double>& exprs_coeff) override
  {
    expressions(x, exprs, exprs_margin, exprs_coeff);
  }
  void CalcCollisions(const DblVec& x, ContactResultMap& dist_results) override { collisions(x, dist_results); }
  VarVector GetVars() override { return detail::concat(vars0_, vars1_); }

private:
  static const TrajOptCollisionConfig& checked(const TrajOptCollisionConfig& c, CollisionExpressionEvaluatorType t)
  {
    if (c.type != CollisionEvaluatorType::CONTINUOUS && c.type != CollisionEvaluatorType::LVS_CONTINUOUS)
      throw std::runtime_error("CastCollisionEvaluator, should be configured with CONTINUOUS or LVS_CONTINUOUS");
    if (t == CollisionExpressionEvaluatorType::SINGLE_TIME_STEP)
      throw std::runtime_error("Invalid CollisionExpressionEvaluatorType for CastCollisionEvaluator!");
    return c;
  }
};
}  // namespace legacy
}  // namespace trajopt_cuda

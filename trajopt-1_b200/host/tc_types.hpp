// tc_types.hpp — stand-ins for the tesseract / trajopt_common types that cross the evaluator seam.
//
// tesseract, Eigen and Boost are not available where this repository is built, so the adaptor in
// tc_evaluators.hpp is written against the small types below.  Each one names the reference type it
// stands in for; field names and meanings are the reference's, so that switching the `using`s to the
// real types is mechanical.  Nothing here runs on the hot path: these records are
// only filled from the arrays libtrajopt_cuda.so returns.
#pragma once
#include <algorithm>
#include <array>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <utility>
#include <vector>

#include "../../include/trajopt_cuda.h"

namespace trajopt_cuda
{
using VectorXd = std::vector<double>;            // Eigen::VectorXd
using Isometry3d = std::array<double, 12>;       // Eigen::Isometry3d, top 3x4 row-major
using LinkIdPair = std::pair<int32_t, int32_t>;  // tesseract::common::LinkIdPair (scene link indices, first < second)

// tesseract::collision::ContinuousCollisionType
enum class ContinuousCollisionType : int32_t
{
  CCType_None = TC_CC_NONE,
  CCType_Time0 = TC_CC_TIME0,
  CCType_Time1 = TC_CC_TIME1,
  CCType_Between = TC_CC_BETWEEN
};

// tesseract::collision::ContactResult — the fields the trajopt path reads
// (trajopt/src/collision_terms.cpp:105-124,215-223)
struct ContactResult
{
  double distance{ 0 };
  std::array<int32_t, 2> link_ids{ -1, -1 };
  std::array<int32_t, 2> shape_id{ -1, -1 };
  std::array<int32_t, 2> subshape_id{ -1, -1 };
  std::array<std::array<double, 3>, 2> nearest_points{};
  std::array<std::array<double, 3>, 2> nearest_points_local{};
  std::array<Isometry3d, 2> transform{};
  std::array<Isometry3d, 2> cc_transform{};
  std::array<double, 3> normal{};
  std::array<double, 2> cc_time{ -1, -1 };
  std::array<ContinuousCollisionType, 2> cc_type{ ContinuousCollisionType::CCType_None,
                                                  ContinuousCollisionType::CCType_None };
};

// tesseract::collision::ContactResultMap: ordered by link pair, contact order within the pair kept
class ContactResultMap
{
public:
  using Container = std::map<LinkIdPair, std::vector<ContactResult>>;
  void clear() { data_.clear(); }
  std::vector<ContactResult>& operator[](const LinkIdPair& k) { return data_[k]; }
  Container::const_iterator begin() const { return data_.begin(); }
  Container::const_iterator end() const { return data_.end(); }
  std::size_t size() const { return data_.size(); }
  bool empty() const { return data_.empty(); }
  // total number of contacts over all pairs (ContactResultMap::count)
  std::size_t count() const
  {
    std::size_t n = 0;
    for (const auto& kv : data_)
      n += kv.second.size();
    return n;
  }

private:
  Container data_;
};

// trajopt_common::LinkGradientResults (trajopt_common/include/trajopt_common/collision_types.h:187-209).
// `jacobian` (3 x n) is not materialised: the device produces gradient = jacobian^T * translation_vector
// directly; a consumer that needs the Jacobian itself is outside the scoped path.
struct LinkGradientResults
{
  bool has_gradient{ false };
  VectorXd gradient;
  std::array<double, 3> translation_vector{};
  double scale{ 1.0 };
  ContinuousCollisionType cc_type{ ContinuousCollisionType::CCType_None };
};

// trajopt_common::GradientResults (collision_types.h:212-235)
struct GradientResults
{
  std::array<LinkGradientResults, 2> gradients;
  std::array<LinkGradientResults, 2> cc_gradients;
  double error{ 0 };
  double error_with_buffer{ 0 };
};

// trajopt_common::LinkMaxError (collision_types.h:237-270)
struct LinkMaxError
{
  std::array<bool, 2> has_error{ false, false };
  std::array<double, 2> error{ std::numeric_limits<double>::lowest(), std::numeric_limits<double>::lowest() };
  std::array<double, 2> error_with_buffer{ std::numeric_limits<double>::lowest(),
                                           std::numeric_limits<double>::lowest() };
  double getMaxError() const { return pick(error); }
  double getMaxErrorWithBuffer() const { return pick(error_with_buffer); }

private:
  double pick(const std::array<double, 2>& e) const
  {
    if (has_error[0] && has_error[1])
      return std::max(e[0], e[1]);
    if (has_error[0])
      return e[0];
    if (has_error[1])
      return e[1];
    throw std::runtime_error("Invalid LinkMaxError");
  }
};

// trajopt_common::GradientResultsSet (collision_types.h:273-328, collision_types.cpp:173-333)
struct GradientResultsSet
{
  LinkIdPair key{ -1, -1 };
  std::pair<std::size_t, std::size_t> shape_key{ 0, 0 };
  double coeff{ 1 };
  bool is_continuous{ false };
  std::array<LinkMaxError, 2> max_error;
  std::vector<GradientResults> results;

  // updates the per-link maxima: slot [0] ignores contacts that only exist at T1, slot [1] those at T0
  void add(GradientResults g)
  {
    for (std::size_t i = 0; i < 2; ++i)
    {
      const LinkGradientResults& lg = g.gradients[i];
      if (!lg.has_gradient)
        continue;
      if (lg.cc_type != ContinuousCollisionType::CCType_Time1)
        raise(max_error[i], 0, g);
      if (lg.cc_type != ContinuousCollisionType::CCType_Time0)
        raise(max_error[i], 1, g);
    }
    results.push_back(std::move(g));
  }
  double getMaxError() const { return get(2, false, "Max error does not exist."); }
  double getMaxErrorT0() const { return get(0, false, "Max error at T0 does not exist."); }
  double getMaxErrorT1() const { return get(1, false, "Max error at T1 does not exist."); }
  double getMaxErrorWithBuffer() const { return get(2, true, "Max error with buffer does not exist."); }
  double getMaxErrorWithBufferT0() const { return get(0, true, "Max error with buffer at T0 does not exist."); }
  double getMaxErrorWithBufferT1() const { return get(1, true, "Max error with buffer at T1 does not exist."); }

private:
  static void raise(LinkMaxError& m, std::size_t k, const GradientResults& g)
  {
    m.has_error[k] = true;
    m.error[k] = std::max(m.error[k], g.error);
    m.error_with_buffer[k] = std::max(m.error_with_buffer[k], g.error_with_buffer);
  }
  // which: 0 = slot T0, 1 = slot T1, 2 = both
  double get(int which, bool buffer, const char* what) const
  {
    double e = std::numeric_limits<double>::lowest();
    bool found = false;
    for (std::size_t i = 0; i < 2; ++i)
      for (std::size_t k = 0; k < 2; ++k)
      {
        if ((which != 2 && which != static_cast<int>(k)) || !max_error[i].has_error[k])
          continue;
        const double v = buffer ? max_error[i].error_with_buffer[k] : max_error[i].error[k];
        if (!found || v > e)
          e = v;
        found = true;
      }
    if (!found)
      throw std::runtime_error(what);
    return e;
  }
};

// trajopt_common::CollisionCacheData (collision_types.h:331-338)
struct CollisionCacheData
{
  using Ptr = std::shared_ptr<CollisionCacheData>;
  ContactResultMap contact_results_map;
  std::vector<GradientResultsSet> gradient_results_sets;
};

// trajopt_common::cantorHash (trajopt_common/src/collision_utils.cpp:64-71)
inline std::size_t cantorHash(int shape_id, int subshape_id)
{
  if (subshape_id < 0)
    return static_cast<std::size_t>(shape_id);
  return static_cast<std::size_t>(0.5 * (shape_id + subshape_id) * (shape_id + subshape_id + 1) + subshape_id);
}

// tesseract::collision::CollisionEvaluatorType
enum class CollisionEvaluatorType : int32_t
{
  NONE = 0,
  DISCRETE = TC_EVAL_DISCRETE,
  LVS_DISCRETE = TC_EVAL_LVS_DISCRETE,
  CONTINUOUS = TC_EVAL_CONTINUOUS,
  LVS_CONTINUOUS = TC_EVAL_LVS_CONTINUOUS
};

// trajopt_common::TrajOptCollisionConfig reduced to what reaches the device
// (trajopt_common/include/trajopt_common/collision_types.h:144-184); the margin / coeff tables themselves
// are baked into the tc_scene at creation
struct TrajOptCollisionConfig
{
  CollisionEvaluatorType type{ CollisionEvaluatorType::DISCRETE };
  double longest_valid_segment_length{ 0.005 };
  double collision_margin_buffer{ 0.01 };
  int max_num_cnt{ 3 };
  int precision{ TC_PRECISION_F64 };
};
}  // namespace trajopt_cuda

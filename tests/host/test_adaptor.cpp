// test_adaptor.cpp — drives the host adaptor (trajopt-1_b200/host/tc_evaluators.hpp) the way the reference's
// constraints drive their evaluators, and prints the fixed-size constraint values / Jacobian rows so that the
// pytest wrapper (tests/test_host_adaptor.py) can compare them with the reference's known answers
// (trajopt_ifopt/test/collision_unit.cpp:97-159) and with the CPU oracle.
//
// The two helpers below restate what the reference's constraints do with a CollisionCacheData
// (discrete_collision_constraint.cpp:91-141, continuous_collision_constraint.cpp:107-236,
// weighted_average_methods.cpp:31-107); in a trajopt build that code is the reference's own and unchanged.
#include <cmath>
#include <cstdio>
#include <cstdlib>

#include "../../trajopt-1_b200/host/tc_evaluators.hpp"
#include "../../trajopt-1_b200/host/tc_legacy_evaluators.hpp"
#include "../../trajopt-1_b200/host/tc_scene_file.hpp"

using namespace trajopt_cuda;

static VectorXd weightedAvg(const GradientResultsSet& set, double max_error_with_buffer, std::size_t n, bool t1)
{
  VectorXd grad(n, 0.0);
  double total = 0;
  long cnt = 0;
  for (const GradientResults& g : set.results)
    for (std::size_t i = 0; i < 2; ++i)
    {
      const LinkGradientResults& lg = t1 ? g.cc_gradients[i] : g.gradients[i];
      const auto skip = t1 ? ContinuousCollisionType::CCType_Time0 : ContinuousCollisionType::CCType_Time1;
      if (!lg.has_gradient || lg.cc_type == skip || !(set.max_error[i].error_with_buffer[t1 ? 1 : 0] > 0))
        continue;
      const double w = std::max(g.error_with_buffer, 0.0) / max_error_with_buffer;
      total += w;
      for (std::size_t j = 0; j < n; ++j)
        grad[j] += w * lg.scale * lg.gradient[j];
      ++cnt;
    }
  if (cnt)
    for (double& v : grad)
      v /= total;
  return grad;
}

struct Rows
{
  std::vector<double> values, jac0, jac1;  // [rows], [rows][n], [rows][n]
};

static Rows discreteRows(const CollisionCacheData& d, std::size_t rows, std::size_t n, double margin_buffer)
{
  Rows r{ std::vector<double>(rows, -margin_buffer), std::vector<double>(rows * n, 0.0), {} };
  const std::size_t cnt = std::min(rows, d.gradient_results_sets.size());
  for (std::size_t i = 0; i < cnt; ++i)
  {
    const GradientResultsSet& s = d.gradient_results_sets[i];
    r.values[i] = s.getMaxErrorT0();
    const VectorXd g = weightedAvg(s, s.getMaxErrorWithBufferT0(), n, false);
    for (std::size_t j = 0; j < n; ++j)
      r.jac0[i * n + j] = -g[j];
  }
  return r;
}

static Rows continuousRows(const CollisionCacheData& d, std::size_t rows, std::size_t n, double margin_buffer, bool f0,
                           bool f1)
{
  Rows r{ std::vector<double>(rows, -margin_buffer), std::vector<double>(rows * n, 0.0),
          std::vector<double>(rows * n, 0.0) };
  const std::size_t cnt = std::min(rows, d.gradient_results_sets.size());
  for (std::size_t i = 0; i < cnt; ++i)
  {
    const GradientResultsSet& s = d.gradient_results_sets[i];
    const bool has0 = s.max_error[0].has_error[0] || s.max_error[1].has_error[0];
    const bool has1 = s.max_error[0].has_error[1] || s.max_error[1].has_error[1];
    if (!f0 && !f1)
      r.values[i] = s.getMaxError();
    else if (!f0)
    {
      if (has0)
        r.values[i] = s.getMaxErrorT0();
    }
    else if (has1)
      r.values[i] = s.getMaxErrorT1();
    if (!f0 && has0)
    {
      const VectorXd g = weightedAvg(s, f1 ? s.getMaxErrorWithBufferT0() : s.getMaxErrorWithBuffer(), n, false);
      for (std::size_t j = 0; j < n; ++j)
        r.jac0[i * n + j] = -g[j];
    }
    if (!f1 && has1)
    {
      const VectorXd g = weightedAvg(s, f0 ? s.getMaxErrorWithBufferT1() : s.getMaxErrorWithBuffer(), n, true);
      for (std::size_t j = 0; j < n; ++j)
        r.jac1[i * n + j] = -g[j];
    }
  }
  return r;
}

static void printRows(const char* tag, int term, const CollisionCacheData& d, const Rows& r, std::size_t n)
{
  std::printf("%s %d contacts %zu sets %zu values", tag, term, d.contact_results_map.count(),
              d.gradient_results_sets.size());
  for (double v : r.values)
    std::printf(" %.17g", v);
  std::printf(" jac0");
  for (double v : r.jac0)
    std::printf(" %.17g", v);
  if (!r.jac1.empty())
  {
    std::printf(" jac1");
    for (double v : r.jac1)
      std::printf(" %.17g", v);
  }
  std::printf("\n");
  (void)n;
}

static bool same(const Rows& a, const Rows& b)
{
  return a.values == b.values && a.jac0 == b.jac0 && a.jac1 == b.jac1;
}

// legacy stack: one line per affine expression, coefficients scattered over the optimiser's full variable vector
static void printExprs(const char* tag, int term, legacy::CollisionEvaluator& ev, const legacy::DblVec& x)
{
  legacy::AffExprVector exprs;
  std::vector<double> margins, coeffs;
  ev.CalcDistExpressions(x, exprs, margins, coeffs);
  ContactResultMap contacts;
  ev.CalcCollisions(x, contacts);
  std::printf("%s %d exprs %zu contacts %zu vars %zu\n", tag, term, exprs.size(), contacts.count(), ev.GetVars().size());
  for (std::size_t i = 0; i < exprs.size(); ++i)
  {
    std::vector<double> dense(x.size(), 0.0);
    for (std::size_t k = 0; k < exprs[i].size(); ++k)
      dense[static_cast<std::size_t>(exprs[i].vars[k].index)] += exprs[i].coeffs[k];
    std::printf("%sX %d %zu %.17g %.17g %.17g %.17g coeffs", tag, term, i, exprs[i].constant, margins[i], coeffs[i],
                exprs[i].value(x));
    for (double v : dense)
      std::printf(" %.17g", v);
    std::printf("\n");
  }
}

int main(int argc, char** argv)
{
  if (argc < 3)
  {
    std::fprintf(stderr, "usage: test_adaptor boxbot.scene arm_around_table.scene\n");
    return 2;
  }
  try
  {
    // ---- boxbot known answers: margin 0.1, coeff 1, buffer 0.01, DiscreteCollisionConstraint(max_num_cnt = 1)
    {
      SceneFile file(argv[1]);
      auto scene = std::make_shared<Scene>(file.desc(0.1, 1.0, 0.01));
      TrajOptCollisionConfig cfg;
      cfg.type = CollisionEvaluatorType::DISCRETE;
      cfg.collision_margin_buffer = 0.01;
      SingleTimestepCollisionEvaluator ev(scene, cfg);
      const double states[4][2] = { { -1.9, 0.01 }, { -1.0, 0.01 }, { 0.01, 1.0 }, { -0.9, 0.01 } };
      CollisionCacheData data;
      for (int k = 0; k < 4; ++k)
      {
        ev.calcCollisions(data, VectorXd{ states[k][0], states[k][1] }, 1);
        printRows("KAT", k, data, discreteRows(data, 1, 2, ev.getCollisionMarginBuffer()), 2);
      }
      // error behaviour of the reference constructors: a mis-typed config throws std::runtime_error
      bool threw = false;
      try
      {
        TrajOptCollisionConfig bad = cfg;
        bad.type = CollisionEvaluatorType::LVS_CONTINUOUS;
        SingleTimestepCollisionEvaluator nope(scene, bad);
      }
      catch (const std::runtime_error&)
      {
        threw = true;
      }
      std::printf("BADCONFIG %d\n", threw ? 1 : 0);
      // legacy stack, same four states: SingleTimestepCollisionEvaluator::CalcDistExpressions
      {
        legacy::SingleTimestepCollisionEvaluator lev(scene, cfg, legacy::VarVector{ { 0 }, { 1 } });
        for (int k = 0; k < 4; ++k)
          printExprs("LEGS", k, lev, legacy::DblVec{ states[k][0], states[k][1] });
        bool lthrew = false;
        try
        {
          legacy::CastCollisionEvaluator nope(scene, cfg, legacy::VarVector{ { 0 }, { 1 } }, legacy::VarVector{ { 2 }, { 3 } },
                                              legacy::CollisionExpressionEvaluatorType::START_FREE_END_FREE);
        }
        catch (const std::runtime_error&)
        {
          lthrew = true;
        }
        std::printf("LEGACY_BADCONFIG %d\n", lthrew ? 1 : 0);
      }
      // numerical-gradient constraint (A18) on the same four states: one batched pass each
      {
        DiscreteCollisionNumericalConstraint num(scene, cfg, 1);
        std::vector<double> values, coeffs, jac;
        for (int k = 0; k < 4; ++k)
        {
          num.evaluate(VectorXd{ states[k][0], states[k][1] }, values, coeffs, jac);
          std::printf("NUMKAT %d value %.12g coeff %.12g jac %.9g %.9g passes %ld\n", k, values[0], coeffs[0], jac[0], jac[1],
                      num.passes());
        }
        bool nthrew = false;
        try
        {
          ContinuousCollisionNumericalConstraint nope(scene, cfg, false, false, 1);
        }
        catch (const std::runtime_error&)
        {
          nthrew = true;
        }
        std::printf("NUM_BADCONFIG %d\n", nthrew ? 1 : 0);
      }
      // max_allowed == 0: contacts only
      ev.calcCollisions(data, VectorXd{ -0.9, 0.01 }, 0);
      std::printf("NOSETS contacts %zu sets %zu\n", data.contact_results_map.count(), data.gradient_results_sets.size());
    }
    // ---- DISCRETE batch with a fixed first timestep: tc_eval skips it, so timestep t is term t - 1 of the pass
    {
      SceneFile file(argv[2]);
      auto scene = std::make_shared<Scene>(file.desc(0.025, 20.0, 0.01));
      const std::size_t n = static_cast<std::size_t>(scene->nDof());
      const std::vector<double>& traj = file.f64("seed_traj");
      const int T = static_cast<int>(traj.size() / n);
      TrajOptCollisionConfig cfg;
      cfg.type = CollisionEvaluatorType::DISCRETE;
      std::vector<uint8_t> fx(static_cast<std::size_t>(T), 0);
      fx[0] = 1;
      auto batch = std::make_shared<CollisionBatch>(scene, cfg, T, fx);
      batch->update(traj.data());
      int bad = 0;
      std::size_t total = 0;
      for (int t = 0; t < T; ++t)
      {
        const VectorXd q(traj.begin() + static_cast<long>(static_cast<std::size_t>(t) * n),
                         traj.begin() + static_cast<long>(static_cast<std::size_t>(t + 1) * n));
        CollisionCacheData alone, sliced;
        SingleTimestepCollisionEvaluator ea(scene, cfg), eb(scene, cfg, batch, t);
        ea.calcCollisions(alone, q, 3);
        eb.calcCollisions(sliced, q, 3);
        total += alone.contact_results_map.count();
        if (!same(discreteRows(alone, 3, n, scene->marginBuffer()), discreteRows(sliced, 3, n, scene->marginBuffer())) ||
            alone.contact_results_map.count() != sliced.contact_results_map.count())
          ++bad;
      }
      std::printf("BATCH_DISCRETE_FIXED mismatches %d contacts %zu passes %ld\n", bad, total, batch->passes());
    }
    // ---- arm_around_table seed trajectory: LVS_CONTINUOUS and LVS_DISCRETE per segment, standalone and batched
    {
      SceneFile file(argv[2]);
      auto scene = std::make_shared<Scene>(file.desc(0.025, 20.0, 0.01));
      const std::size_t n = static_cast<std::size_t>(scene->nDof());
      const std::vector<double>& traj = file.f64("seed_traj");
      const int T = static_cast<int>(traj.size() / n);
      std::vector<uint8_t> fixed(static_cast<std::size_t>(T), 0);
      fixed.front() = fixed.back() = 1;
      const std::size_t max_num_cnt = 3;
      for (int kind = 0; kind < 2; ++kind)
      {
        TrajOptCollisionConfig cfg;
        cfg.type = kind ? CollisionEvaluatorType::LVS_DISCRETE : CollisionEvaluatorType::LVS_CONTINUOUS;
        cfg.longest_valid_segment_length = 0.02;
        auto batch = std::make_shared<CollisionBatch>(scene, cfg, T, fixed);
        batch->update(traj.data());
        batch->update(traj.data());  // same iterate: memoised
        for (int t = 0; t + 1 < T; ++t)
        {
          const VectorXd q0(traj.begin() + static_cast<long>(t * n), traj.begin() + static_cast<long>((t + 1) * n));
          const VectorXd q1(traj.begin() + static_cast<long>((t + 1) * n), traj.begin() + static_cast<long>((t + 2) * n));
          const bool f0 = fixed[static_cast<std::size_t>(t)] != 0, f1 = fixed[static_cast<std::size_t>(t) + 1] != 0;
          CollisionCacheData alone, sliced;
          std::unique_ptr<ContinuousCollisionEvaluator> ev_alone, ev_batch;
          if (kind)
          {
            ev_alone.reset(new LVSDiscreteCollisionEvaluator(scene, cfg));
            ev_batch.reset(new LVSDiscreteCollisionEvaluator(scene, cfg, batch, t));
          }
          else
          {
            ev_alone.reset(new LVSContinuousCollisionEvaluator(scene, cfg));
            ev_batch.reset(new LVSContinuousCollisionEvaluator(scene, cfg, batch, t));
          }
          ev_alone->calcCollisionData(alone, q0, q1, f0, f1, max_num_cnt);
          ev_batch->calcCollisionData(sliced, q0, q1, f0, f1, max_num_cnt);
          const Rows ra = continuousRows(alone, max_num_cnt, n, scene->marginBuffer(), f0, f1);
          const Rows rb = continuousRows(sliced, max_num_cnt, n, scene->marginBuffer(), f0, f1);
          printRows(kind ? "LVSD" : "LVSC", t, alone, ra, n);
          if (!same(ra, rb) || alone.contact_results_map.count() != sliced.contact_results_map.count())
          {
            std::printf("BATCH_MISMATCH %d %d\n", kind, t);
            return 1;
          }
        }
        std::printf("BATCH_PASSES %d %ld\n", kind, batch->passes());
        // fixed steps that make tc_eval skip terms: the first two timesteps fixed -> segment 0 has no term and segment
        // t >= 1 is term t - 1 of the batched pass; every sliced segment must equal the standalone evaluator's, and
        // the skipped one must fall back to a standalone pass (the evaluator itself still answers)
        {
          std::vector<uint8_t> fx(static_cast<std::size_t>(T), 0);
          fx[0] = fx[1] = 1;
          auto b2 = std::make_shared<CollisionBatch>(scene, cfg, T, fx);
          b2->update(traj.data());
          int bad = 0;
          for (int t = 1; t + 1 < T; ++t)
          {
            const VectorXd q0(traj.begin() + static_cast<long>(t * n), traj.begin() + static_cast<long>((t + 1) * n));
            const VectorXd q1(traj.begin() + static_cast<long>((t + 1) * n), traj.begin() + static_cast<long>((t + 2) * n));
            const bool f0 = fx[static_cast<std::size_t>(t)] != 0, f1 = fx[static_cast<std::size_t>(t) + 1] != 0;
            CollisionCacheData alone, sliced;
            std::unique_ptr<ContinuousCollisionEvaluator> ea, eb;
            if (kind)
              ea.reset(new LVSDiscreteCollisionEvaluator(scene, cfg)), eb.reset(new LVSDiscreteCollisionEvaluator(scene, cfg, b2, t));
            else
              ea.reset(new LVSContinuousCollisionEvaluator(scene, cfg)), eb.reset(new LVSContinuousCollisionEvaluator(scene, cfg, b2, t));
            ea->calcCollisionData(alone, q0, q1, f0, f1, max_num_cnt);
            eb->calcCollisionData(sliced, q0, q1, f0, f1, max_num_cnt);
            if (!same(continuousRows(alone, max_num_cnt, n, scene->marginBuffer(), f0, f1),
                      continuousRows(sliced, max_num_cnt, n, scene->marginBuffer(), f0, f1)) ||
                alone.contact_results_map.count() != sliced.contact_results_map.count())
              ++bad;
          }
          CollisionCacheData none;
          const bool skipped_refused = !b2->slice(none, 0, traj.data(), traj.data() + n, true, true, true, max_num_cnt);
          std::printf("BATCH_SKIPPED_TERMS %d mismatches %d skipped_refused %d passes %ld\n", kind, bad, skipped_refused ? 1 : 0,
                      b2->passes());
        }
        // numerical-gradient constraint on segment (1, 2), both ends free, and the batched trajectory check
        {
          const VectorXd q1(traj.begin() + static_cast<long>(n), traj.begin() + static_cast<long>(2 * n));
          const VectorXd q2(traj.begin() + static_cast<long>(2 * n), traj.begin() + static_cast<long>(3 * n));
          ContinuousCollisionNumericalConstraint num(scene, cfg, false, false, static_cast<int>(max_num_cnt));
          std::vector<double> values, coeffs, jac;
          num.evaluate(q1, q2, values, coeffs, jac);
          std::printf("%s values", kind ? "NUMD" : "NUMC");
          for (double v : values)
            std::printf(" %.12g", v);
          std::printf(" jac");
          for (double v : jac)
            std::printf(" %.9g", v);
          std::printf(" passes %ld\n", num.passes());
          Evaluation ev(scene);
          std::vector<double> md;
          const std::vector<bool> hit = checkTrajectories(ev, cfg, traj.data(), 1, T, 0.0, &md);
          std::printf("%s in_collision %d min_distance %.12g\n", kind ? "CHECKD" : "CHECKC", hit[0] ? 1 : 0, md[0]);
        }
        // legacy stack: CastCollisionEvaluator / DiscreteCollisionEvaluator per segment, variables of the whole
        // trajectory in one vector as the legacy problem lays them out (trajopt/src/problem_description.cpp)
        const legacy::DblVec x(traj.begin(), traj.end());
        for (int t = 0; t + 1 < T; ++t)
        {
          legacy::VarVector v0, v1;
          for (std::size_t j = 0; j < n; ++j)
          {
            v0.push_back({ static_cast<int>(static_cast<std::size_t>(t) * n + j) });
            v1.push_back({ static_cast<int>(static_cast<std::size_t>(t + 1) * n + j) });
          }
          using ET = legacy::CollisionExpressionEvaluatorType;
          const ET type = fixed[static_cast<std::size_t>(t)] ? ET::START_FIXED_END_FREE :
                          fixed[static_cast<std::size_t>(t) + 1] ? ET::START_FREE_END_FIXED : ET::START_FREE_END_FREE;
          if (kind)
          {
            legacy::DiscreteCollisionEvaluator lev(scene, cfg, v0, v1, type);
            printExprs("LEGD", t, lev, x);
          }
          else
          {
            legacy::CastCollisionEvaluator lev(scene, cfg, v0, v1, type);
            printExprs("LEGC", t, lev, x);
          }
        }
      }
    }
  }
  catch (const std::exception& e)
  {
    std::fprintf(stderr, "test_adaptor: %s\n", e.what());
    return 1;
  }
  return 0;
}

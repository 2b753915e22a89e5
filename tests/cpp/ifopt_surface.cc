// Exercises include/towr_b200_ifopt.hpp (TEST INFRASTRUCTURE).  `host`: everything that needs no GPU.  `gpu`: also the
// batched evaluation; prints checksums that tests/test_cpp_surface.py compares with the Python binding and the oracle.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <numeric>

#include "towr_b200_ifopt.hpp"

using namespace towr_b200;

static NlpFormulation Make(double gx, bool durations) {
  NlpFormulation f;
  f.SetRobot(TOWR_B200_ANYMAL);
  f.SetGait(1, 2.0);
  f.SetNominalStates(gx, 0.0, 0.0);
  if (durations) f.OptimizePhaseDurations();
  f.AddCost(TOWR_B200_COST_FORCES, 1e-9);      // hopper_example.cc:147-148
  f.AddCost(TOWR_B200_COST_EE_MOTION, 1e-4);
  return f;
}

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  const bool durations = argc > 2 && std::strcmp(argv[2], "durations") == 0;
  try {
    std::vector<NlpFormulation> problems = {Make(1.0, durations), Make(1.5, durations), Make(0.7, durations)};
    BatchedProblem nlp(problems);
    std::printf("sizes %d %d %d\n", nlp.GetNumberOfOptimizationVariables(), nlp.GetNumberOfConstraints(),
                nlp.GetNumberOfJacobianNonzeros());
    int cols = 0, rows = 0;
    for (const auto& n : nlp.GetVariableSetNames()) cols += nlp.GetVariableSet(0, n).GetRows();
    for (const auto& n : nlp.GetConstraintSetNames()) rows += nlp.GetConstraintSet(0, n).GetRows();
    if (cols != nlp.GetNumberOfOptimizationVariables() || rows != nlp.GetNumberOfConstraints()) throw std::runtime_error("set sizes");
    // bounds: per-set views agree with the problem-level vectors; goals differ between problems
    VecBound vb = nlp.GetBoundsOnOptimizationVariables(1);
    VecBound lin = nlp.GetVariableSet(1, "base-lin").GetBounds();
    for (size_t i = 0; i < lin.size(); ++i)
      if (lin[i].lower_ != vb[i].lower_ || lin[i].upper_ != vb[i].upper_) throw std::runtime_error("bounds view");
    VecBound cb = nlp.GetBoundsOnConstraints(0);
    double bsum = 0.0;
    for (const auto& b : cb) bsum += (std::fabs(b.lower_) < 1e19 ? b.lower_ : 0.0) + (std::fabs(b.upper_) < 1e19 ? b.upper_ : 0.0);
    VectorXd x0 = nlp.GetVariableValues(2);
    std::printf("x0sum %.17g boundsum %.17g\n", std::accumulate(x0.begin(), x0.end(), 0.0), bsum);
    // SetVariables through a set view lands in the problem vector
    auto force = nlp.GetVariableSet(2, "ee-force_1");
    VectorXd f = force.GetValues();
    f[3] += 1.0;
    force.SetVariables(f);
    if (nlp.GetVariableSet(2, "ee-force_1").GetValues()[3] != f[3]) throw std::runtime_error("SetVariables view");
    bool threw = false;
    try { nlp.GetConstraintSet(0, "no-such-set"); } catch (const std::runtime_error&) { threw = true; }
    if (!threw) throw std::runtime_error("missing set must throw");
    if (gpu) {
      VectorXd x = nlp.GetVariableValues(1);
      for (size_t i = 0; i < x.size(); ++i) x[i] += 0.01 * std::sin(0.37 * static_cast<double>(i));
      VectorXd g = nlp.EvaluateConstraints(1, x.data());
      const double* v = nlp.GetJacobianOfConstraints(1);
      double gs = 0.0, vs = 0.0;
      for (double a : g) gs += a;
      for (int e = 0; e < nlp.GetNumberOfJacobianNonzeros(); ++e) vs += v[e];
      std::printf("gsum %.17g jacsum %.17g\n", gs, vs);
      // the blocks of one constraint set partition its rows' entries
      auto dyn = nlp.GetConstraintSet(1, "dynamic");
      VectorXd gd = dyn.GetValues();
      size_t n_block = 0;
      double block_sum = 0.0;
      for (const auto& name : nlp.GetVariableSetNames()) {
        std::vector<Triplet> blk;
        dyn.FillJacobianBlock(name, blk);
        n_block += blk.size();
        for (const auto& t : blk) {
          if (t.row < 0 || t.row >= dyn.GetRows()) throw std::runtime_error("block row out of range");
          block_sum += t.value;
        }
      }
      std::printf("dynamic rows %d entries %zu blocksum %.17g g0 %.17g\n", dyn.GetRows(), n_block, block_sum, gd[0]);
    }
    if (gpu) {
      // eval_f / eval_grad_f and the trajectory rows of problem 1 at its current variables
      VectorXd x = nlp.GetVariableValues(1);
      const double f = nlp.EvaluateCostFunction(1, x.data());
      VectorXd gr = nlp.EvaluateCostFunctionGradient(1, x.data());
      double grs = 0.0;
      for (double a : gr) grs += a;
      std::printf("cost %.17g gradsum %.17g\n", f, grs);
      const std::vector<double> rows = nlp.SampleTrajectories(0.05);
      const size_t per = static_cast<size_t>(nlp.GetTrajectorySamples(0.05)) * nlp.GetTrajectoryWidth();
      double rs = 0.0;
      for (size_t i = 0; i < per; ++i) rs += rows[per + i];
      std::printf("trajectory %d %d %.17g\n", nlp.GetTrajectorySamples(0.05), nlp.GetTrajectoryWidth(), rs);
    }
    std::printf("ok\n");
    return 0;
  } catch (const std::exception& e) {
    std::printf("FAILED: %s\n", e.what());
    return 1;
  }
}

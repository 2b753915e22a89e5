// TEST INFRASTRUCTURE.  The reference's own example, towr/test/hopper_example.cc:93-161, stated through the
// reference-named types of include/towr_b200_formulation.hpp (same members, same statements), converted with ToSpec and
// evaluated through the ifopt-surface mirror; every quantity is compared ENTRY BY ENTRY with the oracle built from the
// same spec: x0, variable and constraint bounds, the Jacobian pattern (host), g, the Jacobian values, eval_f,
// eval_grad_f and the SplineHolder view's GetPoint / IsContactPhase (gpu).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "towr_b200_ifopt.hpp"
extern "C" {
#include "../../oracle/oracle_capi.h"
}

using namespace towr_b200;
using namespace towr_b200::ref;
static_assert(sizeof(orc_spec) == sizeof(towr_b200_spec), "the two specs share one layout");

static double Ratio(double a, double b) { return std::fabs(a - b) / (1e-14 + 1e-12 * std::fabs(b)); }
static double Clip(double v) { return std::max(-1e20, std::min(1e20, v)); }

static ref::NlpFormulation Hopper() {
  ref::NlpFormulation formulation;
  // terrain: the five-step stairs defined in the example itself (hopper_example.cc:53-91)
  formulation.terrain_ = HeightMap::MakeTerrain(HeightMap::FiveStepStairsID);
  // Kinematic limits and dynamic parameters of the hopper
  formulation.model_ = RobotModel(RobotModel::Monoped);
  formulation.params_.ee_phase_durations_.push_back({0.5, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4, 0.3, 0.4});
  formulation.params_.ee_stance_position_.push_back({{0.0, 0.0}, {0.4, 0.0}, {0.7, -0.0}, {0.7, 0.0}, {1.3, -0.0}, {1.3, 0.0}, {0.0, 0.0}});
  // set the initial position of the hopper
  formulation.initial_base_.lin.at(kPos) = {0.0, 0.0, 0.6};
  formulation.initial_base_.lin.at(kVel) = {0.0, 0.0, 0.0};
  formulation.initial_base_.ang.at(kPos) = {0.0, 0.0, 0.0};
  formulation.initial_base_.ang.at(kVel) = {0.0, 0.0, 0.0};
  formulation.initial_ee_W_.push_back(Vector3d{0.0, 0.0, 0.0});
  // define the desired goal state of the hopper: FiveStepStairs().GetHeight(0.0, 0.0) = 0
  formulation.final_base_.lin.at(kPos) = {0.0, 0.0, 0.6 + 0.0};
  formulation.final_base_.lin.at(kVel) = {0.0, 0.0, 0.0};
  formulation.final_base_.ang.at(kPos) = {0.0, 0.0, 0.0};
  formulation.final_base_.ang.at(kVel) = {0.0, 0.0, 0.0};
  formulation.params_.ee_in_contact_at_start_.push_back(true);
  formulation.params_.constraints_.push_back(Parameters::Torque);
  formulation.params_.costs_.push_back(std::make_pair(Parameters::ForcesCostID, 1e-9));
  formulation.params_.costs_.push_back(std::make_pair(Parameters::EEMotionCostID, 1e-4));
  formulation.params_.OptimizePhaseDurations();
  return formulation;
}

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  try {
    ref::NlpFormulation formulation = Hopper();
    // what the reference throws for, ToSpec throws for: stance tracking without positions (nlp_formulation.cc:224)
    {
      ref::NlpFormulation bad = Hopper();
      bad.params_.ee_stance_position_.clear();
      bool threw = false;
      try { ToSpec(bad); } catch (const std::out_of_range&) { threw = true; }
      if (!threw) throw std::runtime_error("missing ee_stance_position_ must throw std::out_of_range");
      bad = Hopper();
      bad.params_.angular_rep_ = Parameters::RotationVector;
      threw = false;
      try { ToSpec(bad); } catch (const std::runtime_error& e) { threw = std::strncmp(e.what(), "unsupported", 11) == 0; }
      if (!threw) throw std::runtime_error("RotationVector must be refused as unsupported");
    }
    const towr_b200_spec spec = ToSpec(formulation);
    orc_spec ospec;
    std::memcpy(&ospec, &spec, sizeof(spec));
    orc_problem* op = orc_create(&ospec);
    if (!op) throw std::runtime_error(std::string("oracle: ") + orc_last_error());

    std::vector<towr_b200::NlpFormulation> problems = {towr_b200::NlpFormulation(formulation)};
    BatchedProblem nlp(problems);
    const int nv = nlp.GetNumberOfOptimizationVariables(), nr = nlp.GetNumberOfConstraints(), nnz = nlp.GetNumberOfJacobianNonzeros();
    if (nv != orc_n_vars(op) || nr != orc_n_rows(op)) throw std::runtime_error("sizes differ from the oracle");
    // x0 and bounds, entry by entry
    std::vector<double> x0(nv), lo(nv), up(nv), clo(nr), cup(nr);
    orc_x0(op, x0.data());
    orc_var_bounds(op, lo.data(), up.data());
    orc_con_bounds(op, clo.data(), cup.data());
    VectorXd mine = nlp.GetVariableValues(0);
    VecBound vb = nlp.GetBoundsOnOptimizationVariables(0), cb = nlp.GetBoundsOnConstraints(0);
    double worst = 0.0;
    for (int i = 0; i < nv; ++i) {
      worst = std::max(worst, Ratio(mine[i], x0[i]));
      worst = std::max(worst, std::max(Ratio(Clip(vb[i].lower_), Clip(lo[i])), Ratio(Clip(vb[i].upper_), Clip(up[i]))));
    }
    for (int i = 0; i < nr; ++i)
      worst = std::max(worst, std::max(Ratio(Clip(cb[i].lower_), Clip(clo[i])), Ratio(Clip(cb[i].upper_), Clip(cup[i]))));
    if (worst > 1.0) throw std::runtime_error("x0 / bounds differ from the oracle");
    // the stance bounds of the example are there: x of the second stance is pinned to 0.4
    int pinned = 0;
    for (int i = 0; i < nv; ++i) pinned += (vb[i].lower_ == 0.4 && vb[i].upper_ == 0.4);
    if (pinned < 1) throw std::runtime_error("stance-tracking bounds missing");
    // pattern at a perturbed x
    std::vector<double> x(x0);
    for (int i = 0; i < nv; ++i) x[i] += 0.01 * std::sin(0.37 * i);
    const int onnz = orc_nnz(op, x.data());
    if (onnz != nnz) throw std::runtime_error("stored-entry count differs from the oracle");
    std::vector<int> orow(nnz), ocol(nnz);
    std::vector<double> oval(nnz), og(nr);
    orc_eval_jac(op, x.data(), orow.data(), ocol.data(), oval.data());
    if (orow != nlp.JacobianRows() || ocol != nlp.JacobianCols()) throw std::runtime_error("pattern differs from the oracle");
    std::printf("host_checks sizes %d %d %d worst_ratio %.3g\n", nv, nr, nnz, worst);
    if (gpu) {
      // the class optimises durations: strict band against the oracle with correctly rounded powers
      orc_set_pow_mode(1);
      orc_eval_g(op, x.data(), og.data());
      orc_eval_jac(op, x.data(), orow.data(), ocol.data(), oval.data());
      VectorXd g = nlp.EvaluateConstraints(0, x.data());
      const double* v = nlp.GetJacobianOfConstraints(0);
      double wg = 0.0, wv = 0.0;
      for (int i = 0; i < nr; ++i) wg = std::max(wg, Ratio(g[i], og[i]));
      for (int e = 0; e < nnz; ++e) wv = std::max(wv, Ratio(v[e], oval[e]));
      const double f = nlp.EvaluateCostFunction(0, x.data()), of = orc_eval_f(op, x.data());
      VectorXd gr = nlp.EvaluateCostFunctionGradient(0, x.data());
      std::vector<double> ogr(nv);
      orc_eval_grad_f(op, x.data(), ogr.data());
      double wc = Ratio(f, of);
      for (int i = 0; i < nv; ++i) wc = std::max(wc, Ratio(gr[i], ogr[i]));
      // the solution read-out loop of hopper_example.cc:170-186 through the SplineHolder view
      SplineHolder solution = nlp.GetSplineHolder(0);
      const int W = orc_trajectory_width(op);
      const int n_samples = orc_sample_trajectory(op, x.data(), 0.2, nullptr);
      std::vector<double> rows(static_cast<size_t>(n_samples) * W);
      orc_sample_trajectory(op, x.data(), 0.2, rows.data());
      double ws = 0.0;
      int contact_mismatch = 0;
      double t = 0.0;
      for (int k = 0; k < n_samples; ++k) {
        const double* r = rows.data() + static_cast<size_t>(k) * W;
        if (r[0] != t) throw std::runtime_error("sample time differs");
        for (int d = 0; d < 3; ++d) {
          ws = std::max(ws, Ratio(solution.base_linear_->GetPoint(t).p()[d], r[1 + d]));
          ws = std::max(ws, Ratio(solution.base_angular_->GetPoint(t).p()[d], r[10 + d]));
          ws = std::max(ws, Ratio(solution.ee_motion_.at(0)->GetPoint(t).p()[d], r[19 + d]));
          ws = std::max(ws, Ratio(solution.ee_motion_.at(0)->GetPoint(t).v()[d], r[22 + d]));
          ws = std::max(ws, Ratio(solution.ee_force_.at(0)->GetPoint(t).p()[d], r[19 + 18 + d]));
          ws = std::max(ws, Ratio(solution.ee_torque_.at(0)->GetPoint(t).p()[d], r[19 + 21 + d]));
        }
        contact_mismatch += (solution.phase_durations_.at(0)->IsContactPhase(t) ? 1.0 : 0.0) != r[19 + 24];
        t += 0.2;
      }
      std::vector<double> pd = solution.phase_durations_.at(0)->GetPhaseDurations();
      double pd_sum = 0.0;
      for (double d : pd) pd_sum += d;
      bool force_v_threw = false;
      try { solution.ee_force_.at(0)->GetPoint(0.0).v(); } catch (const std::out_of_range&) { force_v_threw = true; }
      std::printf("gpu_checks g %.3g jac %.3g cost %.3g splines %.3g contact_mismatch %d samples %d phases %zu duration_sum %.17g force_v_threw %d\n",
                  wg, wv, wc, ws, contact_mismatch, n_samples, pd.size(), pd_sum, force_v_threw ? 1 : 0);
    }
    orc_destroy(op);
    std::printf("ok\n");
    return 0;
  } catch (const std::exception& e) {
    std::printf("FAILED: %s\n", e.what());
    return 1;
  }
}

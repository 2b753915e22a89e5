// towr_b200_spline_holder.hpp — towr::SplineHolder (include/towr/variables/spline_holder.h:48-84) as a view over the
// batched trajectory sampler.
//
// After a solve the reference reads the solution through the SplineHolder it got from GetVariableSets():
//     solution.base_linear_->GetPoint(t).p()            test/hopper_example.cc:170-186
//     solution.ee_motion_.at(ee)->GetPoint(t).p()
//     solution.ee_force_.at(ee)->GetPoint(t).p()
//     solution.phase_durations_.at(ee)->IsContactPhase(t)
// The members below keep those names and calls.  GetPoint(t) is answered by towr_b200_batch_sample_points (the same
// spline arithmetic as the evaluation kernel: spline.cc:48-93, polynomial.cc:47-104, phase_spline.cc:54-65 for optimised
// durations); one call evaluates all splines of the problem at t and the row is kept, so the usual loop that reads
// several splines at the same t launches once per t.  Prefetch(times) fetches a whole list of times in one launch.
// There is no CPU fallback: without the CUDA backend the first GetPoint throws.
#pragma once
#include <array>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

#include "towr_b200.h"

namespace towr_b200 {

class SplineHolder {
 public:
  using Vector3d = std::array<double, 3>;
  enum Dx { kPos = 0, kVel, kAcc };

  // towr::State (variables/state.h:49-104).  Force and torque splines carry the value only (the sampler's row has no
  // columns for their derivatives, as SaveTrajectoryToCSV writes none): v(), a() throw for them.
  class State {
   public:
    const Vector3d& at(Dx d) const {
      if (static_cast<int>(d) >= n_) throw std::out_of_range("State::at: derivative not sampled for this spline");
      return x_[d];
    }
    const Vector3d& p() const { return at(kPos); }
    const Vector3d& v() const { return at(kVel); }
    const Vector3d& a() const { return at(kAcc); }

   private:
    friend class SplineHolder;
    std::array<Vector3d, 3> x_{};
    int n_ = 0;
  };

 private:
  struct Impl {
    towr_b200_batch* batch;
    const towr_b200_structure* st;
    towr_b200_spec spec;
    std::vector<double> x;
    int width;
    double total_time;
    std::map<double, std::vector<double>> rows;          // sampled rows by global time
    const std::vector<double>& Row(double t) {
      auto it = rows.find(t);
      if (it != rows.end()) return it->second;
      std::vector<double> r(static_cast<size_t>(width));
      if (towr_b200_batch_sample_points(batch, 1, &t, 1, x.data(), r.data()) != TOWR_B200_OK)
        throw std::runtime_error(towr_b200_last_error());
      return rows.emplace(t, std::move(r)).first->second;
    }
  };

 public:
  // towr::NodeSpline as far as a reader of the solution uses it (node_spline.h / spline.h): GetPoint, GetTotalTime
  class NodeSpline {
   public:
    using Ptr = std::shared_ptr<NodeSpline>;
    const State GetPoint(double t_global) const {
      const std::vector<double>& r = impl_->Row(t_global);
      State s;
      s.n_ = n_out_ / 3;
      for (int d = 0; d < s.n_; ++d)
        for (int k = 0; k < 3; ++k) s.x_[d][k] = r[col_ + 3 * d + k];
      return s;
    }
    double GetTotalTime() const { return impl_->total_time; }

   private:
    friend class SplineHolder;
    NodeSpline(std::shared_ptr<Impl> impl, int col, int n_out) : impl_(std::move(impl)), col_(col), n_out_(n_out) {}
    std::shared_ptr<Impl> impl_;
    int col_, n_out_;
  };

  // towr::PhaseDurations as a reader uses it (phase_durations.h): IsContactPhase, GetPhaseDurations
  class PhaseDurations {
   public:
    using Ptr = std::shared_ptr<PhaseDurations>;
    bool IsContactPhase(double t) const { return impl_->Row(t)[19 + 25 * ee_ + 24] != 0.0; }
    // the foot's phase durations for the current variables: x's ee-schedule entries, the last one t_total_ minus their
    // sum (phase_durations.cc:79-88); the class's fixed durations when they are not optimised
    std::vector<double> GetPhaseDurations() const {
      const towr_b200_spec& s = impl_->spec;
      std::vector<double> d(s.phase_durations[ee_], s.phase_durations[ee_] + s.n_phases[ee_]);
      if (schedule_col0_ < 0) return d;
      double total = 0.0, sum = 0.0;
      for (double v : d) total += v;                     // std::accumulate over the initial timings (phase_durations.cc:45)
      for (size_t i = 0; i + 1 < d.size(); ++i) {
        d[i] = impl_->x[schedule_col0_ + static_cast<int>(i)];
        sum += d[i];
      }
      d.back() = total - sum;
      return d;
    }

   private:
    friend class SplineHolder;
    PhaseDurations(std::shared_ptr<Impl> impl, int ee, int schedule_col0) : impl_(std::move(impl)), ee_(ee), schedule_col0_(schedule_col0) {}
    std::shared_ptr<Impl> impl_;
    int ee_, schedule_col0_;
  };

  // the solution of ONE problem: `x` (n_vars doubles, e.g. what Ipopt returned) for the problem described by `spec`,
  // evaluated on `batch` (any batch of the spec's structure class; the holder keeps its own copy of x)
  SplineHolder(towr_b200_batch* batch, const towr_b200_structure* st, const towr_b200_spec& spec, const double* x) {
    if (!batch || !st || !x) throw std::runtime_error("SplineHolder: null argument");
    auto impl = std::make_shared<Impl>();
    impl->batch = batch;
    impl->st = st;
    impl->spec = spec;
    impl->x.assign(x, x + towr_b200_structure_n_vars(st));
    impl->width = towr_b200_trajectory_width(st);
    {                                                    // Spline::GetTotalTime of the base spline (parameters.cc:114-130)
      double T = 0.0;
      for (int i = 0; i < spec.n_phases[0]; ++i) T += spec.phase_durations[0][i];
      double t_left = T, sum = 0.0;
      while (t_left > 1e-10) {
        sum += t_left > spec.duration_base_polynomial ? spec.duration_base_polynomial : t_left;
        t_left -= spec.duration_base_polynomial;
      }
      impl->total_time = sum;
    }
    impl_ = impl;
    base_linear_.reset(new NodeSpline(impl, 1, 9));
    base_angular_.reset(new NodeSpline(impl, 10, 9));
    std::vector<int> schedule_col0(spec.n_ee, -1);
    int col = 0;
    for (int i = 0; i < towr_b200_structure_n_var_sets(st); ++i) {
      const std::string name = towr_b200_structure_var_set_name(st, i);
      for (int e = 0; e < spec.n_ee; ++e)
        if (name == "ee-schedule" + std::to_string(e)) schedule_col0[e] = col;   // variable_names.h id::EESchedule
      col += towr_b200_structure_var_set_rows(st, i);
    }
    for (int e = 0; e < spec.n_ee; ++e) {
      const int c0 = 19 + 25 * e;                        // row layout of towr_b200_batch_sample_trajectory
      ee_motion_.emplace_back(new NodeSpline(impl, c0, 9));
      ee_ang_.emplace_back(new NodeSpline(impl, c0 + 9, 9));
      ee_force_.emplace_back(new NodeSpline(impl, c0 + 18, 3));
      ee_torque_.emplace_back(new NodeSpline(impl, c0 + 21, 3));
      phase_durations_.emplace_back(new PhaseDurations(impl, e, schedule_col0[e]));
    }
  }

  // one launch for all of `times`; later GetPoint / IsContactPhase calls at these times are served from the result
  void Prefetch(const std::vector<double>& times) {
    if (times.empty()) return;
    std::vector<double> rows(times.size() * static_cast<size_t>(impl_->width));
    if (towr_b200_batch_sample_points(impl_->batch, 1, times.data(), static_cast<int>(times.size()), impl_->x.data(),
                                      rows.data()) != TOWR_B200_OK)
      throw std::runtime_error(towr_b200_last_error());
    for (size_t k = 0; k < times.size(); ++k)
      impl_->rows[times[k]].assign(rows.begin() + k * impl_->width, rows.begin() + (k + 1) * impl_->width);
  }

  NodeSpline::Ptr base_linear_;
  NodeSpline::Ptr base_angular_;
  std::vector<NodeSpline::Ptr> ee_motion_;
  std::vector<NodeSpline::Ptr> ee_ang_;
  std::vector<NodeSpline::Ptr> ee_force_;
  std::vector<NodeSpline::Ptr> ee_torque_;
  std::vector<PhaseDurations::Ptr> phase_durations_;

 private:
  std::shared_ptr<Impl> impl_;
};

}  // namespace towr_b200

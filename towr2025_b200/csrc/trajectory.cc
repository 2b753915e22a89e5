// trajectory.cc — host tables of the trajectory sampler (trajectory.h).
#include "trajectory.h"

#include <cmath>
#include <numeric>
#include <stdexcept>

#include "structure.h"

namespace tb {

std::vector<double> TrajectoryTimes(const HostProgram& hp, double dt) {
  if (!(dt > 0.0)) throw std::runtime_error("the sampling period must be positive");
  const VarSet* lin = hp.Find(kBaseLin);
  // Spline::GetTotalTime (spline.cc:118-123): accumulate over the polynomial durations of base_linear_
  const double T_total = std::accumulate(lin->poly_T.begin(), lin->poly_T.end(), 0.0);
  std::vector<double> ts;
  double t = 0.0;
  while (t <= T_total + 1e-9) {
    ts.push_back(t);
    t += dt;
    if (ts.size() > 10000000) throw std::runtime_error("the sampling period is too small");
  }
  return ts;
}

void BuildTrajTables(const HostProgram& hp, const towr_b200_spec& s, double dt, TrajTables& o) {
  BuildTrajTablesAt(hp, s, TrajectoryTimes(hp, dt), o);
  o.dt = dt;
}

void BuildTrajTablesAt(const HostProgram& hp, const towr_b200_spec& s, const std::vector<double>& times, TrajTables& o) {
  o = TrajTables{};
  o.n_vars = hp.n_vars;
  o.n_ee = hp.n_ee;
  o.width = TrajectoryWidth(hp.n_ee);
  o.rt = hp.optimize_durations;
  o.total_time = hp.total_time;
  o.dt = 0.0;
  // Spline::GetSegmentID asserts t_global >= 0 and runs off the end of the spline beyond its total time (spline.cc:48-66)
  const VarSet* lin = hp.Find(kBaseLin);
  const double T_total = std::accumulate(lin->poly_T.begin(), lin->poly_T.end(), 0.0);
  for (double t : times)
    if (!(t >= 0.0) || !(t <= T_total + 1e-9)) throw std::runtime_error("sample time outside [0, total time]");
  o.times = times;
  o.n_samples = static_cast<int>(o.times.size());
  auto add = [&](const VarSet& v, int out_col, int n_out, int foot) {
    TrajSplineRec r{};
    r.node0 = o.n_nodes;
    r.poly0 = o.n_polys;
    r.n_polys = v.n_polys();
    r.out_col = out_col;
    r.n_out = n_out;
    r.foot = foot;
    r.rt = (foot >= 0 && o.rt) ? 1 : 0;
    std::vector<int> src(static_cast<size_t>(v.n_nodes) * 6, 0xFFFF);
    for (int idx = 0; idx < v.n_vars; ++idx)
      for (const NodeRef& nr : v.refs[idx]) src[nr.node * 6 + nr.deriv * 3 + nr.dim] = v.col0 + idx;
    for (int x : src) o.node_src.push_back(static_cast<uint16_t>(x));
    for (int p = 0; p < v.n_polys(); ++p) {
      const double T = v.poly_T[p];
      o.poly_T.push_back(T);
      o.poly_T.push_back(RuntimePow(T, 2));
      o.poly_T.push_back(RuntimePow(T, 3));
      o.poly_spline.push_back(static_cast<uint16_t>(o.splines.size()));
      o.poly_meta.push_back(r.rt ? (static_cast<uint32_t>(v.poly_phase[p]) | (static_cast<uint32_t>(v.polys_in_phase[p]) << 8))
                                 : kPolyFixed);
    }
    // fixed durations: segment and local time of every sample (spline.cc:48-78), powers with std::pow as
    // Polynomial::GetDerivativeWrtCoeff (polynomial.cc:63-72)
    for (int k = 0; k < o.n_samples; ++k) {
      if (r.rt) { o.samp_poly.push_back(0); o.samp_t.insert(o.samp_t.end(), 3, 0.0); continue; }
      const double tg = o.times[k];
      const int id = SegmentId(tg, v.poly_T);
      double tl = tg;
      for (int i = 0; i < id; ++i) tl -= v.poly_T[i];
      o.samp_poly.push_back(static_cast<uint32_t>(r.poly0 + id));
      o.samp_t.push_back(tl);
      o.samp_t.push_back(RuntimePow(tl, 2));
      o.samp_t.push_back(RuntimePow(tl, 3));
    }
    o.n_nodes += v.n_nodes;
    o.n_polys += v.n_polys();
    o.splines.push_back(r);
  };
  add(*hp.Find(kBaseLin), 1, 9, -1);
  add(*hp.Find(kBaseAng), 10, 9, -1);
  for (int e = 0; e < hp.n_ee; ++e) {
    const int c0 = 19 + 25 * e;
    add(*hp.Find(kEEMotion, e), c0, 9, e);
    add(*hp.Find(kEEAng, e), c0 + 9, 9, e);
    add(*hp.Find(kEEForce, e), c0 + 18, 3, e);
    add(*hp.Find(kEETorque, e), c0 + 21, 3, e);
  }
  for (int e = 0; e < hp.n_ee; ++e) {
    TrajFootRec f{};
    f.phase0 = o.n_phases;
    f.n_phases = s.n_phases[e];
    f.col0 = o.rt ? hp.Find(kEESchedule, e)->col0 : 0;
    f.contact_at_start = s.in_contact_at_start[e] != 0;
    for (int i = 0; i < s.n_phases[e]; ++i) f.total += s.phase_durations[e][i];   // phase_durations.cc:40-50
    o.feet.push_back(f);
    o.n_phases += s.n_phases[e];
    // fixed durations: PhaseDurations::IsContactPhase (phase_durations.cc:113-118) on the class's phase durations
    std::vector<double> pd(s.phase_durations[e], s.phase_durations[e] + s.n_phases[e]);
    for (int k = 0; k < o.n_samples; ++k) {
      const int ph = o.rt ? 0 : SegmentId(o.times[k], pd);
      o.contact.push_back(static_cast<uint8_t>((ph % 2 == 0) ? f.contact_at_start : !f.contact_at_start));
    }
  }
}

}  // namespace tb

// trajectory.h — batched trajectory sampling: every spline of a problem's SplineHolder evaluated on a time grid, one
// CSV row of towr/src/utils/save_data.cpp:9-132 (SaveTrajectoryToCSV) per sample.  The step on the far side of the
// solver; same spline arithmetic as the evaluation kernel (spline.cc:48-93, polynomial.cc:47-104).
#pragma once
#include <cstdint>
#include <vector>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#endif

#include "program.h"
#include "../../include/towr_b200.h"

namespace tb {

struct TrajSplineRec {   // one spline of the SplineHolder, CSV column order
  int node0, poly0, n_polys;
  int out_col, n_out;    // first column of the row, 9 (pos, vel, acc) or 3 (pos)
  int foot;              // -1: base spline
  int rt;                // 1: polynomial durations follow the ee-schedule variables (PhaseSpline)
  int pad;
};
struct TrajFootRec { int phase0, n_phases, col0, contact_at_start; double total; };   // total: this foot's own t_total_

struct TrajTables {      // host side, one per (structure class, dt)
  int n_vars = 0, n_ee = 0, width = 0, n_samples = 0, n_nodes = 0, n_polys = 0, n_phases = 0;
  bool rt = false;
  double total_time = 0.0, dt = 0.0;
  std::vector<double> times;
  std::vector<TrajSplineRec> splines;
  std::vector<TrajFootRec> feet;
  std::vector<uint16_t> node_src;   // 6 per node, 0xFFFF = not optimised (0)
  std::vector<double> poly_T;       // 3 per polynomial: T, pow(T,2), pow(T,3)
  std::vector<uint32_t> poly_meta;  // phase | polys_in_phase << 8 ; kPolyFixed: class-constant duration
  std::vector<uint16_t> poly_spline;// spline of every polynomial
  std::vector<uint32_t> samp_poly;  // fixed durations: [spline][sample] polynomial (global index)
  std::vector<double> samp_t;       // fixed durations: [spline][sample] t_local, pow(t,2), pow(t,3)
  std::vector<uint8_t> contact;     // fixed durations: [foot][sample] PhaseDurations::IsContactPhase
};

struct TrajDevice {
  int n_vars, n_ee, width, n_samples, n_splines, n_nodes, n_polys, n_phases, rt;
  double total_time;
  const double* times;
  const TrajSplineRec* splines;
  const TrajFootRec* feet;
  const uint16_t* node_src;
  const double* poly_T;
  const uint32_t* poly_meta;
  const uint16_t* poly_spline;
  const uint32_t* samp_poly;
  const double* samp_t;
  const uint8_t* contact;
};

inline int TrajectoryWidth(int n_ee) { return 19 + 25 * n_ee; }
// sample times of SaveTrajectoryToCSV: t = 0, t += dt while t <= T_total + 1e-9 (save_data.cpp:55,118)
std::vector<double> TrajectoryTimes(const HostProgram& hp, double dt);
void BuildTrajTables(const HostProgram& hp, const towr_b200_spec& s, double dt, TrajTables& out);
// the same for an explicit list of global times in [0, T_total] (SplineHolder::...->GetPoint(t) at arbitrary t)
void BuildTrajTablesAt(const HostProgram& hp, const towr_b200_spec& s, const std::vector<double>& times, TrajTables& out);
size_t TrajSharedBytes(const TrajDevice& t);
#ifdef __CUDACC__
cudaError_t LaunchSampleTrajectory(const TrajDevice& t, const double* x, double* out, int n_problems, int sms,
                                   cudaStream_t stream);
#endif

}  // namespace tb

// structure.h — internal host API of the structure-class builder.
#pragma once
#include <vector>

#include "../../include/towr_b200.h"
#include "program.h"

namespace tb {

// layout.cc
double TotalTime(const towr_b200_spec& s);
std::vector<double> BasePolyDurations(const towr_b200_spec& s);
bool OptimizesDurations(const towr_b200_spec& s);
void ValidateSpec(const towr_b200_spec& s);
void BuildLayout(const towr_b200_spec& s, HostProgram& hp);

// structure.cc
// sample times of a TimeDiscretizationConstraint (time_discretization_constraint.cc:41-49)
std::vector<double> SampleTimes(double T, double dt);
// Hermite basis d{pos,vel,acc}(t)/d{start p, start v, end p, end v}: out[dx*4 + k] (polynomial.cc:106-234)
void HermiteBasis(double T, double t, double out[12]);
// segment lookup with the 1e-10 junction rule (spline.cc:48-78)
int SegmentId(double t_global, const std::vector<double>& durations);
// std::pow(x, n) evaluated by libm at run time.  The reference takes most powers with an exponent that is a run-time
// value (polynomial.cc:67-69: std::pow(t, c)), and glibc's pow(x, 2.0) is not always x * x - but g++ folds a literal
// std::pow(x, 2) into x * x.  The class tables must hold the same bits as the oracle's powers (one ulp of t^2 shows as 6e-14
// where a contact force of 0.007 N is the rest of terms near 400; tools/fuzz_parity.py), so every power of the host
// builder goes through here.
double RuntimePow(double x, int n);
double LocalTime(double t_global, int id, const std::vector<double>& durations);
// g_cap > 0 overrides the G slots per segment (the batch lowers it when a class's tables leave less shared memory)
void BuildProgram(const towr_b200_spec& s, HostProgram& hp, int g_cap = 0);
// curvature_dims: the class's TerrainCurvatureDims (a problem on a curvature-free terrain may join a class that declares
// the curvature block; it then writes explicit zeros there)
std::string StructureSignature(const towr_b200_spec& s, int curvature_dims);

// pack.cc
void PackProgram(HostProgram& hp);

// formulation.cc
void ProblemX0(const HostProgram& hp, const towr_b200_spec& s, double* x);
void ProblemVarBounds(const HostProgram& hp, const towr_b200_spec& s, double* lo, double* up);
void ProblemConBounds(const HostProgram& hp, const towr_b200_spec& s, double* lo, double* up);
void FillProblemConst(const towr_b200_spec& s, ProblemConst& pc);
// vals[nnz]: the x-independent stored values of a problem (+-B, mass * B; flat_basis: also terrain-basis component * B on
// a terrain without slope), 0.0 at the positions of hp.dyn_index (flat_basis: hp.dyn_index_flat)
void ProblemConstantValues(const HostProgram& hp, double mass, double torque_k_friction, bool flat_basis, double* vals);
double TerrainHeight(int terrain_id, double x, double y);

}  // namespace tb

// kernels.h — launcher interface between capi.cu and kernels.cu
#pragma once
#include <cuda_runtime.h>

#include "../../include/towr_b200.h"
#include "program.h"

namespace tb {
// A kernel variant (eval_kernel.cuh): `lanes` problems in lock step per stream, `streams` independent streams per CTA,
// ring = value stores through a shared-memory ring and cp.async.bulk instead of the LSU.
struct EvalVariant {
  int lanes = 2, streams = 1;
  bool ring = false;
  int skew = 0;          // clocks by which stream s starts late (s * skew): puts the streams in anti-phase
  bool split = false;    // producer / consumer warp groups with two G buffers (eval_kernel_split.cuh)
  int mode() const { return split ? 2 : ring ? 1 : 0; }
};
// is (lanes, streams, ring) instantiated for this program's kind (fixed / optimised durations)?
bool EvalVariantBuilt(const DeviceProgram& P, const EvalVariant& v);
// dynamic shared memory of the evaluation kernel
size_t EvalSharedBytes(const DeviceProgram& P, int lanes, int streams = 1, int mode = 0);
// evaluates problems 0..n_problems-1; x, g, vals, status are device pointers on the current device
cudaError_t LaunchEval(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                       double* vals, int* status, int n_problems, const EvalVariant& v, int max_ctas, cudaStream_t stream);
// resident CTAs per SM of the evaluation kernel for this program and variant (0: does not fit)
int EvalOccupancy(const DeviceProgram& P, const EvalVariant& v);
}  // namespace tb

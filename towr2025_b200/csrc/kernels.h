// kernels.h — launcher interface between capi.cu and kernels.cu
#pragma once
#include <cuda_runtime.h>

#include "../../include/towr_b200.h"
#include "program.h"

namespace tb {
// dynamic shared memory of the evaluation kernel when `lanes` problems share one CTA
size_t EvalSharedBytes(const DeviceProgram& P, int lanes);
// evaluates problems 0..n_problems-1; x, g, vals, status are device pointers on the current device
cudaError_t LaunchEval(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                       double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream);
// resident CTAs per SM of the evaluation kernel for this program and lane count (0: does not fit)
int EvalOccupancy(const DeviceProgram& P, int lanes);
// warp-specialised kernel (producer warps: stages N, C, A; consumer warps: stage B; double-buffered G)
size_t EvalSharedBytesWS(const DeviceProgram& P, int lanes);
bool EvalFitsWS(const DeviceProgram& P, int lanes, size_t smem_optin);
cudaError_t LaunchEvalWS(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                         double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream);
// warp-specialised kernel whose consumer warps stage sub-blocks in shared memory and store them with cp.async.bulk
size_t EvalSharedBytesWSBulk(const DeviceProgram& P, int lanes);
bool EvalFitsWSBulk(const DeviceProgram& P, int lanes, size_t smem_optin);
cudaError_t LaunchEvalWSBulk(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                             double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream);
// bulk-store kernel (kernels_tma.cu): stage B goes through a shared-memory staging buffer and cp.async.bulk stores
size_t EvalSharedBytesBulk(const DeviceProgram& P, int lanes);
bool EvalFitsBulk(const DeviceProgram& P, int lanes, size_t smem_optin);
cudaError_t LaunchEvalBulk(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                           double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream);
}  // namespace tb

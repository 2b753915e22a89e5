// kernels.cu — dispatch from (lanes, durations, what) to the kernel instantiations (kernels_lanes.cu, kernels_ws.cu,
// kernels_tma.cu) and the shared-memory sizing they share.
#include <cuda_runtime.h>

#include "kernels.h"

namespace tb {

#define TB_DECL(L, R)                                                                                                  \
  cudaError_t LaunchLanes##L##_##R(const DeviceProgram&, const ProblemConst*, int, const double*, double*, double*, int*, \
                                   int, int, cudaStream_t);                                                            \
  int OccupancyLanes##L##_##R(const DeviceProgram&);
TB_DECL(1, 0) TB_DECL(1, 1) TB_DECL(2, 0) TB_DECL(2, 1) TB_DECL(4, 0) TB_DECL(4, 1)
#undef TB_DECL

size_t EvalSharedBytes(const DeviceProgram& P, int lanes) {
  const size_t rt = P.blob.rt ? 8 * static_cast<size_t>(P.blob.n_polys) + P.blob.n_pdur : 0;   // kRtd (device_math.cuh)
  return static_cast<size_t>(P.blob.bytes) + sizeof(double) * lanes *
         (static_cast<size_t>(P.blob.n_node_values) + 6 * static_cast<size_t>(P.blob.n_polys) + rt + P.g_capacity);
}

cudaError_t LaunchEval(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                       double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  const bool rt = P.blob.rt != 0;
#define TB_GO(L) return (rt ? LaunchLanes##L##_1 : LaunchLanes##L##_0)(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream)
  switch (lanes) {
    case 1: TB_GO(1);
    case 2: TB_GO(2);
    case 4: TB_GO(4);
  }
#undef TB_GO
  return cudaErrorInvalidValue;
}

int EvalOccupancy(const DeviceProgram& P, int lanes) {
  const bool rt = P.blob.rt != 0;
  switch (lanes) {
    case 1: return rt ? OccupancyLanes1_1(P) : OccupancyLanes1_0(P);
    case 2: return rt ? OccupancyLanes2_1(P) : OccupancyLanes2_0(P);
    case 4: return rt ? OccupancyLanes4_1(P) : OccupancyLanes4_0(P);
  }
  return 0;
}

}  // namespace tb

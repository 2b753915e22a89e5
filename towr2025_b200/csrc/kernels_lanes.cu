// kernels_lanes.cu — explicit instantiations of the lock-step kernel for one (lanes, durations) pair.  Compiled once per
// pair with -DTB_LANES=<1|2|4> -DTB_RT=<0|1> (Makefile) so that the variants build in parallel.
#if defined(TB_TIMING) && TB_LANES == 2 && TB_RT == 0
#define TB_TIMING_HERE
#endif
#include "eval_kernel.cuh"

namespace tb {

template <int WHAT>
static cudaError_t LaunchT(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  constexpr int LANES = TB_LANES;
  const size_t smem = EvalSharedBytes(P, LANES);
  auto kern = EvalKernel<WHAT, LANES, TB_RT != 0>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int grid = (n_problems + LANES - 1) / LANES;
  if (grid > max_ctas) grid = max_ctas;
  kern<<<grid, kThreads * LANES, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
  return cudaGetLastError();
}

#define TB_CAT3(a, b, c) a##b##_##c
#define TB_NAME(a, b, c) TB_CAT3(a, b, c)

cudaError_t TB_NAME(LaunchLanes, TB_LANES, TB_RT)(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x,
                                                  double* g, double* vals, int* status, int n_problems, int max_ctas,
                                                  cudaStream_t stream) {
  switch (what) {
    case TOWR_B200_EVAL_G: return LaunchT<1>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_JAC: return LaunchT<2>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchT<3>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

int TB_NAME(OccupancyLanes, TB_LANES, TB_RT)(const DeviceProgram& P) {
  int n = 0;
  const size_t smem = EvalSharedBytes(P, TB_LANES);
  auto kern = EvalKernel<3, TB_LANES, TB_RT != 0>;
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, kThreads * TB_LANES, smem) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // namespace tb

TB_PHASE_CLOCK_READER(towr_b200_debug_phase_clocks)

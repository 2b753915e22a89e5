// kernels_variant.cu — explicit instantiations of the evaluation kernel for one variant.  Compiled once per variant with
// -DTB_LANES=<1|2|4> -DTB_RT=<0|1> -DTB_STREAMS=<1|2> -DTB_RING=<0|1|2: split> -DTB_CURV=<0|1> (Makefile) so that the variants build in parallel.
// timing build (make timing [TIMING_CODE=<lanes><rt><streams><ring>]): the one variant that owns the phase counters
#ifndef TB_TIMING_CODE
#define TB_TIMING_CODE 2010
#endif
#ifndef TB_CURV
#define TB_CURV 0
#endif
#if defined(TB_TIMING) && TB_CURV == 0 && (TB_LANES * 1000 + TB_RT * 100 + TB_STREAMS * 10 + TB_RING) == TB_TIMING_CODE
#define TB_TIMING_HERE
#endif
#define TB_PREZERO_G 1   // the kernels zero each segment's G buffer cooperatively (device_math.cuh)
#include "eval_kernel_split.cuh"

namespace tb {

template <int WHAT>
static cudaError_t LaunchT(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, int skew, cudaStream_t stream) {
  constexpr int LANES = TB_LANES, STREAMS = TB_STREAMS;
  const size_t smem = EvalSharedBytes(P, LANES, STREAMS, TB_RING);
  int grid = (n_problems + LANES * STREAMS - 1) / (LANES * STREAMS);
  if (grid > max_ctas) grid = max_ctas;
#if TB_RING == 2
  // g only: nothing for the consumer warps to do, the lock-step kernel of the same lane count runs instead (kernels.cu)
  auto kern = EvalKernelSplit<WHAT, LANES, TB_RT != 0, TB_CURV != 0>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  kern<<<grid, (kThreads + kSplitConsumersPerLane) * LANES, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
#else
  auto kern = EvalKernel<WHAT, LANES, TB_RT != 0, STREAMS, TB_RING != 0, TB_CURV != 0>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  kern<<<grid, kThreads * LANES * STREAMS, smem, stream>>>(P, pcs, x, g, vals, status, n_problems, skew);
#endif
  return cudaGetLastError();
}

#define TB_CAT6(a, b, c, d, e, f) a##b##_##c##_##d##_##e##_##f
#define TB_NAME6(a, b, c, d, e, f) TB_CAT6(a, b, c, d, e, f)
#define TB_NAME(a, b, c, d, e) TB_NAME6(a, b, c, d, e, TB_CURV)

cudaError_t TB_NAME(LaunchV, TB_LANES, TB_RT, TB_STREAMS, TB_RING)(const DeviceProgram& P, const ProblemConst* pcs, int what,
                                                                   const double* x, double* g, double* vals, int* status,
                                                                   int n_problems, int max_ctas, int skew, cudaStream_t stream) {
  switch (what) {
    case TOWR_B200_EVAL_G: return LaunchT<1>(P, pcs, x, g, vals, status, n_problems, max_ctas, skew, stream);
    case TOWR_B200_EVAL_JAC: return LaunchT<2>(P, pcs, x, g, vals, status, n_problems, max_ctas, skew, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchT<3>(P, pcs, x, g, vals, status, n_problems, max_ctas, skew, stream);
  }
  return cudaErrorInvalidValue;
}

int TB_NAME(OccupancyV, TB_LANES, TB_RT, TB_STREAMS, TB_RING)(const DeviceProgram& P) {
  int n = 0;
  const size_t smem = EvalSharedBytes(P, TB_LANES, TB_STREAMS, TB_RING);
#if TB_RING == 2
  auto kern = EvalKernelSplit<3, TB_LANES, TB_RT != 0, TB_CURV != 0>;
  const int threads = (kThreads + kSplitConsumersPerLane) * TB_LANES;
#else
  auto kern = EvalKernel<3, TB_LANES, TB_RT != 0, TB_STREAMS, TB_RING != 0, TB_CURV != 0>;
  const int threads = kThreads * TB_LANES * TB_STREAMS;
#endif
  if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, threads, smem) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // namespace tb

#ifdef TB_TIMING_HERE
TB_PHASE_CLOCK_READER(towr_b200_debug_phase_clocks)
#endif

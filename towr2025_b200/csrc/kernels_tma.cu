// kernels_tma.cu — bulk-store variant of the lock-step kernel (EvalKernel<..., BULK = true>, eval_kernel.cuh): the
// Jacobian values of a segment are staged in shared memory in storage order and written by the TMA engine
// (cp.async.bulk shared -> global), asynchronously to stage A of the next segment.
#if defined(TB_TIMING)
#define TB_TIMING_HERE
#endif
#include "eval_kernel.cuh"

namespace tb {

size_t EvalSharedBytesBulk(const DeviceProgram& P, int lanes) {
  DeviceProgram t = P;
  t.g_capacity = P.joint_capacity;     // G and the staging buffer share the joint region
  return EvalSharedBytes(t, lanes) + 16;
}

bool EvalFitsBulk(const DeviceProgram& P, int lanes, size_t smem_optin) {
  if (lanes != 1 && lanes != 2) return false;
  return EvalSharedBytesBulk(P, lanes) + 1024 <= smem_optin;   // + static shared memory of the kernel
}

template <int WHAT, int LANES, bool RT>
static cudaError_t LaunchB(const DeviceProgram& P, const ProblemConst* pcs, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  const size_t smem = EvalSharedBytesBulk(P, LANES);
  auto kern = EvalKernel<WHAT, LANES, RT, true>;
  cudaError_t err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int grid = (n_problems + LANES - 1) / LANES;
  if (grid > max_ctas) grid = max_ctas;
  kern<<<grid, kThreads * LANES, smem, stream>>>(P, pcs, x, g, vals, status, n_problems);
  return cudaGetLastError();
}

template <int LANES, bool RT>
static cudaError_t LaunchW(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g, double* vals,
                           int* status, int n_problems, int max_ctas, cudaStream_t stream) {
  switch (what) {
    // values only: nothing to stage, the lock-step kernel is the same program
    case TOWR_B200_EVAL_G: return LaunchEval(P, pcs, what, x, g, vals, status, n_problems, LANES, max_ctas, stream);
    case TOWR_B200_EVAL_JAC: return LaunchB<2, LANES, RT>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
    case TOWR_B200_EVAL_BOTH: return LaunchB<3, LANES, RT>(P, pcs, x, g, vals, status, n_problems, max_ctas, stream);
  }
  return cudaErrorInvalidValue;
}

cudaError_t LaunchEvalBulk(const DeviceProgram& P, const ProblemConst* pcs, int what, const double* x, double* g,
                           double* vals, int* status, int n_problems, int lanes, int max_ctas, cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  const bool rt = P.blob.rt != 0;
#ifdef TB_BULK_MINIMAL   // development builds: the headline instantiation only
  if (lanes == 2 && !rt) return LaunchW<2, false>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream);
  return cudaErrorInvalidValue;
#else
  if (lanes == 2) return rt ? LaunchW<2, true>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream)
                            : LaunchW<2, false>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream);
  if (lanes == 1) return rt ? LaunchW<1, true>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream)
                            : LaunchW<1, false>(P, pcs, what, x, g, vals, status, n_problems, max_ctas, stream);
  return cudaErrorInvalidValue;
#endif
}

}  // namespace tb

TB_PHASE_CLOCK_READER(towr_b200_debug_phase_clocks_bulk)

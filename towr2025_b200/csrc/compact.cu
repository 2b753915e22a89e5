// compact.cu — gathers the x-dependent Jacobian values of a batch into a dense buffer (TOWR_B200_EVAL_JAC_DYNAMIC).
//
// The evaluation kernel writes every stored value (n x nnz, what eval_jac_g hands to the solver).  On the host interface
// 40-90 % of those bytes are constants of the problem (+-B, mass * B: HostProgram::dyn_index lists the others), and the
// PCIe link, not the kernel, bounds the end-to-end step.  dyn[p][k] = vals[p][idx[k]]: the D2H copy then carries the
// dynamic entries only; towr_b200_scatter_dynamic puts them back between the constants on the host.
#include <cuda_runtime.h>

#include "kernels.h"

namespace tb {

constexpr int kCompactThreads = 256;

__global__ void __launch_bounds__(kCompactThreads)
CompactKernel(const double* __restrict__ vals, const int* __restrict__ idx, double* __restrict__ dyn, int nnz, int nd, int n) {
  const int k = blockIdx.x * kCompactThreads + threadIdx.x;
  if (k >= nd) return;
  const int src = __ldg(idx + k);
  for (int p = blockIdx.y; p < n; p += gridDim.y)
    __stcs(dyn + static_cast<size_t>(p) * nd + k, __ldcs(vals + static_cast<size_t>(p) * nnz + src));
}

cudaError_t LaunchCompact(const double* vals, const int* idx, double* dyn, int nnz, int nd, int n, cudaStream_t stream) {
  if (n <= 0 || nd <= 0) return cudaSuccess;
  dim3 grid((nd + kCompactThreads - 1) / kCompactThreads, n < 16384 ? n : 16384);
  CompactKernel<<<grid, kCompactThreads, 0, stream>>>(vals, idx, dyn, nnz, nd, n);
  return cudaGetLastError();
}

}  // namespace tb

// cost_kernel.cu — eval_f / eval_grad_f for the NodeCost terms of NlpFormulation::GetCosts (nlp_formulation.cc:605-681,
// costs/node_cost.cc:36-76): f = sum over terms of weight * value^2 over the nodes of a set, gradient entry
// weight * 2 * value at every optimised variable that writes such a node value (a stance variable that writes two
// nodes is counted twice, as in the reference).  One CTA per problem; thread c owns column c of the dense gradient.
#include <cuda_runtime.h>

#include "cost.h"

namespace tb {

constexpr int kCostThreads = 128;

__global__ void __launch_bounds__(kCostThreads)
CostKernel(const __grid_constant__ CostDevice Cd, int what, const double* __restrict__ x, double* __restrict__ f_out,
           double* __restrict__ grad_out, int n_problems) {
  __shared__ double part[kCostThreads / 32];
  const int tid = threadIdx.x;
  for (int prob = blockIdx.x; prob < n_problems; prob += gridDim.x) {
    const double* xp = x + static_cast<size_t>(prob) * Cd.n_vars;
    double fsum = 0.0;
    for (int c = tid; c < Cd.n_vars; c += kCostThreads) {
      const int k0 = Cd.col_ptr[c], k1 = Cd.col_ptr[c + 1];
      double g = 0.0;
      if (k1 > k0) {
        const double val = __ldg(xp + c);
        for (int k = k0; k < k1; ++k) {
          const double w = Cd.weights[k];
          g = g + (w * 2.0) * val;             // jac.coeffRef(0, i) += weight_*2.0*val   (node_cost.cc:73)
          fsum = fsum + w * (val * val);       // cost += weight_*std::pow(val,2)         (node_cost.cc:59)
        }
      }
      if (what & 2) grad_out[static_cast<size_t>(prob) * Cd.n_vars + c] = g;
    }
    if (what & 1) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) fsum += __shfl_xor_sync(0xFFFFFFFFu, fsum, o);
      if ((tid & 31) == 0) part[tid >> 5] = fsum;
      __syncthreads();
      if (tid == 0) {
        double t = 0.0;
        for (int i = 0; i < kCostThreads / 32; ++i) t += part[i];
        f_out[prob] = t;
      }
      __syncthreads();
    }
  }
}

cudaError_t LaunchCost(const CostDevice& c, int what, const double* x, double* f, double* grad, int n_problems, int max_ctas,
                       cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  const int grid = n_problems < max_ctas ? n_problems : max_ctas;
  CostKernel<<<grid, kCostThreads, 0, stream>>>(c, what, x, f, grad, n_problems);
  return cudaGetLastError();
}

}  // namespace tb

// cost_kernel.cu — eval_f / eval_grad_f for the NodeCost terms of NlpFormulation::GetCosts (nlp_formulation.cc:605-681,
// costs/node_cost.cc:36-76): f = sum over terms of weight * value^2 over the nodes of a set, gradient entry
// weight * 2 * value at every optimised variable that writes such a node value (a stance variable that writes two
// nodes is counted twice, as in the reference).  One CTA per problem; thread c owns column c of the dense gradient.
// A second kernel (SampledCostKernel, below) adds the time-discretised terms: EnergyCost, AngularMomentumCost,
// EEBasePosCost (costs/energy_cost.cc, angular_momentum_cost.cc, ee_base_pos_cost.cc).
#include <cuda_runtime.h>

#include <mutex>

#include "cost.h"
#include "sampled_cost.cuh"

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

// ---- time-discretised cost terms (EnergyCost, AngularMomentumCost): arithmetic in sampled_cost.cuh ------------------
// One CTA per problem (grid-stride), class tables (cost.h blob) staged in shared memory once per CTA.  Steps:
//   1. item = (spline, sample): segment and local time on the problem's own polynomial durations (spline.cc:48-78), the
//      spline's value there, and what the gradient needs of it (the row vector m of energy_cost.cc:112 / the chain-rule
//      factors of angular_momentum_cost.cc:144-203);
//   2. f: one thread per sample forms its summand in the reference's order, one thread per term adds them in time order;
//   3. gradient: one thread per column walks the samples of the column's two polynomials in time order, like the
//      reference's accumulation into jac.coeffRef(0, col) over its sample loop.
// All terms of the class go through each step together (one barrier per step).
// Adds to f / grad written by CostKernel earlier on the same stream.
constexpr int kSampledThreads = 256;

// make cost_timing: per-step clocks of CTA 0's first problem (tools/cost_timing.py); not in the product build
#ifdef TB_COST_TIMING
__device__ long long g_cost_clocks[16];
#define TB_COST_CLK(i) do { if (blockIdx.x == 0 && threadIdx.x == 0 && prob == 0) g_cost_clocks[i] = clock64(); } while (0)
#else
#define TB_COST_CLK(i) do { } while (0)
#endif

__global__ void __launch_bounds__(kSampledThreads, 2)
SampledCostKernel(const unsigned char* __restrict__ blob_g, const ProblemConst* __restrict__ pcs,
                  const double* __restrict__ pref, int what, const double* __restrict__ x, double* __restrict__ f_out, double* __restrict__ grad_out, int n_problems) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int tid = threadIdx.x;
  const int blob_bytes = reinterpret_cast<const SampledHeader*>(blob_g)->bytes;
  for (int i = tid; i < blob_bytes / 16; i += kSampledThreads)
    reinterpret_cast<int4*>(smem)[i] = __ldg(reinterpret_cast<const int4*>(blob_g) + i);
  __syncthreads();
  const SampledView V = MakeSampledView(smem);
  const SampledHeader& H = *V.h;
  const SampledWork W = SampledCarve(H, smem + ((H.bytes + 15) & ~15));
  for (int prob = blockIdx.x; prob < n_problems; prob += gridDim.x) {
    const double* xp = x + static_cast<size_t>(prob) * H.n_vars;
    TB_COST_CLK(0);
#pragma unroll 4
    for (int i = tid; i < 6 * H.n_nodes; i += kSampledThreads) SampledNode(V, W, xp, i);
    if (H.rt && tid < H.n_ee) SampledPhaseDurations(V, W, xp, tid);
    __syncthreads();
    TB_COST_CLK(1);
    for (int poly = tid; poly < H.n_polys + H.n_phase_slots; poly += kSampledThreads) SampledPolyRecord(V, W, poly);
    __syncthreads();
    TB_COST_CLK(2);
    for (int i = tid; i < 3 * H.n_polys + H.n_splines; i += kSampledThreads) SampledCoeff(V, W, i);
    __syncthreads();
    TB_COST_CLK(3);
    double I[9];
#pragma unroll
    for (int i = 0; i < 9; ++i) I[i] = pcs ? __ldg(&pcs[prob].I[i]) : H.I[i];
    const double* p_ref_B = pref ? pref + 12 * static_cast<size_t>(prob) : H.p_ref_B;
    for (int it = tid; it < H.n_items; it += kSampledThreads) SampledItem(V, W, I, p_ref_B, it);
    __syncthreads();
    TB_COST_CLK(4);
    if (what & 1) {
      for (int k = tid; k < H.n_times; k += kSampledThreads) SampledSummand(V, W, k);
      __syncthreads();
    }
    TB_COST_CLK(5);
    // one thread per term adds its samples in time order: the last lanes of the last warps, which have the fewest jobs
    if ((what & 1) && (tid & 31) == 31 && (kSampledThreads / 32 - 1 - (tid >> 5)) < H.n_terms)
      SampledTermSum(V, W, kSampledThreads / 32 - 1 - (tid >> 5));
    TB_COST_CLK(6);
    if (what & 2) {
      double* gp = grad_out + static_cast<size_t>(prob) * H.n_vars;
      for (int j = tid; j < H.n_jobs; j += kSampledThreads) SampledJobGradient(V, W, gp, j);
    }
    TB_COST_CLK(7);
    __syncthreads();
    if ((what & 2) && H.n_comb > 0) {                  // columns that several terms write: parked sums, added in term order
      double* gp = grad_out + static_cast<size_t>(prob) * H.n_vars;
      for (int i = tid; i < H.n_comb; i += kSampledThreads) SampledCombine(V, W, gp, i);
      __syncthreads();                                 // the scratch is reused by the next problem's jobs
    }
    TB_COST_CLK(8);
    if ((what & 1) && tid == 0) {
      double f = f_out[prob];
      for (int ti = 0; ti < H.n_terms; ++ti) f = f + W.fterm[ti];   // the terms in the order of params_.costs_
      f_out[prob] = f;
    }
  }
}

cudaError_t LaunchSampledCost(const unsigned char* blob_dev, size_t shared_bytes, const ProblemConst* pcs, const double* pref,
                              int what, const double* x, double* f, double* grad, int n_problems, int sms, cudaStream_t stream) {
  if (n_problems <= 0 || !blob_dev) return cudaSuccess;
  cudaError_t err = cudaFuncSetAttribute(SampledCostKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         static_cast<int>(shared_bytes));
  if (err != cudaSuccess) return err;
  int occ = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, SampledCostKernel, kSampledThreads, shared_bytes) != cudaSuccess ||
      occ < 1) {
    cudaGetLastError();
    occ = 1;
  }
  const int max_ctas = sms * occ;
  const int grid = n_problems < max_ctas ? n_problems : max_ctas;
  SampledCostKernel<<<grid, kSampledThreads, shared_bytes, stream>>>(blob_dev, pcs, pref, what, x, f, grad, n_problems);
  return cudaGetLastError();
}

// ---- allocation-free variant (towr_b200_eval_cost_static, include/towr_b200.h) --------------------------------------
// The work tables of the cost kernel (column pointers and weights), which towr_b200_batch_eval_cost[_device] keeps in
// cudaMalloc'ed memory, live here in a file-scope device array: every GPU that loads the library holds these
// kCostStaticBytes until the process exits.  All calls share the array and each call rewrites the part it uses (nothing
// in it has to start at zero), so calls that overlap on the GPU race.
constexpr int kCostStaticMaxVars = TOWR_B200_STATIC_COST_MAX_VARS;
constexpr int kCostStaticMaxWeights = TOWR_B200_STATIC_COST_MAX_WEIGHTS;
constexpr size_t kCostStaticWeightsOff = ((sizeof(int) * (kCostStaticMaxVars + 1)) + 255) & ~static_cast<size_t>(255);
constexpr size_t kCostStaticSampledOff = (kCostStaticWeightsOff + sizeof(double) * kCostStaticMaxWeights + 255) & ~static_cast<size_t>(255);
constexpr size_t kCostStaticBytes = kCostStaticSampledOff + TOWR_B200_STATIC_COST_MAX_SAMPLED_BYTES;
static __device__ __align__(256) unsigned char g_cost_static[kCostStaticBytes];

constexpr int kMaxDevices = 64;
static unsigned char* g_cost_static_addr[kMaxDevices];     // device address of g_cost_static per device, looked up once
static int g_cost_static_sms[kMaxDevices];
static std::mutex g_cost_static_mu;

cudaError_t LaunchCostStatic(const CostTables& h, const std::vector<SampledCost>& sc, int device, int what, const double* x, double* f,
                             double* grad, int n_problems, cudaStream_t stream, const char** why) {
  *why = "";
  if (device < 0 || device >= kMaxDevices) { *why = "device index out of range"; return cudaErrorInvalidDevice; }
  if (h.n_vars > kCostStaticMaxVars || static_cast<int>(h.weights.size()) > kCostStaticMaxWeights) {
    *why = "the class exceeds TOWR_B200_STATIC_COST_MAX_VARS / _MAX_WEIGHTS";
    return cudaErrorInvalidValue;
  }
  size_t sampled_bytes = 0;
  for (const SampledCost& part : sc) sampled_bytes += (part.blob.size() + 255) & ~static_cast<size_t>(255);
  if (sampled_bytes > TOWR_B200_STATIC_COST_MAX_SAMPLED_BYTES) {
    *why = "the class's EnergyCost / AngularMomentumCost tables exceed TOWR_B200_STATIC_COST_MAX_SAMPLED_BYTES";
    return cudaErrorInvalidValue;
  }
  cudaError_t err = cudaSetDevice(device);
  if (err != cudaSuccess) return err;
  unsigned char* base = nullptr;
  int sms = 0;
  {
    std::lock_guard<std::mutex> lock(g_cost_static_mu);
    if (!g_cost_static_addr[device]) {
      // in host code `g_cost_static` is only a stand-in for the symbol: the kernel needs the device address
      void* p = nullptr;
      err = cudaGetSymbolAddress(&p, g_cost_static);
      if (err != cudaSuccess) { *why = "cudaGetSymbolAddress(g_cost_static)"; return err; }
      err = cudaDeviceGetAttribute(&g_cost_static_sms[device], cudaDevAttrMultiProcessorCount, device);
      if (err != cudaSuccess) return err;
      g_cost_static_addr[device] = static_cast<unsigned char*>(p);
    }
    base = g_cost_static_addr[device];
    sms = g_cost_static_sms[device];
  }
  if (n_problems <= 0) return cudaSuccess;
  int* col_ptr = reinterpret_cast<int*>(base);
  double* weights = reinterpret_cast<double*>(base + kCostStaticWeightsOff);
  err = cudaMemcpyAsync(col_ptr, h.col_ptr.data(), sizeof(int) * h.col_ptr.size(), cudaMemcpyHostToDevice, stream);
  if (err != cudaSuccess) return err;
  if (!h.weights.empty()) {
    err = cudaMemcpyAsync(weights, h.weights.data(), sizeof(double) * h.weights.size(), cudaMemcpyHostToDevice, stream);
    if (err != cudaSuccess) return err;
  }
  CostDevice c;
  c.n_vars = h.n_vars;
  c.col_ptr = col_ptr;
  c.weights = weights;
  err = LaunchCost(c, what, x, f, grad, n_problems, 16 * sms, stream);
  // the time-discretised terms: tables after the weights (one blob per launch), inertia and swing reference positions from
  // the class spec (no per-problem constants here)
  unsigned char* blob = base + kCostStaticSampledOff;
  for (const SampledCost& part : sc) {
    if (err != cudaSuccess) return err;
    err = cudaMemcpyAsync(blob, part.blob.data(), part.blob.size(), cudaMemcpyHostToDevice, stream);
    if (err != cudaSuccess) return err;
    err = LaunchSampledCost(blob, part.shared_bytes, nullptr, nullptr, what, x, f, grad, n_problems, sms, stream);
    blob += (part.blob.size() + 255) & ~static_cast<size_t>(255);
  }
  return err;
}

}  // namespace tb

#ifdef TB_COST_TIMING
extern "C" int tb_cost_clocks(long long* out) {
  return cudaMemcpyFromSymbol(out, tb::g_cost_clocks, sizeof(long long) * 16) == cudaSuccess ? 0 : -1;
}
#endif

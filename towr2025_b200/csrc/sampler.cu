// sampler.cu — batched trajectory sampling kernel (sm_100a): SplineHolder::GetPoint of every spline on a time grid.
//
// One CTA per problem (grid-stride).  Shared memory: the node values of all splines (gathered from x), the Hermite
// coefficients c, d of every polynomial (polynomial.cc:97-104) and, with optimised durations, the polynomial durations
// (phase_durations.cc:78-100, nodes_variables_phase_based.cc:75-86).  Work item = (sample, spline): thread i evaluates
// pos / vel / acc (polynomial.cc:47-72) and writes its columns of the sample's CSV row (save_data.cpp:57-112) into a
// shared-memory tile of rows, which the CTA then stores coalesced.
// HBM traffic per problem: n_vars doubles in, n_samples * width doubles out; the sample tables are L2-resident.
#include <cuda_runtime.h>

#include "device_math.cuh"
#include "trajectory.h"

namespace tb {

constexpr int kSampleThreads = 256;
constexpr int kRowChunk = 32;          // samples staged per pass (one warp = one spline's samples of the chunk)
constexpr int kRec = 7;                // doubles per node / coefficient record in shared memory: 6 used; with a stride of 6
                                       // doubles the records of a warp's splines fall onto 8 bank groups (9 wavefronts per LDS.64)

__global__ void __launch_bounds__(kSampleThreads, 3)
SampleKernel(const __grid_constant__ TrajDevice Tj, const double* __restrict__ x, double* __restrict__ out, int n_problems) {
  extern __shared__ double sm[];
  double* nodes = sm;                         // 6 per node
  double* coef = nodes + kRec * Tj.n_nodes;   // 6 per polynomial: c[3], d[3]
  double* pT = coef + kRec * Tj.n_polys;         // 3 per polynomial: T, T^2, T^3
  double* pdur = pT + 3 * Tj.n_polys;         // optimised durations: phase durations of every foot
  double* rowbuf = pdur + Tj.n_phases;        // kRowChunk rows
  // class tables, staged once per CTA (every lookup below would otherwise be an L2 round trip)
  TrajSplineRec* splines = reinterpret_cast<TrajSplineRec*>(rowbuf + kRowChunk * Tj.width);
  TrajFootRec* feet = reinterpret_cast<TrajFootRec*>(splines + Tj.n_splines);
  uint32_t* pmeta = reinterpret_cast<uint32_t*>(feet + Tj.n_ee);
  uint16_t* pspl = reinterpret_cast<uint16_t*>(pmeta + Tj.n_polys);
  uint16_t* nsrc = pspl + Tj.n_polys;
  const int tid = threadIdx.x;
  for (int i = tid; i < Tj.n_splines; i += kSampleThreads) splines[i] = Tj.splines[i];
  for (int i = tid; i < Tj.n_ee; i += kSampleThreads) feet[i] = Tj.feet[i];
  for (int i = tid; i < Tj.n_polys; i += kSampleThreads) { pmeta[i] = Tj.poly_meta[i]; pspl[i] = Tj.poly_spline[i]; }
  for (int i = tid; i < 6 * Tj.n_nodes; i += kSampleThreads) nsrc[i] = Tj.node_src[i];
  __syncthreads();
  for (int prob = blockIdx.x; prob < n_problems; prob += gridDim.x) {
    const double* xp = x + static_cast<size_t>(prob) * Tj.n_vars;
    // node values (NodesVariables::SetVariables, nodes_variables.cc:68-76): values that are not optimised stay 0
    for (int i = tid; i < 6 * Tj.n_nodes; i += kSampleThreads) {
      const unsigned src = nsrc[i];
      nodes[(i / 6) * kRec + i % 6] = src != 0xFFFFu ? __ldg(xp + src) : 0.0;
    }
    if (Tj.rt && tid < Tj.n_ee) {             // PhaseDurations::SetVariables (phase_durations.cc:78-100)
      const TrajFootRec ft = feet[tid];
      double sum = 0.0;
      for (int i = 0; i < ft.n_phases - 1; ++i) {
        const double d = __ldg(xp + ft.col0 + i);
        pdur[ft.phase0 + i] = d;
        sum += d;
      }
      pdur[ft.phase0 + ft.n_phases - 1] = ft.total - sum;
    }
    __syncthreads();
    for (int poly = tid; poly < Tj.n_polys; poly += kSampleThreads) {   // polynomial durations
      const uint32_t pm = pmeta[poly];
      if (pm == kPolyFixed) {
        pT[3 * poly] = __ldg(Tj.poly_T + 3 * poly); pT[3 * poly + 1] = __ldg(Tj.poly_T + 3 * poly + 1);
        pT[3 * poly + 2] = __ldg(Tj.poly_T + 3 * poly + 2);
      } else {                                 // ConvertPhaseToPolyDurations (nodes_variables_phase_based.cc:75-86)
        const double Td = pdur[feet[splines[pspl[poly]].foot].phase0 + (pm & 0xFF)] / static_cast<double>((pm >> 8) & 0xFF);
        pT[3 * poly] = Td; pT[3 * poly + 1] = Td * Td; pT[3 * poly + 2] = Cube(Td);
      }
    }
    __syncthreads();
    // CubicHermitePolynomial::UpdateCoeff (polynomial.cc:97-104)
    for (int i = tid; i < 3 * Tj.n_polys; i += kSampleThreads) {
      const int poly = i / 3, d = i - 3 * poly;
      const TrajSplineRec sp = splines[pspl[poly]];
      const double* n0 = nodes + kRec * (sp.node0 + (poly - sp.poly0));
      const double p0 = n0[d], v0 = n0[3 + d], p1 = n0[kRec + d], v1 = n0[kRec + 3 + d];
      const double T = pT[3 * poly], T2 = pT[3 * poly + 1], T3 = pT[3 * poly + 2];
      coef[kRec * poly + d] = -(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)) / T2;
      coef[kRec * poly + 3 + d] = (2.0 * (p0 - p1) + T * (v0 + v1)) / T3;
    }
    __syncthreads();
    // rows are staged in shared memory kRowChunk samples at a time and written with coalesced stores (a problem's rows
    // are one contiguous block of n_samples * width doubles)
    double* rows = out + static_cast<size_t>(prob) * Tj.n_samples * Tj.width;
    for (int k0 = 0; k0 < Tj.n_samples; k0 += kRowChunk) {
      const int kc = min(kRowChunk, Tj.n_samples - k0);
      const int n_items = kc * Tj.n_splines;
      for (int i = tid; i < n_items; i += kSampleThreads) {
        // consecutive threads take consecutive samples of one spline: they mostly share the polynomial (broadcast reads
        // of its node / coefficient record), read the sample tables coalesced and write rows with an odd stride
        const int s = i / kc, kl = i - s * kc, k = k0 + kl;
        const TrajSplineRec sp = splines[s];
        const double tg = __ldg(Tj.times + k);
        int poly;
        double t, t2, t3;
        if (!sp.rt) {
          const size_t si = static_cast<size_t>(s) * Tj.n_samples + k;
          poly = static_cast<int>(__ldg(Tj.samp_poly + si));
          t = __ldg(Tj.samp_t + 3 * si); t2 = __ldg(Tj.samp_t + 3 * si + 1); t3 = __ldg(Tj.samp_t + 3 * si + 2);
        } else {
          // Spline::GetSegmentID / GetLocalTime (spline.cc:48-78) on the problem's own durations
          int id = sp.n_polys - 1;
          double acc = 0.0;
          for (int j = 0; j < sp.n_polys; ++j) {
            acc += pT[3 * (sp.poly0 + j)];
            if (acc >= tg - 1e-10) { id = j; break; }
          }
          t = tg;
          for (int j = 0; j < id; ++j) t -= pT[3 * (sp.poly0 + j)];
          poly = sp.poly0 + id;
          t2 = t * t; t3 = Cube(t);
        }
        const double* n0 = nodes + kRec * (sp.node0 + (poly - sp.poly0));
        double* row = rowbuf + kl * Tj.width;
        double* o = row + sp.out_col;
#pragma unroll
        for (int d = 0; d < 3; ++d) {          // Polynomial::GetPoint (polynomial.cc:47-61), coefficients A..D in order
          const double a = n0[d], b = n0[3 + d], c = coef[kRec * poly + d], dd = coef[kRec * poly + 3 + d];
          o[d] = ((a + t * b) + t2 * c) + t3 * dd;
          if (sp.n_out == 9) {
            o[3 + d] = (b + (2 * t) * c) + (3 * t2) * dd;
            o[6 + d] = 2.0 * c + (6 * t) * dd;
          }
        }
        if (s == 0) row[0] = tg;
        if (sp.foot >= 0 && sp.n_out == 9 && sp.out_col == 19 + 25 * sp.foot) {   // the foot's ee-motion item also writes is_contact
          double contact;
          if (!Tj.rt) {
            contact = __ldg(Tj.contact + static_cast<size_t>(sp.foot) * Tj.n_samples + k) ? 1.0 : 0.0;
          } else {                             // PhaseDurations::IsContactPhase (phase_durations.cc:113-118)
            const TrajFootRec ft = feet[sp.foot];
            int ph = ft.n_phases - 1;
            double acc = 0.0;
            for (int j = 0; j < ft.n_phases; ++j) {
              acc += pdur[ft.phase0 + j];
              if (acc >= tg - 1e-10) { ph = j; break; }
            }
            const bool c0 = ft.contact_at_start != 0;
            contact = ((ph % 2 == 0) ? c0 : !c0) ? 1.0 : 0.0;
          }
          row[19 + 25 * sp.foot + 24] = contact;
        }
      }
      __syncthreads();
      double* dst = rows + static_cast<size_t>(k0) * Tj.width;
#pragma unroll 4
      for (int i = tid; i < kc * Tj.width; i += kSampleThreads) __stcs(dst + i, rowbuf[i]);   // streaming: rows are never read back
      __syncthreads();
    }
  }
}

size_t TrajSharedBytes(const TrajDevice& t) {
  return sizeof(double) * (kRec * static_cast<size_t>(t.n_nodes) + (kRec + 3) * static_cast<size_t>(t.n_polys) + t.n_phases +
                           static_cast<size_t>(kRowChunk) * t.width) +
         sizeof(TrajSplineRec) * t.n_splines + sizeof(TrajFootRec) * t.n_ee + 6 * static_cast<size_t>(t.n_polys) +
         12 * static_cast<size_t>(t.n_nodes) + 16;
}

cudaError_t LaunchSampleTrajectory(const TrajDevice& t, const double* x, double* out, int n_problems, int sms,
                                   cudaStream_t stream) {
  if (n_problems <= 0) return cudaSuccess;
  const size_t smem = TrajSharedBytes(t);
  cudaError_t err = cudaFuncSetAttribute(SampleKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  if (err != cudaSuccess) return err;
  int occ = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, SampleKernel, kSampleThreads, smem) != cudaSuccess || occ < 1) {
    cudaGetLastError();
    occ = 1;
  }
  const int max_ctas = sms * occ;
  const int grid = n_problems < max_ctas ? n_problems : max_ctas;
  SampleKernel<<<grid, kSampleThreads, smem, stream>>>(t, x, out, n_problems);
  return cudaGetLastError();
}

}  // namespace tb

// sampled_cost.cuh — per-problem arithmetic of the time-discretised cost terms (cost.h): EnergyCost
// (costs/energy_cost.cc:38-163) and AngularMomentumCost (costs/angular_momentum_cost.cc:39-208).  The kernel
// (cost_kernel.cu) strides these functions over a CTA's threads with a barrier between the steps; tests/emu replays them
// serially on the host.
#pragma once
#include "cost.h"
#include "device_math.cuh"

namespace tb {

struct SampledView {
  const SampledHeader* h;
  const SampledTerm* terms;
  const SampledSpline* splines;
  const SampledFoot* feet;
  const double* poly_T;
  const double* times;
  const SampledJob* jobs;
  const uint32_t* poly_meta;
  const uint16_t* node_src;
  const uint16_t* poly_spline;
};

// d (pos | vel)(t) / d (node value) of a cubic Hermite polynomial (polynomial.cc:106-234); side 0 start, 1 end
__device__ __forceinline__ double HermiteWeight(int dxdt, int side, int deriv, double t, double T) {
  const double t2 = t * t, t3 = Cube(t), T2 = T * T, T3 = Cube(T);
  if (dxdt == 0) {
    if (side == 0) return deriv == 0 ? (2 * t3) / T3 - (3 * t2) / T2 + 1 : t - (2 * t2) / T + t3 / T2;
    return deriv == 0 ? (3 * t2) / T2 - (2 * t3) / T3 : t3 / T2 - t2 / T;
  }
  if (side == 0) return deriv == 0 ? (6 * t2) / T3 - (6 * t) / T2 : (3 * t2) / T2 - (4 * t) / T + 1;
  return deriv == 0 ? (6 * t) / T2 - (6 * t2) / T3 : (3 * t2) / T2 - (2 * t) / T;
}

__device__ __forceinline__ void Mat3Mul(const double A[9], const double B[9], double C[9]) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = (A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j]) + A[3 * i + 2] * B[6 + j];
}
__device__ __forceinline__ double Dot3(const double a[3], const double b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }


// shared-memory work arrays of one CTA (one problem at a time), carved by SampledCarve
struct SampledWork {
  double *nodes, *coef, *pT, *pdur, *itl, *im, *ix, *ie, *inst;
  int* iseg;             // per item: polynomial within the spline | phase << 16
};

__device__ __forceinline__ SampledView MakeSampledView(const unsigned char* blob) {
  SampledView V;
  V.h = reinterpret_cast<const SampledHeader*>(blob);
  const SampledHeader& H = *V.h;
  V.terms = reinterpret_cast<const SampledTerm*>(blob + H.off_terms);
  V.splines = reinterpret_cast<const SampledSpline*>(blob + H.off_splines);
  V.feet = reinterpret_cast<const SampledFoot*>(blob + H.off_feet);
  V.poly_T = reinterpret_cast<const double*>(blob + H.off_poly_T);
  V.times = reinterpret_cast<const double*>(blob + H.off_times);
  V.jobs = reinterpret_cast<const SampledJob*>(blob + H.off_jobs);
  V.poly_meta = reinterpret_cast<const uint32_t*>(blob + H.off_poly_meta);
  V.node_src = reinterpret_cast<const uint16_t*>(blob + H.off_node_src);
  V.poly_spline = reinterpret_cast<const uint16_t*>(blob + H.off_poly_spline);
  return V;
}

// `base`: the bytes after the staged blob (SampledSharedBytes, cost.h)
__device__ __forceinline__ SampledWork SampledCarve(const SampledHeader& H, unsigned char* base) {
  SampledWork W;
  W.nodes = reinterpret_cast<double*>(base);
  W.coef = W.nodes + kSampledRec * H.n_nodes;            // c[3], d[3] of every polynomial (polynomial.cc:97-104)
  W.pT = W.coef + kSampledRec * H.n_polys;
  W.pdur = W.pT + H.n_polys;
  W.itl = W.pdur + H.n_phases;                           // per item: local time
  W.im = W.itl + H.max_items;                            //   3: energy m(r) / ang-mom d cost / d euler angle
  W.ix = W.im + 3 * H.max_items;                         //   3: energy (m . dx/dT, m . vel, -) / ang-mom d cost / d euler rate
  W.ie = W.ix + 3 * H.max_items;                         //   squared norm
  W.inst = W.ie + H.max_items;                           // per sample: the summand of f
  W.iseg = reinterpret_cast<int*>(W.inst + H.max_samples + 4);
  return W;
}

// ---- set-up of one problem; a barrier separates the four steps
// node value i of 6 * n_nodes (nodes_variables.cc:68-76)
__device__ __forceinline__ void SampledNode(const SampledView& V, const SampledWork& W, const double* __restrict__ xp, int i) {
  const unsigned src = V.node_src[i];
  W.nodes[(i / 6) * kSampledRec + i % 6] = src != 0xFFFFu ? __ldg(xp + src) : 0.0;
}
// phase durations of foot e (phase_durations.cc:78-100), optimised durations only
__device__ __forceinline__ void SampledPhaseDurations(const SampledView& V, const SampledWork& W, const double* __restrict__ xp, int e) {
  const SampledFoot ft = V.feet[e];
  double sum = 0.0;
  for (int i = 0; i < ft.n_phases - 1; ++i) {
    const double d = __ldg(xp + ft.col0 + i);
    W.pdur[ft.phase0 + i] = d;
    sum += d;
  }
  W.pdur[ft.phase0 + ft.n_phases - 1] = ft.total - sum;
}
// polynomial durations (nodes_variables_phase_based.cc:75-86)
__device__ __forceinline__ void SampledPolyDuration(const SampledView& V, const SampledWork& W, int poly) {
  const uint32_t pm = V.poly_meta[poly];
  W.pT[poly] = pm == kPolyFixed ? V.poly_T[poly]
                                : W.pdur[V.feet[V.splines[V.poly_spline[poly]].foot].phase0 + (pm & 0xFF)] /
                                      static_cast<double>((pm >> 8) & 0xFF);
}
// Hermite coefficients c, d of (polynomial, dim) i (polynomial.cc:97-104)
__device__ __forceinline__ void SampledCoeff(const SampledView& V, const SampledWork& W, int i) {
  const int poly = i / 3, d = i - 3 * poly;
  const SampledSpline& sp = V.splines[V.poly_spline[poly]];
  const double* n0 = W.nodes + kSampledRec * (sp.node0 + (poly - sp.poly0));
  const double p0 = n0[d], v0 = n0[3 + d], p1 = n0[kSampledRec + d], v1 = n0[kSampledRec + 3 + d];
  const double T = W.pT[poly], T2 = T * T, T3 = Cube(T);
  W.coef[kSampledRec * poly + d] = -(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)) / T2;
  W.coef[kSampledRec * poly + 3 + d] = (2.0 * (p0 - p1) + T * (v0 + v1)) / T3;
}

// ---- step 1: item `it` = (spline, sample) of term tm; I: the problem's body inertia
__device__ __forceinline__ void SampledItem(const SampledView& V, const SampledWork& W, const SampledTerm& tm, const double I[9], int it) {
  const SampledHeader& H = *V.h;
  double *nodes = W.nodes, *coef = W.coef, *pT = W.pT, *pdur = W.pdur, *itl = W.itl, *im = W.im, *ix = W.ix, *ie = W.ie;
  int* iseg = W.iseg;
  const int K = tm.n_samples;
  const double wdt = tm.weight * tm.dtf;
  (void)H; (void)pdur; (void)wdt;
  const int s = it / K, k = it - s * K;
  const SampledSpline& sp = V.splines[tm.spline0 + s];
  const double tg = V.times[tm.time0 + k];
  int id = sp.n_polys - 1;                         // Spline::GetSegmentID / GetLocalTime (spline.cc:48-78)
  double acc = 0.0;
  for (int j = 0; j < sp.n_polys; ++j) {
    acc += pT[sp.poly0 + j];
    if (acc >= tg - 1e-10) { id = j; break; }
  }
  double t = tg;
  for (int j = 0; j < id; ++j) t -= pT[sp.poly0 + j];
  const int poly = sp.poly0 + id;
  const double t2 = t * t, t3 = Cube(t);
  const double* n0 = nodes + kSampledRec * (sp.node0 + id);
  double pos[3], vel[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {                    // Polynomial::GetPoint (polynomial.cc:47-61)
    const double a = n0[d], b = n0[3 + d], c = coef[kSampledRec * poly + d], dd = coef[kSampledRec * poly + 3 + d];
    pos[d] = ((a + t * b) + t2 * c) + t3 * dd;
    vel[d] = (b + (2 * t) * c) + (3 * t2) * dd;
  }
  int phase = 0;
  itl[it] = t;
  if (tm.kind == kSampledEnergy) {
    ie[it] = Dot3(pos, pos);                       // squaredNorm (energy_cost.cc:84)
    double m[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) { m[d] = sp.mscale * pos[d]; im[3 * it + d] = m[d]; }   // :112, :121
    if (sp.rt) {
      // PhaseSpline::GetJacobianOfPosWrtDurations (phase_spline.cc:67-93): dx/dT of the active polynomial
      // (polynomial.cc:236-257) through its share of the phase, and the phase the global time falls into
      const uint32_t pm = V.poly_meta[poly];
      const int nin = (pm >> 8) & 0xFF, prev = (pm >> 16) & 0xFF;
      const double T = pT[poly], T2 = T * T, T3 = Cube(T), T4 = Pow4(T);
      const double inner = 1.0 / nin;
      double dxdT[3];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const double x0 = n0[d], v0 = n0[3 + d], x1 = n0[kSampledRec + d], v1 = n0[kSampledRec + 3 + d];
        const double dx = (((t3 * (v0 + v1)) / T3 - (t2 * (2.0 * v0 + v1)) / T2) -
                           ((3 * t3) * (((2.0 * x0 - 2.0 * x1) + T * v0) + T * v1)) / T4) +
                          ((2 * t2) * (((3.0 * x0 - 3.0 * x1) + (2 * T) * v0) + T * v1)) / T3;
        dxdT[d] = inner * (dx - prev * vel[d]);
      }
      const SampledFoot ft = V.feet[sp.foot];
      phase = ft.n_phases - 1;
      double pacc = 0.0;
      for (int j = 0; j < ft.n_phases; ++j) {
        pacc += pdur[ft.phase0 + j];
        if (pacc >= tg - 1e-10) { phase = j; break; }
      }
      ix[3 * it] = Dot3(m, dxdT);
      ix[3 * it + 1] = Dot3(m, vel);
    }
  } else {
    // AngularMomentumCost (angular_momentum_cost.cc:67-101, 103-206): L = R I_b R^T omega, omega = M(euler) euler_rate
    const Trig tr = MakeTrig(pos);
    double R[9], RI[9], Rt[9], Iw[9];
    RotationMatrix(tr, R);
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) Rt[3 * i + j] = R[3 * j + i];
    Mat3Mul(R, I, RI);
    Mat3Mul(RI, Rt, Iw);
    // euler_converter.cc:59-71,133-146
    const double Mm[9] = {tr.cy * tr.cz, -tr.sz, 0.0, tr.cy * tr.sz, tr.cz, 0.0, -tr.sy, 0.0, 1.0};
    double omega[3], L[3];
    omega[0] = Mm[0] * vel[0] + Mm[1] * vel[1];
    omega[1] = Mm[3] * vel[0] + Mm[4] * vel[1];
    omega[2] = Mm[6] * vel[0] + vel[2];
    MatVec3(Iw, omega, L);
    ie[it] = Dot3(L, L);
    double m[3], u[3], v2[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) m[d] = (2.0 * wdt) * L[d];
    MatTVec3(R, omega, u);
    MatVec3(I, u, v2);
    // d omega / d euler_j = (dM / d euler_j) euler_rate  (euler_converter.cc:168-198)
    const double dw[3][3] = {
        {0.0, 0.0, 0.0},
        {-tr.sy * tr.cz * vel[0], -tr.sy * tr.sz * vel[0], -tr.cy * vel[0]},
        {-tr.cy * tr.sz * vel[0] - tr.cz * vel[1], tr.cy * tr.cz * vel[0] - tr.sz * vel[1], 0.0}};
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double D[9], a1[3], a2[3], a3[3], q[3], Iq[3], dL[3], col[3];
      RotationMatrixDeriv(tr, j, D);
      MatVec3(D, v2, a1);                          // d(R v2)/d euler_j at fixed v2      (:170, J_R_v2)
      MatTVec3(D, omega, a2);                      // d(R^T omega)/d euler_j at fixed omega (:171, term_a)
      MatTVec3(R, dw[j], a3);                      // R^T d omega / d euler_j             (term_b)
#pragma unroll
      for (int d = 0; d < 3; ++d) q[d] = a2[d] + a3[d];
      MatVec3(I, q, Iq);
      MatVec3(R, Iq, dL);                          // J2 = R (I_b term_u)
      im[3 * it + j] = Dot3(m, a1) + Dot3(m, dL);
      col[0] = Mm[j]; col[1] = Mm[3 + j]; col[2] = Mm[6 + j];   // d omega / d euler_rate_j
      MatTVec3(R, col, a3);
      MatVec3(I, a3, Iq);
      MatVec3(R, Iq, dL);
      ix[3 * it + j] = Dot3(m, dL);
    }
  }
  iseg[it] = id | (phase << 16);
}

// ---- step 2: the summand of sample k (energy_cost.cc:80-87, angular_momentum_cost.cc:97)
__device__ __forceinline__ void SampledSummand(const SampledView& V, const SampledWork& W, const SampledTerm& tm, int k) {
  const double* ie = W.ie;
  double* inst = W.inst;
  const int K = tm.n_samples;
  const double wdt = tm.weight * tm.dtf;
  double v;
  if (tm.kind == kSampledEnergy) {
    double in = 0.0;
    for (int e = 0; e < tm.n_splines / 2; ++e)   // inst += f.squaredNorm() + torque_weight_*tau.squaredNorm()  (:84)
      in += ie[(2 * e) * K + k] + V.splines[tm.spline0 + 2 * e + 1].scale * ie[(2 * e + 1) * K + k];
    v = tm.weight * in * tm.dtf;                 // :87
  } else {
    v = wdt * ie[k];                             // angular_momentum_cost.cc:97
  }
  inst[k] = v;
}

// ---- step 3: gradient column of job j, added to gp (the problem's dense gradient)
__device__ __forceinline__ void SampledJobGradient(const SampledView& V, const SampledWork& W, const SampledTerm& tm, double* __restrict__ gp, int j) {
  const double *pT = W.pT, *itl = W.itl, *im = W.im, *ix = W.ix;
  const int* iseg = W.iseg;
  const int K = tm.n_samples;
  const SampledJob jb = V.jobs[tm.job0 + j];
  double g = 0.0;
  if (jb.kind == kJobNode) {
    const int s = jb.spline, node = jb.node, deriv = jb.deriv_dim >> 2, dim = jb.deriv_dim & 3;
    const SampledSpline& sp = V.splines[tm.spline0 + s];
    for (int k = 0; k < K; ++k) {
      const int it = s * K + k;
      const int id = iseg[it] & 0xFFFF;
      if (id != node && id + 1 != node) continue;
      const int side = id == node ? 0 : 1;
      const double t = itl[it], T = pT[sp.poly0 + id];
      if (tm.kind == kSampledEnergy)
        g += im[3 * it + dim] * HermiteWeight(0, side, deriv, t, T);
      else
        g += im[3 * it + dim] * HermiteWeight(0, side, deriv, t, T) + ix[3 * it + dim] * HermiteWeight(1, side, deriv, t, T);
    }
  } else {
    // PhaseDurations::GetJacobianOfPos (phase_durations.cc:126-154), column `ph`
    const int ph = jb.node;
    const SampledFoot ft = V.feet[V.splines[tm.spline0 + jb.spline].foot];
    for (int k = 0; k < K; ++k)
      for (int q = 0; q <= jb.deriv_dim; ++q) {
        const int it = (jb.spline + q) * K + k;
        const int cur = iseg[it] >> 16;
        const bool last = cur == ft.n_phases - 1;
        const double A = ix[3 * it], B = ix[3 * it + 1];
        if (ph == cur && !last) g += A;
        if (ph < cur) g += last ? -B - A : -B;
      }
  }
  gp[jb.col] = gp[jb.col] + g;
}

}  // namespace tb

// sampled_cost.cuh — per-problem arithmetic of the time-discretised cost terms (cost.h): EnergyCost
// (costs/energy_cost.cc:38-163), AngularMomentumCost (costs/angular_momentum_cost.cc:39-208) and EEBasePosCost
// (costs/ee_base_pos_cost.cc:38-160).  The kernel
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
  const double* fix_t;
  const uint16_t* fix_seg;
  const int *comb_col, *comb_ptr, *comb_src;
};

__device__ __forceinline__ void Mat3Mul(const double A[9], const double B[9], double C[9]) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = (A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j]) + A[3 * i + 2] * B[6 + j];
}
__device__ __forceinline__ double Dot3(const double a[3], const double b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }

// shared-memory work arrays of one CTA (one problem at a time), carved by SampledCarve
struct SampledWork {
  double *nodes, *coef, *prec, *pdur, *items, *inst, *fterm, *scratch;
  int *iseg, *k0, *k1;   // iseg per item: polynomial within the spline | phase << 16; k0 / k1 per polynomial: its samples
  int* pk0;              // per (spline, phase) slot: the first sample that falls into the phase
};
// per-polynomial record (kSampledPolyRec doubles): T, T^2, T^3, T^4, their correctly rounded reciprocals (ExactDiv) and,
// for optimised durations, the polynomial's end time within its spline (accumulated like Spline::GetSegmentID's `t += d`)
enum { kPrT = 0, kPrT2, kPrT3, kPrR1, kPrR2, kPrR3, kPrR4, kPrT4, kPrEnd };

// the item arrays of one term (N = item_splines * n_samples items, tm.rec doubles each):
//   energy:      P[3N] the spline's value, t[N] local time, with optimised durations also X[6N]: the sample's products
//                m(r) * J(r, col) for the ee-schedule column of its own phase [3] and for every earlier column [3]
//   ang-mom:     P[3N] d summand / d euler angle, t[N], X[3N] d summand / d euler rate, then e[N] = |L|^2
//   ee-base-pos: P[3N] m b_R_w (the row vector of ee_base_pos_cost.cc:125), t[N] local time in the foot's motion
//                spline, X[3N] d summand / d euler angle, then tb[N] local time in the base splines, e[N] = |e|^2
struct SampledTermArrays { double *P, *t, *X; };
__device__ __forceinline__ SampledTermArrays TermArrays(const SampledWork& W, const SampledTerm& tm) {
  const int N = tm.item_splines * tm.n_samples;
  SampledTermArrays A;
  A.P = W.items + tm.scratch0;
  A.t = A.P + 3 * N;
  A.X = A.t + N;
  return A;
}

// d (pos | vel)(t) / d (node value) of a cubic Hermite polynomial (polynomial.cc:106-234) from the polynomial's record;
// side 0 start node, 1 end node; deriv: the node's position or velocity
__device__ __forceinline__ double HermiteWeight(int dxdt, int side, int deriv, double t, const double* q) {
  const double T = q[kPrT], T2 = q[kPrT2], T3 = q[kPrT3], r1 = q[kPrR1], r2 = q[kPrR2], r3 = q[kPrR3];
  const double t2 = t * t;
  if (dxdt == 0) {
    const double t3 = Cube(t);
    if (side == 0)
      return deriv == 0 ? ExactDiv(2 * t3, T3, r3) - ExactDiv(3 * t2, T2, r2) + 1 : t - ExactDiv(2 * t2, T, r1) + ExactDiv(t3, T2, r2);
    return deriv == 0 ? ExactDiv(3 * t2, T2, r2) - ExactDiv(2 * t3, T3, r3) : ExactDiv(t3, T2, r2) - ExactDiv(t2, T, r1);
  }
  if (side == 0)
    return deriv == 0 ? ExactDiv(6 * t2, T3, r3) - ExactDiv(6 * t, T2, r2) : ExactDiv(3 * t2, T2, r2) - ExactDiv(4 * t, T, r1) + 1;
  return deriv == 0 ? ExactDiv(6 * t, T2, r2) - ExactDiv(6 * t2, T3, r3) : ExactDiv(3 * t2, T2, r2) - ExactDiv(2 * t, T, r1);
}

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
  V.fix_t = reinterpret_cast<const double*>(blob + H.off_fix_t);
  V.fix_seg = reinterpret_cast<const uint16_t*>(blob + H.off_fix_seg);
  V.comb_col = reinterpret_cast<const int*>(blob + H.off_comb_col);
  V.comb_ptr = reinterpret_cast<const int*>(blob + H.off_comb_ptr);
  V.comb_src = reinterpret_cast<const int*>(blob + H.off_comb_src);
  return V;
}

// `base`: the bytes after the staged blob (SampledSharedBytes, cost.h)
__device__ __forceinline__ SampledWork SampledCarve(const SampledHeader& H, unsigned char* base) {
  SampledWork W;
  W.nodes = reinterpret_cast<double*>(base);
  W.coef = W.nodes + kSampledRec * H.n_nodes;            // c[3], d[3] of every polynomial (polynomial.cc:97-104)
  W.prec = W.coef + kSampledRec * H.n_polys;
  W.pdur = W.prec + kSampledPolyRec * H.n_polys;
  W.items = W.pdur + H.n_phases;
  W.inst = W.items + H.item_doubles;                     // per sample of every term: the summand of f
  W.fterm = W.inst + H.n_times;
  W.scratch = W.fterm + 8;                               // sums of the jobs whose columns other terms write too
  W.iseg = reinterpret_cast<int*>(W.scratch + H.n_scratch);
  W.k0 = W.iseg + H.n_items;
  W.k1 = W.k0 + H.n_polys;
  W.pk0 = W.k1 + H.n_polys;
  return W;
}

// ---- set-up of one problem; a barrier separates the three steps
// step A: node value i of 6 * n_nodes (nodes_variables.cc:68-76) ...
__device__ __forceinline__ void SampledNode(const SampledView& V, const SampledWork& W, const double* __restrict__ xp, int i) {
  const unsigned src = V.node_src[i];
  W.nodes[(i / 6) * kSampledRec + i % 6] = src != 0xFFFFu ? __ldg(xp + src) : 0.0;
}
// ... and the phase durations of foot e (phase_durations.cc:78-100), optimised durations only
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
// step B, poly in [0, n_polys + n_phase_slots): polynomial durations (nodes_variables_phase_based.cc:75-86), their powers
// and reciprocals; the last indices clear the phase slots' sample ranges
__device__ __forceinline__ void SampledPolyRecord(const SampledView& V, const SampledWork& W, int poly) {
  if (poly >= V.h->n_polys) {
    W.pk0[poly - V.h->n_polys] = 0x7FFFFFFF;
    return;
  }
  const uint32_t pm = V.poly_meta[poly];
  const double T = pm == kPolyFixed ? V.poly_T[poly]
                                    : W.pdur[V.feet[V.splines[V.poly_spline[poly]].foot].phase0 + (pm & 0xFF)] /
                                          static_cast<double>((pm >> 8) & 0xFF);
  double* q = W.prec + kSampledPolyRec * poly;
  const double T2 = T * T, T3 = Cube(T), T4 = Pow4(T);
  q[kPrT] = T; q[kPrT2] = T2; q[kPrT3] = T3; q[kPrT4] = T4;
  q[kPrR1] = 1.0 / T; q[kPrR2] = 1.0 / T2; q[kPrR3] = 1.0 / T3; q[kPrR4] = 1.0 / T4;
  W.k0[poly] = 0x7FFFFFFF;
  W.k1[poly] = 0;
}
// step C, i in [0, 3 n_polys + n_splines): Hermite coefficients c, d of (polynomial, dim) i (polynomial.cc:97-104); the
// last n_splines indices: end times of a spline's polynomials (optimised durations)
__device__ __forceinline__ void SampledCoeff(const SampledView& V, const SampledWork& W, int i) {
  if (i >= 3 * V.h->n_polys) {
    const SampledSpline& sp = V.splines[i - 3 * V.h->n_polys];
    if (!sp.rt) return;
    double acc = 0.0;
    for (int j = 0; j < sp.n_polys; ++j) {
      double* q = W.prec + kSampledPolyRec * (sp.poly0 + j);
      acc += q[kPrT];
      q[kPrEnd] = acc;
    }
    return;
  }
  const int poly = i / 3, d = i - 3 * poly;
  const SampledSpline& sp = V.splines[V.poly_spline[poly]];
  const double* n0 = W.nodes + kSampledRec * (sp.node0 + (poly - sp.poly0));
  const double p0 = n0[d], v0 = n0[3 + d], p1 = n0[kSampledRec + d], v1 = n0[kSampledRec + 3 + d];
  const double* q = W.prec + kSampledPolyRec * poly;
  const double T = q[kPrT];
  W.coef[kSampledRec * poly + d] = ExactDiv(-(3.0 * (p0 - p1) + T * (2.0 * v0 + v1)), q[kPrT2], q[kPrR2]);
  W.coef[kSampledRec * poly + 3 + d] = ExactDiv(2.0 * (p0 - p1) + T * (v0 + v1), q[kPrT3], q[kPrR3]);
}

// the term an index of a concatenated per-term range belongs to (items, samples, jobs)
__device__ __forceinline__ int TermOfItem(const SampledView& V, int it) {
  int ti = 0;
  while (ti + 1 < V.h->n_terms && it >= V.terms[ti + 1].item0) ++ti;
  return ti;
}
__device__ __forceinline__ int TermOfSample(const SampledView& V, int k) {
  int ti = 0;
  while (ti + 1 < V.h->n_terms && k >= V.terms[ti + 1].time0) ++ti;
  return ti;
}

// Spline::GetSegmentID / GetLocalTime (spline.cc:48-78) of sample k (global time tg) in spline sp: a table read where the
// durations are class constants, the reference's search and one-by-one subtraction on the problem's own durations
// otherwise.  Returns the polynomial within the spline; bit 15: the foot is in contact (class-constant durations only).
__device__ __forceinline__ int LocateSample(const SampledView& V, const SampledWork& W, const SampledSpline& sp, int k, double tg,
                                            double* t_local) {
  if (!sp.rt) {
    *t_local = V.fix_t[sp.fix0 + k];
    return V.fix_seg[sp.fix0 + k];
  }
  const double* pr = W.prec + kSampledPolyRec * sp.poly0;
  const double lim = tg - 1e-10;
  int id = sp.n_polys - 1;
#pragma unroll 4
  for (int j = 0; j < sp.n_polys; ++j)
    if (pr[j * kSampledPolyRec + kPrEnd] >= lim) { id = j; break; }
  double t = tg;
  for (int j = 0; j < id; ++j) t -= pr[j * kSampledPolyRec];
  *t_local = t;
  return id;
}
// the phase of foot ft the global time falls into, on the problem's own phase durations (phase_durations.cc:113-118)
__device__ __forceinline__ int PhaseAt(const SampledWork& W, const SampledFoot& ft, double tg) {
  int phase = ft.n_phases - 1;
  double pacc = 0.0;
  for (int j = 0; j < ft.n_phases; ++j) {
    pacc += W.pdur[ft.phase0 + j];
    if (pacc >= tg - 1e-10) { phase = j; break; }
  }
  return phase;
}
// Polynomial::GetPoint (polynomial.cc:47-61), position only
__device__ __forceinline__ void SplinePos(const SampledWork& W, const SampledSpline& sp, int id, double t, double pos[3]) {
  const double t2 = t * t, t3 = Cube(t);
  const double* n0 = W.nodes + kSampledRec * (sp.node0 + id);
  const double* cf = W.coef + kSampledRec * (sp.poly0 + id);
#pragma unroll
  for (int d = 0; d < 3; ++d) pos[d] = ((n0[d] + t * n0[3 + d]) + t2 * cf[d]) + t3 * cf[3 + d];
}

// EEBasePosCost (ee_base_pos_cost.cc:67-160), one item per sample: e = b_R_w (p_ee - p_base) - p_ref_B during swing
__device__ __forceinline__ void SampledItemEEBasePos(const SampledView& V, const SampledWork& W, const SampledTerm& tm,
                                                     const double* __restrict__ p_ref_B, int git) {
  const SampledTermArrays A = TermArrays(W, tm);
  const int K = tm.n_samples, k = git - tm.item0;
  const SampledSpline& sm = V.splines[tm.spline0];       // the foot's motion, base-lin, base-ang
  const SampledSpline& sl = V.splines[tm.spline0 + 1];
  const SampledSpline& sa = V.splines[tm.spline0 + 2];
  const double tg = V.times[tm.time0 + k];
  double t_m, t_b;
  int id_m = LocateSample(V, W, sm, k, tg, &t_m);
  const int id_b = LocateSample(V, W, sl, k, tg, &t_b) & 0x7FFF;   // base-lin and base-ang share their durations
  bool contact;
  if (!sm.rt) {
    contact = (id_m & 0x8000) != 0;
    id_m &= 0x7FFF;
  } else {                                               // PhaseDurations::IsContactPhase (phase_durations.cc:113-118)
    const SampledFoot ft = V.feet[sm.foot];
    const int ph = PhaseAt(W, ft, tg);
    contact = (ph % 2 == 0) ? ft.contact_at_start != 0 : ft.contact_at_start == 0;
  }
  atomicMin(W.k0 + sm.poly0 + id_m, k);
  atomicMax(W.k1 + sm.poly0 + id_m, k + 1);
  atomicMin(W.k0 + sl.poly0 + id_b, k);
  atomicMax(W.k1 + sl.poly0 + id_b, k + 1);
  atomicMin(W.k0 + sa.poly0 + id_b, k);
  atomicMax(W.k1 + sa.poly0 + id_b, k + 1);
  W.iseg[git] = id_m | (id_b << 12) | (contact ? 0 : 1 << 24);   // bit 24: a swing sample (the only ones that count, :78-80)
  A.t[k] = t_m;
  double* tb_ = A.X + 3 * K;
  double* e2 = tb_ + K;
  tb_[k] = t_b;
  if (contact) {
    e2[k] = 0.0;
    return;
  }
  double p_ee[3], p_base[3], eul[3], r_W[3], p_B[3], e[3], m[3], mW[3];
  SplinePos(W, sm, id_m, t_m, p_ee);
  SplinePos(W, sl, id_b, t_b, p_base);
  SplinePos(W, sa, id_b, t_b, eul);
#pragma unroll
  for (int d = 0; d < 3; ++d) r_W[d] = p_ee[d] - p_base[d];
  const Trig tr = MakeTrig(eul);
  double R[9];
  RotationMatrix(tr, R);
  MatTVec3(R, r_W, p_B);                                 // b_R_w = w_R_b^T
#pragma unroll
  for (int d = 0; d < 3; ++d) { e[d] = p_B[d] - p_ref_B[3 * tm.foot + d]; m[d] = sm.mscale * e[d]; }
  e2[k] = Dot3(e, e);
  MatVec3(R, m, mW);                                     // (m b_R_w)(c) = sum_r m(r) R(c, r)   (:125)
#pragma unroll
  for (int d = 0; d < 3; ++d) A.P[3 * k + d] = mW[d];
#pragma unroll
  for (int j = 0; j < 3; ++j) {                          // m . d(b_R_w r_W)/d euler_j  (DerivOfRotVecMult, inverse; :143-150)
    double D[9], a[3];
    RotationMatrixDeriv(tr, j, D);
    MatTVec3(D, r_W, a);
    A.X[3 * k + j] = Dot3(m, a);
  }
}

// ---- step 1: item `git` (all terms) = (spline, sample) of a term; I: the problem's body inertia
__device__ __forceinline__ void SampledItem(const SampledView& V, const SampledWork& W, const double I[9],
                                            const double* __restrict__ p_ref_B, int git) {
  const SampledTerm& tm = V.terms[TermOfItem(V, git)];
  if (tm.kind == kSampledEEBasePos) {
    SampledItemEEBasePos(V, W, tm, p_ref_B, git);
    return;
  }
  const SampledTermArrays A = TermArrays(W, tm);
  const int it = git - tm.item0;
  const int K = tm.n_samples;
  const int s = it / K, k = it - s * K;
  const SampledSpline& sp = V.splines[tm.spline0 + s];
  const double tg = V.times[tm.time0 + k];
  double t;
  const int id = LocateSample(V, W, sp, k, tg, &t) & 0x7FFF;
  const int poly = sp.poly0 + id;
  atomicMin(W.k0 + poly, k);                         // the polynomial's samples are k0 .. k1 - 1 (time order)
  atomicMax(W.k1 + poly, k + 1);
  const double* q = W.prec + kSampledPolyRec * poly;
  const double t2 = t * t, t3 = Cube(t);
  const double* n0 = W.nodes + kSampledRec * (sp.node0 + id);
  double pos[3], vel[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {                      // Polynomial::GetPoint (polynomial.cc:47-61)
    const double a = n0[d], b = n0[3 + d], c = W.coef[kSampledRec * poly + d], dd = W.coef[kSampledRec * poly + 3 + d];
    pos[d] = ((a + t * b) + t2 * c) + t3 * dd;
    vel[d] = (b + (2 * t) * c) + (3 * t2) * dd;
  }
  A.t[it] = t;
  int phase = 0;
  if (tm.kind == kSampledEnergy) {
#pragma unroll
    for (int d = 0; d < 3; ++d) A.P[3 * it + d] = pos[d];
    if (sp.rt) {
      // PhaseSpline::GetJacobianOfPosWrtDurations (phase_spline.cc:67-93): dx/dT of the active polynomial
      // (polynomial.cc:236-257) through its share of the phase, and the phase the global time falls into
      const uint32_t pm = V.poly_meta[poly];
      const int nin = (pm >> 8) & 0xFF, prev = (pm >> 16) & 0xFF;
      const double T = q[kPrT], T2 = q[kPrT2], T3 = q[kPrT3], T4 = q[kPrT4], r2 = q[kPrR2], r3 = q[kPrR3], r4 = q[kPrR4];
      const double inner = 1.0 / nin;
      double dxdT[3];
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const double x0 = n0[d], v0 = n0[3 + d], x1 = n0[kSampledRec + d], v1 = n0[kSampledRec + 3 + d];
        const double dx = ((ExactDiv(t3 * (v0 + v1), T3, r3) - ExactDiv(t2 * (2.0 * v0 + v1), T2, r2)) -
                           ExactDiv((3 * t3) * (((2.0 * x0 - 2.0 * x1) + T * v0) + T * v1), T4, r4)) +
                          ExactDiv((2 * t2) * (((3.0 * x0 - 3.0 * x1) + (2 * T) * v0) + T * v1), T3, r3);
        dxdT[d] = inner * (dx - prev * vel[d]);
      }
      const SampledFoot ft = V.feet[sp.foot];
      phase = PhaseAt(W, ft, tg);
      // PhaseDurations::GetJacobianOfPos (phase_durations.cc:126-154): +dx/dT in the column of the sample's own phase, -vel
      // in every earlier column (-vel, then -= dx/dT when the sample is in the last phase); times m(r) (energy_cost.cc:112)
      const bool in_last = phase == ft.n_phases - 1;
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const double m = sp.mscale * pos[d];
        A.X[6 * it + d] = m * dxdT[d];
        A.X[6 * it + 3 + d] = m * (in_last ? (-1 * vel[d]) - dxdT[d] : -1 * vel[d]);
      }
      atomicMin(W.pk0 + sp.ps0 + phase, k);           // the phase's samples start at pk0 (time order)
    }
  } else {
    // AngularMomentumCost (angular_momentum_cost.cc:67-101, 103-206): L = R I_b R^T omega, omega = M(euler) euler_rate
    const double wdt = tm.weight * tm.dtf;
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
    A.X[3 * tm.n_samples + it] = Dot3(L, L);
    double m[3], u[3], v2[3], mRI[3], mIw[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) m[d] = (2.0 * wdt) * L[d];
    MatTVec3(R, omega, u);
    MatVec3(I, u, v2);
    // m . (R I_b q) = (m^T R I_b) . q and m . (I_w c) = (m^T I_w) . c: the row vectors once per sample instead of a
    // 3 x n product per Jacobian (angular_momentum_cost.cc:196-203 forms J2 = R (I_b term_u), then m * J2)
    MatTVec3(RI, m, mRI);
    MatTVec3(Iw, m, mIw);
    // d omega / d euler_j = (dM / d euler_j) euler_rate  (euler_converter.cc:168-198)
    const double dw[3][3] = {
        {0.0, 0.0, 0.0},
        {-tr.sy * tr.cz * vel[0], -tr.sy * tr.sz * vel[0], -tr.cy * vel[0]},
        {-tr.cy * tr.sz * vel[0] - tr.cz * vel[1], tr.cy * tr.cz * vel[0] - tr.sz * vel[1], 0.0}};
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double D[9], a1[3], a2[3], a3[3], qv[3], col[3];
      RotationMatrixDeriv(tr, j, D);
      MatVec3(D, v2, a1);                            // d(R v2)/d euler_j at fixed v2      (:170, J_R_v2)
      MatTVec3(D, omega, a2);                        // d(R^T omega)/d euler_j at fixed omega (:171, term_a)
      MatTVec3(R, dw[j], a3);                        // R^T d omega / d euler_j             (term_b)
#pragma unroll
      for (int d = 0; d < 3; ++d) qv[d] = a2[d] + a3[d];
      A.P[3 * it + j] = Dot3(m, a1) + Dot3(mRI, qv);
      col[0] = Mm[j]; col[1] = Mm[3 + j]; col[2] = Mm[6 + j];   // d omega / d euler_rate_j
      A.X[3 * it + j] = Dot3(mIw, col);
    }
  }
  W.iseg[git] = id | (phase << 16);
}

// ---- step 2: the summand of sample `gk` (all terms; energy_cost.cc:80-87, angular_momentum_cost.cc:97)
__device__ __forceinline__ void SampledSummand(const SampledView& V, const SampledWork& W, int gk) {
  const SampledTerm& tm = V.terms[TermOfSample(V, gk)];
  const SampledTermArrays A = TermArrays(W, tm);
  const int K = tm.n_samples, k = gk - tm.time0;
  double v;
  if (tm.kind == kSampledEnergy) {
    double in = 0.0;
    for (int e = 0; e < tm.n_splines / 2; ++e) {     // inst += f.squaredNorm() + torque_weight_*tau.squaredNorm()  (:84)
      const double* f = A.P + 3 * ((2 * e) * K + k);
      const double* tau = A.P + 3 * ((2 * e + 1) * K + k);
      in += Dot3(f, f) + V.splines[tm.spline0 + 2 * e + 1].scale * Dot3(tau, tau);
    }
    v = tm.weight * in * tm.dtf;                     // :87
  } else if (tm.kind == kSampledAngMom) {
    v = (tm.weight * tm.dtf) * A.X[3 * K + k];       // angular_momentum_cost.cc:97
  } else {
    v = tm.weight * A.X[4 * K + k];                  // ee_base_pos_cost.cc:92; 0 for the samples in contact (skipped there)
  }
  W.inst[gk] = v;
}
// f of term ti: the samples added in time order
__device__ __forceinline__ void SampledTermSum(const SampledView& V, const SampledWork& W, int ti) {
  const SampledTerm& tm = V.terms[ti];
  double c = 0.0;
  for (int k = 0; k < tm.n_samples; ++k) c += W.inst[tm.time0 + k];
  W.fterm[ti] = c;
}

// a job's sum for its i-th column: added to the gradient, or parked for SampledCombine when other terms write the column too
__device__ __forceinline__ void EmitColumn(const SampledJob& jb, const SampledWork& W, double* __restrict__ gp, int i, double v) {
  if (jb.col[i] < 0) return;
  if (jb.slot < 0) gp[jb.col[i]] = gp[jb.col[i]] + v;
  else W.scratch[jb.slot + i] = v;
}
// ---- step 4: column i of the shared columns: the parked sums added in term order
__device__ __forceinline__ void SampledCombine(const SampledView& V, const SampledWork& W, double* __restrict__ gp, int i) {
  double g = gp[V.comb_col[i]];
  for (int k = V.comb_ptr[i]; k < V.comb_ptr[i + 1]; ++k) g = g + W.scratch[V.comb_src[k]];
  gp[V.comb_col[i]] = g;
}

// ---- step 3: gradient columns of job j (all terms), added to gp (the problem's dense gradient)
__device__ __forceinline__ void SampledJobGradient(const SampledView& V, const SampledWork& W, double* __restrict__ gp, int j) {
  const SampledJob jb = V.jobs[j];
  const SampledTerm& tm = V.terms[jb.kind >> 8];
  const SampledTermArrays A = TermArrays(W, tm);
  const int K = tm.n_samples;
  if (tm.kind == kSampledEEBasePos) {
    // EEBasePosCost: swing samples only (:78-80, :112-114); base-lin gets -m b_R_w, base-ang m . d(b_R_w r_W)/d euler
    // (:133-150), the foot's motion variables m b_R_w (:123-131); position weights only, no ee-schedule block
    const SampledSpline& sp = V.splines[tm.spline0 + jb.spline];
    const int* seg = W.iseg + tm.item0;
    const double* tb_ = A.X + 3 * K;
    if ((jb.kind & 0xFF) == kJobNode) {
      double gpos[3] = {0.0, 0.0, 0.0}, gvel[3] = {0.0, 0.0, 0.0};
      for (int side = 1; side >= 0; --side) {
        const int id = jb.node - side;
        if (id < 0 || id >= sp.n_polys) continue;
        const double* q = W.prec + kSampledPolyRec * (sp.poly0 + id);
        for (int k = W.k0[sp.poly0 + id]; k < W.k1[sp.poly0 + id]; ++k) {
          if (!(seg[k] >> 24 & 1)) continue;
          const double wp = HermiteWeight(0, side, 0, tb_[k], q), wv = HermiteWeight(0, side, 1, tb_[k], q);
#pragma unroll
          for (int dim = 0; dim < 3; ++dim) {
            const double v = jb.spline == 1 ? -A.P[3 * k + dim] : A.X[3 * k + dim];
            gpos[dim] += v * wp;
            gvel[dim] += v * wv;
          }
        }
      }
#pragma unroll
      for (int dim = 0; dim < 3; ++dim) {
        EmitColumn(jb, W, gp, dim, gpos[dim]);
        EmitColumn(jb, W, gp, 3 + dim, gvel[dim]);
      }
    } else {                                         // kJobColumn: a variable of the foot's motion, its node values in order
      double g = 0.0;
      for (int r = 0; r < jb.aux; ++r) {
        const int node = jb.col[1 + r] & 0xFFFF, deriv = (jb.col[1 + r] >> 16) & 0xF, dim = (jb.col[1 + r] >> 20) & 0xF;
        for (int side = 1; side >= 0; --side) {
          const int id = node - side;
          if (id < 0 || id >= sp.n_polys) continue;
          const double* q = W.prec + kSampledPolyRec * (sp.poly0 + id);
          for (int k = W.k0[sp.poly0 + id]; k < W.k1[sp.poly0 + id]; ++k) {
            if (!(seg[k] >> 24 & 1)) continue;
            g += A.P[3 * k + dim] * HermiteWeight(0, side, deriv, A.t[k], q);
          }
        }
      }
      EmitColumn(jb, W, gp, 0, g);
    }
    return;
  }
  if ((jb.kind & 0xFF) == kJobNode) {
    // the samples of the two polynomials that meet in the node, in time order: the reference's accumulation into
    // jac.coeffRef(0, col) over its sample loop.  gpos / gvel: the columns of the node's position / velocity values
    const int s = jb.spline, node = jb.node;
    const SampledSpline& sp = V.splines[tm.spline0 + s];
    double gpos[3] = {0.0, 0.0, 0.0}, gvel[3] = {0.0, 0.0, 0.0};
    for (int side = 1; side >= 0; --side) {          // end node of polynomial node-1, then start node of polynomial node
      const int id = node - side;
      if (id < 0 || id >= sp.n_polys) continue;
      const int ka = W.k0[sp.poly0 + id], kb = W.k1[sp.poly0 + id];
      const double* q = W.prec + kSampledPolyRec * (sp.poly0 + id);
      const double T = q[kPrT], T2 = q[kPrT2], T3 = q[kPrT3], r1 = q[kPrR1], r2 = q[kPrR2], r3 = q[kPrR3];
      for (int k = ka; k < kb; ++k) {
        const int it = s * K + k;
        const double t = A.t[it], t2 = t * t, t3 = Cube(t);
        // d pos / d (node position, node velocity)  (polynomial.cc:106-160, kPos)
        const double a = ExactDiv(2 * t3, T3, r3), b = ExactDiv(3 * t2, T2, r2), d = ExactDiv(t3, T2, r2);
        const double wp = side == 0 ? a - b + 1 : b - a;
        const double wv = side == 0 ? t - ExactDiv(2 * t2, T, r1) + d : d - ExactDiv(t2, T, r1);
        if (tm.kind == kSampledEnergy) {             // m(r) * J(r, col), m = (2 wdt [torque_weight]) * value  (:112-123)
#pragma unroll
          for (int dim = 0; dim < 3; ++dim) {
            const double m = sp.mscale * A.P[3 * it + dim];
            gpos[dim] += m * wp;
            gvel[dim] += m * wv;
          }
        } else {                                     // ... and d vel / d (node position, node velocity)  (kVel)
          const double e = ExactDiv(6 * t2, T3, r3), f = ExactDiv(6 * t, T2, r2), c = ExactDiv(3 * t2, T2, r2);
          const double up = side == 0 ? e - f : f - e;
          const double uv = side == 0 ? c - ExactDiv(4 * t, T, r1) + 1 : c - ExactDiv(2 * t, T, r1);
#pragma unroll
          for (int dim = 0; dim < 3; ++dim) {
            const double da = A.P[3 * it + dim], dr = A.X[3 * it + dim];
            gpos[dim] += da * wp + dr * up;
            gvel[dim] += da * wv + dr * uv;
          }
        }
      }
    }
#pragma unroll
    for (int dim = 0; dim < 3; ++dim) {
      EmitColumn(jb, W, gp, dim, gpos[dim]);
      EmitColumn(jb, W, gp, 3 + dim, gvel[dim]);
    }
  } else {
    // column `ph` of the foot's ee-schedule (energy_cost.cc:129-157): per sample in time order, force spline then torque
    // spline, m(r) * J(r, ph) for r = 0, 1, 2 with J = GetJacobianOfPosWrtDurations (the products are formed in the item
    // step).  The same operations in the same order as the reference: these sums cancel, a regrouped sum leaves the
    // 1e-12 band.
    const int ph = jb.node;
    const SampledSpline& sf = V.splines[tm.spline0 + jb.spline];
    const int last = V.feet[sf.foot].n_phases - 1;
    int k_first = K;                                   // samples of earlier phases contribute nothing to this column
    for (int c = ph; c <= last; ++c) k_first = W.pk0[sf.ps0 + c] < k_first ? W.pk0[sf.ps0 + c] : k_first;
    double g = 0.0;
    const int* seg = W.iseg + tm.item0;
    // the products of the item step: own-phase column or earlier column; a later column gets m(r) * 0
    auto add = [&](int it) {
      const int cur = seg[it] >> 16;
      const double* X = A.X + 6 * it + (ph == cur ? 0 : 3);
      const bool in = ph <= cur;
      g += in ? X[0] : 0.0;
      g += in ? X[1] : 0.0;
      g += in ? X[2] : 0.0;
    };
    const int it_f = jb.spline * K, it_t = it_f + K;
    if (jb.aux) {
#pragma unroll 4
      for (int k = k_first; k < K; ++k) { add(it_f + k); add(it_t + k); }   // force spline, then torque spline (:131-156)
    } else {
#pragma unroll 4
      for (int k = k_first; k < K; ++k) add(it_f + k);
    }
    EmitColumn(jb, W, gp, 0, g);
  }
}

}  // namespace tb

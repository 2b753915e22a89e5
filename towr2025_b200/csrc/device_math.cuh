// device_math.cuh — per-sample math of the hot path (FP64, CUDA cores; nothing here is a dense contraction).
//
//   spline point evaluation      towr/src/helpers/polynomial.cc:47-61 (same association: sum_c t^c * coeff_c)
//   Euler ZYX kinematics         towr/src/helpers/euler_converter.cc:133-221,241-304
//   single-rigid-body residual   towr/src/models/single_rigid_body_dynamics.cc:76-102
// The Jacobian side does not replay Eigen's sparse row algebra: it forms the dense 3x3 partial derivatives of a
// constraint w.r.t. the spline's (pos, vel, acc) once per sample; stage B multiplies them with Hermite basis values.
#pragma once
#include "program.h"
#include "terrain.h"

namespace tb {

struct Pt3 { double p[3], v[3], a[3]; };

enum { kNeedPos = 1, kNeedVel = 2, kNeedAcc = 4 };

// 1: the caller zeroes a segment's whole G buffer cooperatively before stage A (the evaluation kernels, classes with
// optimised durations), so the items skip their serial zero fill of the all-column weight vectors - 60 to 100 shared-memory
// stores per vector and thread.  0 (host emulation): every item zeroes what it owns, G starts as NaN to catch stray reads.
#ifndef TB_PREZERO_G
#define TB_PREZERO_G 0
#endif

// class tables in shared memory (BlobLayout, program.h)
struct Tables {
  const double* B;
  const double* times;      // 3 per entry
  const double* durs;       // 3 per entry
  const double* rdurs;      // 2 per entry
  const SegmentRec* segments;
  const ItemC* items;
  const uint32_t* samples;
  const uint32_t* polys;
  const uint16_t* node_src;
  const RtSpline* rt_splines;   // optimised durations only
  const RtFoot* rt_feet;
  const uint32_t* poly_meta;
};

__device__ __forceinline__ Tables MakeTables(const unsigned char* blob, const BlobLayout& L) {
  Tables t;
  t.B = reinterpret_cast<const double*>(blob + L.off_B);
  t.times = reinterpret_cast<const double*>(blob + L.off_times);
  t.durs = reinterpret_cast<const double*>(blob + L.off_durs);
  t.rdurs = reinterpret_cast<const double*>(blob + L.off_rdurs);
  t.segments = reinterpret_cast<const SegmentRec*>(blob + L.off_segments);
  t.items = reinterpret_cast<const ItemC*>(blob + L.off_items);
  t.samples = reinterpret_cast<const uint32_t*>(blob + L.off_samples);
  t.polys = reinterpret_cast<const uint32_t*>(blob + L.off_polys);
  t.node_src = reinterpret_cast<const uint16_t*>(blob + L.off_node_src);
  t.rt_splines = reinterpret_cast<const RtSpline*>(blob + L.off_rt_splines);
  t.rt_feet = reinterpret_cast<const RtFoot*>(blob + L.off_rt_feet);
  t.poly_meta = reinterpret_cast<const uint32_t*>(blob + L.off_poly_meta);
  return t;
}

// a / b, correctly rounded, from the correctly rounded reciprocal y = RN(1/b) of a class constant b (Markstein:
// q = RN(a y), r = a - b q exactly by FMA, RN(q + r y) = RN(a / b)): the result is bit-identical to the reference's
// true division without the ~45-instruction software divide.  (fma is explicit: -fmad=false does not touch it.)
__device__ __forceinline__ double ExactDiv(double a, double b, double y) {
  const double q = a * y;
  const double r = fma(-b, q, a);
  return fma(r, y, q);
}

// Per-problem arrays are interleaved over the P problems a CTA evaluates in lock step: element i of this thread's
// problem is base[i * P] (the base pointers are already offset by the problem's lane).
// nodes: element slot*6 + deriv*3 + dim ; coef: element poly*6 + {0:c,3:d} + dim
template <int NEED, int P>
__device__ __forceinline__ void EvalSpline(const Tables& tb_, const double* __restrict__ nodes,
                                           const double* __restrict__ coef, int sample, Pt3& o) {
  const uint32_t sr = tb_.samples[sample];
  const int poly = sr & 0xFFFF;
  const double* tt = tb_.times + 3 * (sr >> 16);
  const double t = tt[0], t2 = tt[1], t3 = tt[2];
  const int n0 = tb_.polys[poly] & 0xFFFF;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double a = nodes[(n0 * 6 + d) * P], b = nodes[(n0 * 6 + 3 + d) * P];
    const double c = coef[(poly * 6 + d) * P], dd = coef[(poly * 6 + 3 + d) * P];
    if (NEED & kNeedPos) o.p[d] = ((a + t * b) + t2 * c) + t3 * dd;
    if (NEED & kNeedVel) o.v[d] = (b + (2 * t) * c) + (3 * t2) * dd;
    if (NEED & kNeedAcc) o.a[d] = 2.0 * c + (6 * t) * dd;
  }
}

// t^3 rounded (almost always) correctly, like the reference's std::pow(t, 3): t*t with its rounding error carried by FMA
__device__ __forceinline__ double Cube(double t) {
  const double p = t * t;
  const double e = fma(t, t, -p);
  return fma(p, t, e * t);
}

// t^4 likewise (std::pow(T, 4) in polynomial.cc:240)
__device__ __forceinline__ double Pow4(double t) {
  const double p = t * t;
  const double e = fma(t, t, -p);
  const double hi = p * p;
  return hi + (fma(p, p, -hi) + 2.0 * p * e);
}

// ---- PhaseSpline with optimised durations: the polynomial a global time falls into and its local time depend on
// the problem's current durations.  rtd: per-problem polynomial records of kRtd doubles (T, T^2, T^3, 1/T^2, 1/T^3,
// 1/T, 1/T^4 and the end time of the polynomial within its spline, accumulated left to right like Spline::GetSegmentID's
// `t += d`), element i at rtd[i * NS].  The reciprocals are true divisions (correctly rounded), so ExactDiv with them is
// bit-identical to the reference's divisions at 3 instructions instead of a ~45-instruction software divide.
constexpr int kRtd = 8;
enum { kRtT = 0, kRtT2, kRtT3, kRtR2, kRtR3, kRtR1, kRtR4, kRtEnd };
struct RtPoint {
  int poly;            // global polynomial index
  double t, T;         // local time, polynomial duration
};

// Spline::GetSegmentID / GetLocalTime (spline.cc:48-78): first i with sum_{j<=i} d_j >= t - 1e-10; the local time
// subtracts the earlier durations one by one.  Then polynomial.cc:47-61 with the reference's association.
template <int NEED, int NS>
__device__ __forceinline__ RtPoint EvalSplineRt(const Tables& tb_, const double* __restrict__ nodes,
                                                const double* __restrict__ coef, const double* __restrict__ rtd,
                                                int sample, Pt3& o) {
  const uint32_t sr = tb_.samples[sample];
  const RtSpline sp = tb_.rt_splines[sr & 0xFFFF];
  const double tg = tb_.times[3 * (sr >> 16)];
  // the polynomial end times are prefix sums made once per problem (stage D): the search only loads and compares
  const double* pr = rtd + sp.poly0 * kRtd * NS;
  const double lim = tg - 1e-10;
  int id = sp.n_polys - 1;
#pragma unroll 4
  for (int i = 0; i < sp.n_polys; ++i) {
    if (pr[(i * kRtd + kRtEnd) * NS] >= lim) { id = i; break; }
  }
  double t = tg;
  for (int i = 0; i < id; ++i) t -= pr[i * kRtd * NS];
  RtPoint r;
  r.poly = sp.poly0 + id;
  r.t = t;
  r.T = pr[id * kRtd * NS];
  const double t2 = t * t, t3 = Cube(t);
  const int n0 = tb_.polys[r.poly] & 0xFFFF;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double a = nodes[(n0 * 6 + d) * NS], b = nodes[(n0 * 6 + 3 + d) * NS];
    const double c = coef[(r.poly * 6 + d) * NS], dd = coef[(r.poly * 6 + 3 + d) * NS];
    if (NEED & kNeedPos) o.p[d] = ((a + t * b) + t2 * c) + t3 * dd;
    if (NEED & kNeedVel) o.v[d] = (b + (2 * t) * c) + (3 * t2) * dd;
    if (NEED & kNeedAcc) o.a[d] = 2.0 * c + (6 * t) * dd;
  }
  return r;
}

// W[c] = d pos(t) / d (variable c of the spline's set), all columns (NodeSpline::FillJacobianWrtNodes,
// node_spline.cc:84-112, on the pre-filled structure of phase_spline.cc:45-51): zero except the variables mapped to
// the active polynomial's two nodes; closed forms polynomial.cc:106-160 (kPos).  W: element c at W[c * GS].
template <int GS, int NS>
__device__ __forceinline__ void WriteHermiteWeights(const Tables& tb_, const double* __restrict__ rtd, const RtSpline& sp,
                                                    const RtPoint& r, double* W) {
  if (!TB_PREZERO_G)
    for (int c = 0; c < sp.n_cols; ++c) W[c * GS] = 0.0;
  const double* q = rtd + r.poly * kRtd * NS;
  const double t = r.t, T = r.T, T2 = q[kRtT2 * NS], T3 = q[kRtT3 * NS];
  const double r1 = q[kRtR1 * NS], r2 = q[kRtR2 * NS], r3 = q[kRtR3 * NS];
  const double t2 = t * t, t3 = Cube(t);
  double b[4];
  b[0] = ExactDiv(2 * t3, T3, r3) - ExactDiv(3 * t2, T2, r2) + 1;     // start node, position
  b[1] = t - ExactDiv(2 * t2, T, r1) + ExactDiv(t3, T2, r2);          // start node, velocity
  b[2] = ExactDiv(3 * t2, T2, r2) - ExactDiv(2 * t3, T3, r3);         // end node, position
  b[3] = ExactDiv(t3, T2, r2) - ExactDiv(t2, T, r1);                  // end node, velocity
  const int n0 = tb_.polys[r.poly] & 0xFFFF;
#pragma unroll
  for (int side = 0; side < 2; ++side)
#pragma unroll
    for (int deriv = 0; deriv < 2; ++deriv)
#pragma unroll
      for (int d = 0; d < 3; ++d) {
        const unsigned src = tb_.node_src[(n0 + side) * 6 + deriv * 3 + d];
        if (src != 0xFFFFu) {
          double* w = W + (static_cast<int>(src) - sp.col0) * GS;
          *w = *w + b[side * 2 + deriv];
        }
      }
}

// PhaseSpline::GetDerivativeOfPosWrtPhaseDuration (phase_spline.cc:77-93) with
// CubicHermitePolynomial::GetDerivativeOfPosWrtDuration (polynomial.cc:236-257); returns the current phase
template <int NS>
__device__ __forceinline__ int PosWrtPhaseDuration(const Tables& tb_, const double* __restrict__ nodes,
                                                   const double* __restrict__ rtd, const RtPoint& r, const double vel[3],
                                                   double dxdT[3]) {
  const uint32_t pm = tb_.poly_meta[r.poly];
  const int phase = pm & 0xFF, nin = (pm >> 8) & 0xFF, prev = (pm >> 16) & 0xFF;
  const double* q = rtd + r.poly * kRtd * NS;
  const double t = r.t, T = r.T, T2 = q[kRtT2 * NS], T3 = q[kRtT3 * NS], T4 = Pow4(T);
  const double r2 = q[kRtR2 * NS], r3 = q[kRtR3 * NS], r4 = q[kRtR4 * NS];
  const double t2 = t * t, t3 = Cube(t);
  const int n0 = tb_.polys[r.poly] & 0xFFFF;
  const double inner = 1.0 / nin;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double x0 = nodes[(n0 * 6 + d) * NS], v0 = nodes[(n0 * 6 + 3 + d) * NS];
    const double x1 = nodes[(n0 * 6 + 6 + d) * NS], v1 = nodes[(n0 * 6 + 9 + d) * NS];
    const double dx = ((ExactDiv(t3 * (v0 + v1), T3, r3) - ExactDiv(t2 * (2.0 * v0 + v1), T2, r2)) -
                       ExactDiv((3 * t3) * (((2.0 * x0 - 2.0 * x1) + T * v0) + T * v1), T4, r4)) +
                      ExactDiv((2 * t2) * (((3.0 * x0 - 3.0 * x1) + (2 * T) * v0) + T * v1), T3, r3);
    dxdT[d] = inner * (dx - prev * vel[d]);
  }
  return phase;
}

struct Trig { double sx, cx, sy, cy, sz, cz; };

__device__ __forceinline__ Trig MakeTrig(const double e[3]) {
  Trig t;
  sincos(e[0], &t.sx, &t.cx);
  sincos(e[1], &t.sy, &t.cy);
  sincos(e[2], &t.sz, &t.cz);
  return t;
}

// base-to-world rotation (euler_converter.cc:207-221), row-major R[3*i+j]
__device__ __forceinline__ void RotationMatrix(const Trig& t, double R[9]) {
  R[0] = t.cy * t.cz; R[1] = t.cz * t.sx * t.sy - t.cx * t.sz; R[2] = t.sx * t.sz + t.cx * t.cz * t.sy;
  R[3] = t.cy * t.sz; R[4] = t.cx * t.cz + t.sx * t.sy * t.sz; R[5] = t.cx * t.sy * t.sz - t.cz * t.sx;
  R[6] = -t.sy;       R[7] = t.cy * t.sx;                      R[8] = t.cx * t.cy;
}

// dR/d(euler c), c = 0 roll, 1 pitch, 2 yaw (coefficients of jac_x, jac_y, jac_z in euler_converter.cc:241-268)
__device__ __forceinline__ void RotationMatrixDeriv(const Trig& t, int c, double D[9]) {
  if (c == 0) {
    D[0] = 0.0; D[1] = t.sx * t.sz + t.cx * t.cz * t.sy;  D[2] = t.cx * t.sz - t.cz * t.sx * t.sy;
    D[3] = 0.0; D[4] = t.cx * t.sy * t.sz - t.cz * t.sx;  D[5] = -t.cx * t.cz - t.sx * t.sy * t.sz;
    D[6] = 0.0; D[7] = t.cx * t.cy;                       D[8] = -t.cy * t.sx;
  } else if (c == 1) {
    D[0] = -t.cz * t.sy; D[1] = t.cy * t.cz * t.sx; D[2] = t.cx * t.cy * t.cz;
    D[3] = -t.sy * t.sz; D[4] = t.cy * t.sx * t.sz; D[5] = t.cx * t.cy * t.sz;
    D[6] = -t.cy;        D[7] = -t.sx * t.sy;       D[8] = -t.cx * t.sy;
  } else {
    D[0] = -t.cy * t.sz; D[1] = -t.cx * t.cz - t.sx * t.sy * t.sz; D[2] = t.cz * t.sx - t.cx * t.sy * t.sz;
    D[3] = t.cy * t.cz;  D[4] = t.cz * t.sx * t.sy - t.cx * t.sz;  D[5] = t.sx * t.sz + t.cx * t.cz * t.sy;
    D[6] = 0.0;          D[7] = 0.0;                               D[8] = 0.0;
  }
}

__device__ __forceinline__ void MatVec3(const double A[9], const double v[3], double o[3]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) o[i] = (A[3 * i] * v[0] + A[3 * i + 1] * v[1]) + A[3 * i + 2] * v[2];
}
__device__ __forceinline__ void MatTVec3(const double A[9], const double v[3], double o[3]) {
#pragma unroll
  for (int i = 0; i < 3; ++i) o[i] = (A[i] * v[0] + A[3 + i] * v[1]) + A[6 + i] * v[2];
}
__device__ __forceinline__ void Cross3(const double a[3], const double b[3], double o[3]) {
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}

}  // namespace tb

// items.cuh — stage A: one thread evaluates one item of one problem (a time sample or a node of one constraint
// family), writes the item's constraint rows into g and its partial-derivative slots into the shared G buffer.
// Level-0 items depend only on the splines; level-1 items also read the scratch slots a level-0 item wrote.
#pragma once
#include "device_math.cuh"

namespace tb {

// Memory layout of the per-problem arrays an item touches: element i of this thread's problem is at
//   G[i * GS] (G buffer and scratch), nodes[i * NS] / coef[i * NS] (spline tables), g[i * RS] (constraint rows).
// Lockstep<P>: the fused kernel's shared-memory arrays, interleaved over the P problems of a CTA (and the host
// emulation, P = 1).  The tiled pipeline (tiled.cu) uses Tiled<W>.
// RT: the structure class optimises the phase durations (PhaseSpline path): ee-motion / ee-force / ee-torque samples are
// resolved per problem, their Jacobian weights and the ee-schedule columns are written by the items.
// CURV: the class declares the curvature block of the force / torque constraints (Gap; ItemCurvature below).  A template
// flag, not a run-time branch: with the code merely present the other classes' kernels lose registers (measured: +19 %).
template <int P, bool RT_ = false, bool CURV_ = false>
struct Lockstep { static constexpr int GS = P, NS = P, RS = 1; static constexpr bool RT = RT_, CURV = CURV_; };

template <class L>
struct ItemCtx {
  Tables T;
  const double* nodes;   // interleaved, offset by the problem lane (see EvalSpline)
  const double* coef;
  double* G;             // segment-relative G buffer (shared memory), element i at G[i * L::GS]
  double* g;             // this problem's constraint values (global), or nullptr
  const ProblemConst* pc;
  const double* rtd;     // RT: per-problem polynomial durations (device_math.cuh), interleaved like nodes
  const double* pdur;    // RT: per-problem phase durations, element RtFoot::phase0 + phase
  int n_ee;
};

// the problem's height map at (x, y).  Terrains given as data (terrain id 8, ProblemConst::table) are read only by the
// CURV instantiations: a class that admits them is built with that flag (TerrainClassBits, terrain.h), so the analytic
// terrains' kernels carry no table code
template <class L>
__device__ __forceinline__ TerrainSample TerrainAt(const ItemCtx<L>& cx, double x, double y) {
  if constexpr (L::CURV) {
    if (cx.pc->terrain_id == kTerrainTableId) return EvalTerrainTable(cx.pc->table, x);
  }
  return EvalTerrain(cx.pc->terrain_id, x, y);
}

// column j of PhaseDurations::GetJacobianOfPos (phase_durations.cc:126-154): d pos / d (phase duration j)
__device__ __forceinline__ void DurationColumn(int j, int cur, int n_phases, const double dxdT[3], const double xd[3], double c[3]) {
  const bool in_last = cur == n_phases - 1;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    double v = 0.0;
    if (!in_last && j == cur) v = dxdT[d];
    if (j < cur) {
      v = -1.0 * xd[d];
      if (in_last) v = v - dxdT[d];
    }
    c[d] = v;
  }
}

#define TB_G(i) G[(i) * L::GS]

// ------------------------------------------------------------------------------------------------ dynamic, level 0
// Everything of a dynamic sample that depends on the euler spline only (euler_converter.cc:133-221):
// thetadot, thetaddot, trig, R, R I_b, I_w = R I_b R^T ((R I_b) first, inner index ascending), M, Mdot,
// omega = M thetadot, omegadot = Mdot thetadot + M thetaddot, I_w omega, I_b R^T omegadot, I_b R^T omega.
template <class L>
__device__ __forceinline__ void ItemDynPre(const ItemCtx<L>& cx, const ItemC& it) {
  Pt3 ang;
  EvalSpline<kNeedPos | kNeedVel | kNeedAcc, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, ang);
  const Trig tr = MakeTrig(ang.p);
  double* G = cx.G + it.gslot * L::GS;
  double R[9];
  RotationMatrix(tr, R);
  const double yd = ang.v[1], zd = ang.v[2];
  const double M00 = tr.cy * tr.cz, M01 = -tr.sz, M10 = tr.cy * tr.sz, M11 = tr.cz, M20 = -tr.sy;
  const double Md00 = -tr.cz * tr.sy * yd - tr.cy * tr.sz * zd, Md01 = -tr.cz * zd;
  const double Md10 = tr.cy * tr.cz * zd - tr.sy * tr.sz * yd, Md11 = -tr.sz * zd;
  const double Md20 = -tr.cy * yd;
  double w[3], wd[3];
  w[0] = M00 * ang.v[0] + M01 * ang.v[1];
  w[1] = M10 * ang.v[0] + M11 * ang.v[1];
  w[2] = M20 * ang.v[0] + 1.0 * ang.v[2];
  wd[0] = (Md00 * ang.v[0] + Md01 * ang.v[1]) + (M00 * ang.a[0] + M01 * ang.a[1]);
  wd[1] = (Md10 * ang.v[0] + Md11 * ang.v[1]) + (M10 * ang.a[0] + M11 * ang.a[1]);
  wd[2] = (Md20 * ang.v[0]) + (M20 * ang.a[0] + 1.0 * ang.a[2]);
#pragma unroll
  for (int d = 0; d < 3; ++d) { TB_G(kDpV + d) = ang.v[d]; TB_G(kDpA + d) = ang.a[d]; TB_G(kDpW + d) = w[d]; TB_G(kDpWd + d) = wd[d]; }
  TB_G(kDpTrig + 0) = tr.sx; TB_G(kDpTrig + 1) = tr.cx; TB_G(kDpTrig + 2) = tr.sy;
  TB_G(kDpTrig + 3) = tr.cy; TB_G(kDpTrig + 4) = tr.sz; TB_G(kDpTrig + 5) = tr.cz;
  TB_G(kDpM + 0) = M00; TB_G(kDpM + 1) = M01; TB_G(kDpM + 2) = M10; TB_G(kDpM + 3) = M11; TB_G(kDpM + 4) = M20;
  TB_G(kDpMd + 0) = Md00; TB_G(kDpMd + 1) = Md01; TB_G(kDpMd + 2) = Md10; TB_G(kDpMd + 3) = Md11; TB_G(kDpMd + 4) = Md20;
  const double* Ib = cx.pc->I;
  double RI[9], Iw[9];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) RI[3 * i + j] = (R[3 * i] * Ib[j] + R[3 * i + 1] * Ib[3 + j]) + R[3 * i + 2] * Ib[6 + j];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) Iw[3 * i + j] = (RI[3 * i] * R[3 * j] + RI[3 * i + 1] * R[3 * j + 1]) + RI[3 * i + 2] * R[3 * j + 2];
#pragma unroll
  for (int i = 0; i < 9; ++i) { TB_G(kDpR + i) = R[i]; TB_G(kDpRI + i) = RI[i]; TB_G(kDpIw + i) = Iw[i]; }
  double Iww[3], tmp[3], v11[3], v21[3];
  MatVec3(Iw, w, Iww);
  MatTVec3(R, wd, tmp);  MatVec3(Ib, tmp, v11);     // I_b R^T wd
  MatTVec3(R, w, tmp);   MatVec3(Ib, tmp, v21);     // I_b R^T w
#pragma unroll
  for (int d = 0; d < 3; ++d) { TB_G(kDpIww + d) = Iww[d]; TB_G(kDpV11 + d) = v11[d]; TB_G(kDpV21 + d) = v21[d]; }
}

// ------------------------------------------------------------------------------------------------ dynamic, per foot
// f_e and (c - p_e) slots (ee-motion / ee-force blocks) and the foot's torque term Cross(f_e)(c - p_e) + tau_e
// (dynamic_constraint.cc:63-68, single_rigid_body_dynamics.cc:80-86).
template <class L>
__device__ __forceinline__ void ItemDynFoot(const ItemCtx<L>& cx, const ItemC& it) {
  const int E = it.n_ee, e = it.ee;
  Pt3 lin, f, p, tq;
  EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, lin);
  RtPoint rp, rf, rt;
  if constexpr (L::RT) {
    rp = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 1, p);
    rf = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 2, f);
    rt = EvalSplineRt<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 3, tq);
  } else {
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 1, p);
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 2, f);
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 3, tq);
  }
  double* G = cx.G + it.gslot * L::GS;
  double r[3] = {lin.p[0] - p.p[0], lin.p[1] - p.p[1], lin.p[2] - p.p[2]};
  double cr[3];
  Cross3(f.p, r, cr);
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    TB_G(kDynF + 3 * e + d) = f.p[d];
    TB_G(kDynF + 3 * E + 3 * e + d) = r[d];
    TB_G(kDynF + 6 * E + 3 * e + d) = cr[d] + tq.p[d];
  }
  if constexpr (L::RT) {
    // negated copies (the Cross() entries with a minus sign), the Hermite weights of the three splines and the finished
    // ee-schedule block 6 x (P-1): GetJacobianWrtForce(J_f,dT) + GetJacobianWrtEEPos(J_p,dT)
    // (dynamic_constraint.cc:99-104, single_rigid_body_dynamics.cc:168-204)
    const int neg = DynNegOffset(E);
#pragma unroll
    for (int d = 0; d < 3; ++d) { TB_G(neg + 3 * e + d) = -f.p[d]; TB_G(neg + 3 * E + 3 * e + d) = -r[d]; }
    const RtSpline sm = cx.T.rt_splines[e], sf = cx.T.rt_splines[E + e], st = cx.T.rt_splines[2 * E + e];
    double* X = cx.G + it.aux * L::GS;                 // this foot's block: W_p | W_f | W_tau | schedule[6][P-1]
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, sm, rp, X);
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, sf, rf, X + sm.n_cols * L::GS);
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, st, rt, X + (sm.n_cols + sf.n_cols) * L::GS);
    double* S = X + (sm.n_cols + sf.n_cols + st.n_cols) * L::GS;
    const RtFoot ft = cx.T.rt_feet[e];
    double dpT[3], dfT[3];
    const int cur_p = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rp, p.v, dpT);
    const int cur_f = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rf, f.v, dfT);
    const int nc = ft.n_phases - 1;
    for (int j = 0; j < nc; ++j) {
      double cf[3], cp[3];
      DurationColumn(j, cur_f, ft.n_phases, dfT, f.v, cf);
      DurationColumn(j, cur_p, ft.n_phases, dpT, p.v, cp);
      // -((-Cross(r)) J_f,dT) rows, then + -(Cross(f) (-J_p,dT)) rows; stored off-diagonal entries in column order
      const double af0 = -((r[2]) * cf[1] + (-r[1]) * cf[2]);
      const double af1 = -((-r[2]) * cf[0] + (r[0]) * cf[2]);
      const double af2 = -((r[1]) * cf[0] + (-r[0]) * cf[1]);
      const double ap0 = -((-f.p[2]) * (-cp[1]) + (f.p[1]) * (-cp[2]));
      const double ap1 = -((f.p[2]) * (-cp[0]) + (-f.p[0]) * (-cp[2]));
      const double ap2 = -((-f.p[1]) * (-cp[0]) + (f.p[0]) * (-cp[1]));
      S[(0 * nc + j) * L::GS] = af0 + ap0;
      S[(1 * nc + j) * L::GS] = af1 + ap1;
      S[(2 * nc + j) * L::GS] = af2 + ap2;
#pragma unroll
      for (int d = 0; d < 3; ++d) S[((3 + d) * nc + j) * L::GS] = -cf[d];
    }
  }
}

// ------------------------------------------------------------------------------------------------ dynamic, level 1
// g rows (dynamic_constraint.cc:63-68,128-148 + single_rigid_body_dynamics.cc:76-102, same operation order; the feet
// are summed in the reference's order from the DynFoot slots).
template <class L>
__device__ __forceinline__ void ItemDynWrench(const ItemCtx<L>& cx, const ItemC& it) {
  if (!cx.g) return;
  const int E = it.n_ee;
  Pt3 lin;
  EvalSpline<kNeedAcc, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, lin);
  const double* G = cx.G + it.gslot * L::GS;
  const double* S = cx.G + it.aux * L::GS;            // scratch of the level-0 parent
  double f_sum[3] = {0.0, 0.0, 0.0}, tau_sum[3] = {0.0, 0.0, 0.0};
  for (int e = 0; e < E; ++e) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      tau_sum[d] = tau_sum[d] + TB_G(kDynF + 6 * E + 3 * e + d);
      f_sum[d] = f_sum[d] + TB_G(kDynF + 3 * e + d);
    }
  }
  double w[3], wd[3], Iww[3], Iw[9], Iw_wd[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) { w[d] = S[(kDpW + d) * L::GS]; wd[d] = S[(kDpWd + d) * L::GS]; Iww[d] = S[(kDpIww + d) * L::GS]; }
#pragma unroll
  for (int i = 0; i < 9; ++i) Iw[i] = S[(kDpIw + i) * L::GS];
  MatVec3(Iw, wd, Iw_wd);
  double gy[3];   // Cross(w) * (I_w w): off-diagonal entries in column order
  gy[0] = ((-w[2]) * Iww[1]) + w[1] * Iww[2];
  gy[1] = (w[2] * Iww[0]) + (-w[0]) * Iww[2];
  gy[2] = ((-w[1]) * Iww[0]) + w[0] * Iww[1];
  const double m = cx.pc->mass;
  const double grav = 9.80665;
  double* gr = cx.g + it.row * L::RS;
#pragma unroll
  for (int d = 0; d < 3; ++d) gr[(d) * L::RS] = (Iw_wd[d] + gy[d]) - tau_sum[d];
  gr[(3) * L::RS] = (m * lin.a[0] - f_sum[0]) - 0.0;
  gr[(4) * L::RS] = (m * lin.a[1] - f_sum[1]) - 0.0;
  gr[(5) * L::RS] = (m * lin.a[2] - f_sum[2]) - (-m * grav);
}

// rows of the base-lin angular block (sub = r: one row, sub = 3: all): entry(r, (sd,d)) = -(sum_e Cross(f_e)(r,d) * Bpos[sd]), feet summed
// in order
template <class L>
__device__ __forceinline__ void ItemDynJlin(const ItemCtx<L>& cx, const ItemC& it) {
  const int E = it.n_ee;
  double* G = cx.G + it.gslot * L::GS;
  const double* Bl = cx.T.B + it.b0;
  const double b4[4] = {Bl[0], Bl[1], Bl[2], Bl[3]};
  // sub = 0..2: that row only; sub = 3: all three rows (one item per sample)
  const int r0 = it.sub < 3 ? it.sub : 0, r1 = it.sub < 3 ? it.sub + 1 : 3;
  for (int r = r0; r < r1; ++r) {
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (d == r) {              // no entry uses these slots; keep them finite for the status check
#pragma unroll
        for (int sd = 0; sd < 4; ++sd) TB_G(kDynJlin + r * 12 + sd * 3 + d) = 0.0;
        continue;
      }
      const int k = 3 - r - d;
      const double sg = ((d - r + 3) % 3 == 2) ? 1.0 : -1.0;
      double acc[4];
      {
        const double sf = sg * TB_G(kDynF + k);
#pragma unroll
        for (int sd = 0; sd < 4; ++sd) acc[sd] = sf * b4[sd];
      }
      for (int e = 1; e < E; ++e) {
        const double sf = sg * TB_G(kDynF + 3 * e + k);
#pragma unroll
        for (int sd = 0; sd < 4; ++sd) acc[sd] = acc[sd] + sf * b4[sd];
      }
#pragma unroll
      for (int sd = 0; sd < 4; ++sd) TB_G(kDynJlin + r * 12 + sd * 3 + d) = -acc[sd];
    }
  }
}

// One euler column c of d(angular rows)/d(base-ang variables):
//   value(r, (sd,c)) = A_pos[r][c]*Bp[sd] + A_vel[r][c]*Bv[sd] + A_acc[r][c]*Ba[sd]
// with A_* the partial derivatives of I_w wd + w x (I_w w) w.r.t. the euler spline's pos, vel, acc
// (single_rigid_body_dynamics.cc:124-166 through euler_converter.cc:85-131,223-304, regrouped).
template <class L>
__device__ __forceinline__ void ItemDynCol(const ItemCtx<L>& cx, const ItemC& it) {
  const int c = it.sub;
  const double* S = cx.G + it.aux * L::GS;
#define TB_S(i) S[(i) * L::GS]
  Trig tr;
  tr.sx = TB_S(kDpTrig); tr.cx = TB_S(kDpTrig + 1); tr.sy = TB_S(kDpTrig + 2);
  tr.cy = TB_S(kDpTrig + 3); tr.sz = TB_S(kDpTrig + 4); tr.cz = TB_S(kDpTrig + 5);
  const double v0 = TB_S(kDpV), v1 = TB_S(kDpV + 1), yd = v1, zd = TB_S(kDpV + 2);
  const double a0 = TB_S(kDpA), a1 = TB_S(kDpA + 1);
  // d(omega)/d(theta_c) = dM_c thetadot ; d(omegadot)/d(theta_c) = dMdot_c thetadot + dM_c thetaddot
  // d(omega)/d(thetadot_c) = M(:,c)     ; d(omegadot)/d(thetadot_c) = dMdot/d(thetadot_c) thetadot + Mdot(:,c)
  double Wp[3], Wdp[3], Mc[3], Wdv[3];
  if (c == 0) {
    Wp[0] = Wp[1] = Wp[2] = 0.0;
    Wdp[0] = Wdp[1] = Wdp[2] = 0.0;
    Mc[0] = TB_S(kDpM); Mc[1] = TB_S(kDpM + 2); Mc[2] = TB_S(kDpM + 4);
    Wdv[0] = TB_S(kDpMd); Wdv[1] = TB_S(kDpMd + 2); Wdv[2] = TB_S(kDpMd + 4);
  } else if (c == 1) {
    const double dM00 = -tr.sy * tr.cz, dM10 = -tr.sy * tr.sz, dM20 = -tr.cy;
    Wp[0] = dM00 * v0; Wp[1] = dM10 * v0; Wp[2] = dM20 * v0;
    const double dMd00 = -tr.cy * tr.cz * yd + tr.sy * tr.sz * zd;
    const double dMd10 = -tr.cy * tr.sz * yd - tr.cz * tr.sy * zd;
    const double dMd20 = tr.sy * yd;
    Wdp[0] = dMd00 * v0 + dM00 * a0;
    Wdp[1] = dMd10 * v0 + dM10 * a0;
    Wdp[2] = dMd20 * v0 + dM20 * a0;
    Mc[0] = TB_S(kDpM + 1); Mc[1] = TB_S(kDpM + 3); Mc[2] = 0.0;
    Wdv[0] = (-tr.cz * tr.sy) * v0 + TB_S(kDpMd + 1);
    Wdv[1] = (-tr.sy * tr.sz) * v0 + TB_S(kDpMd + 3);
    Wdv[2] = (-tr.cy) * v0;
  } else {
    const double dM00 = -tr.cy * tr.sz, dM01 = -tr.cz, dM10 = tr.cy * tr.cz, dM11 = -tr.sz;
    Wp[0] = dM00 * v0 + dM01 * v1;
    Wp[1] = dM10 * v0 + dM11 * v1;
    Wp[2] = 0.0;
    const double dMd00 = tr.sy * tr.sz * yd - tr.cy * tr.cz * zd, dMd01 = tr.sz * zd;
    const double dMd10 = -tr.cz * tr.sy * yd - tr.cy * tr.sz * zd, dMd11 = -tr.cz * zd;
    Wdp[0] = (dMd00 * v0 + dMd01 * v1) + (dM00 * a0 + dM01 * a1);
    Wdp[1] = (dMd10 * v0 + dMd11 * v1) + (dM10 * a0 + dM11 * a1);
    Wdp[2] = 0.0;
    Mc[0] = 0.0; Mc[1] = 0.0; Mc[2] = 1.0;
    Wdv[0] = (-tr.cy * tr.sz) * v0 + (-tr.cz) * v1;
    Wdv[1] = (tr.cy * tr.cz) * v0 + (-tr.sz) * v1;
    Wdv[2] = 0.0;
  }
  double dR[9];
  RotationMatrixDeriv(tr, c, dR);
  double w[3], wd[3], Iww[3], v11[3], v21[3];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    w[d] = TB_S(kDpW + d); wd[d] = TB_S(kDpWd + d); Iww[d] = TB_S(kDpIww + d);
    v11[d] = TB_S(kDpV11 + d); v21[d] = TB_S(kDpV21 + d);
  }
  double t1[3], u1[3], q1[3], q2[3];
  MatVec3(dR, v11, t1);                                // dR I_b R^T wd
  MatVec3(dR, v21, u1);                                // dR I_b R^T w
  MatTVec3(dR, wd, q1);                                // dR^T wd
  MatTVec3(dR, w, q2);                                 // dR^T w
  double t2[3], u2[3], t3[3], u3[3], IwM[3], a1v[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {                        // rows of R I_b and I_w come from shared memory on demand
    const double ri0 = TB_S(kDpRI + 3 * i), ri1 = TB_S(kDpRI + 3 * i + 1), ri2 = TB_S(kDpRI + 3 * i + 2);
    const double iw0 = TB_S(kDpIw + 3 * i), iw1 = TB_S(kDpIw + 3 * i + 1), iw2 = TB_S(kDpIw + 3 * i + 2);
    t2[i] = (ri0 * q1[0] + ri1 * q1[1]) + ri2 * q1[2];     // R I_b dR^T wd
    u2[i] = (ri0 * q2[0] + ri1 * q2[1]) + ri2 * q2[2];     // R I_b dR^T w
    t3[i] = (iw0 * Wdp[0] + iw1 * Wdp[1]) + iw2 * Wdp[2];  // I_w d(wd)
    u3[i] = (iw0 * Wp[0] + iw1 * Wp[1]) + iw2 * Wp[2];     // I_w d(w)
    IwM[i] = (iw0 * Mc[0] + iw1 * Mc[1]) + iw2 * Mc[2];
    a1v[i] = (iw0 * Wdv[0] + iw1 * Wdv[1]) + iw2 * Wdv[2];
  }
  double inner[3] = {u1[0] + u2[0] + u3[0], u1[1] + u2[1] + u3[1], u1[2] + u2[2] + u3[2]};
  double c1[3], c2[3], b1[3], b2[3];
  Cross3(w, inner, c1);                                // w x d(I_w w)
  Cross3(Iww, Wp, c2);                                 // (I_w w) x dw
  Cross3(w, IwM, b1);
  Cross3(Iww, Mc, b2);
  const double* Ba = cx.T.B + it.b0 + 4;               // pos[4], vel[4], acc[4] basis by sd
  double* G = cx.G + it.gslot * L::GS;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double Ap = (t1[r] + t2[r] + t3[r]) + (c1[r] - c2[r]);
    const double Av = a1v[r] + (b1[r] - b2[r]);
    const double Aa = IwM[r];
#pragma unroll
    for (int sd = 0; sd < 4; ++sd) TB_G(kDynJang + r * 12 + sd * 3 + c) = (Ap * Ba[sd] + Av * Ba[4 + sd]) + Aa * Ba[8 + sd];
  }
#undef TB_S
}

// ------------------------------------------------------------------------------------------ range of motion
// level 0: base position, R and dR/d(euler) of a RoM sample
template <class L>
__device__ __forceinline__ void ItemRomPre(const ItemCtx<L>& cx, const ItemC& it) {
  Pt3 lin, ang;
  EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, lin);
  EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 1, ang);
  const Trig tr = MakeTrig(ang.p);
  double* G = cx.G + it.gslot * L::GS;
  double R[9];
  RotationMatrix(tr, R);
#pragma unroll
  for (int d = 0; d < 3; ++d) TB_G(kRpLin + d) = lin.p[d];
#pragma unroll
  for (int i = 0; i < 9; ++i) TB_G(kRpR + i) = R[i];
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    RotationMatrixDeriv(tr, c, R);
#pragma unroll
    for (int i = 0; i < 9; ++i) TB_G(kRpDR + 9 * c + i) = R[i];
  }
}

// level 1: g = R^T (p_e - c) (range_of_motion_constraint.cc:72-83) and D[r][c] = (dR_c^T (p_e - c))[r]
template <class L>
__device__ __forceinline__ void ItemRomFoot(const ItemCtx<L>& cx, const ItemC& it) {
  Pt3 p;
  RtPoint rp;
  if constexpr (L::RT) rp = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0, p);
  else EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, p);
  const double* S = cx.G + it.aux * L::GS;
  double* G = cx.G + it.gslot * L::GS;
  if constexpr (L::RT) {
    // Hermite weights of the ee-motion spline and the finished ee-schedule block R^T J_p,dT
    // (range_of_motion_constraint.cc:126-130): G + 9 = W_p[n_cols] | schedule[3][P-1]
    const RtSpline sm = cx.T.rt_splines[it.ee];
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, sm, rp, G + 9 * L::GS);
    double* Sc = G + (9 + sm.n_cols) * L::GS;
    const RtFoot ft = cx.T.rt_feet[it.ee];
    double dpT[3];
    const int cur = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rp, p.v, dpT);
    const int nc = ft.n_phases - 1;
    for (int j = 0; j < nc; ++j) {
      double c[3];
      DurationColumn(j, cur, ft.n_phases, dpT, p.v, c);
#pragma unroll
      for (int r = 0; r < 3; ++r)
        Sc[(r * nc + j) * L::GS] = (S[(kRpR + r) * L::GS] * c[0] + S[(kRpR + 3 + r) * L::GS] * c[1]) + S[(kRpR + 6 + r) * L::GS] * c[2];
    }
  }
  const double rw[3] = {p.p[0] - S[(kRpLin + 0) * L::GS], p.p[1] - S[(kRpLin + 1) * L::GS], p.p[2] - S[(kRpLin + 2) * L::GS]};
  if (cx.g) {
    double* gr = cx.g + it.row * L::RS;
#pragma unroll
    for (int r = 0; r < 3; ++r)
      gr[(r) * L::RS] = (S[(kRpR + r) * L::GS] * rw[0] + S[(kRpR + 3 + r) * L::GS] * rw[1]) + S[(kRpR + 6 + r) * L::GS] * rw[2];
  }
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const double* D = S + (kRpDR + 9 * c) * L::GS;
      TB_G(3 * r + c) = (D[r * L::GS] * rw[0] + D[(3 + r) * L::GS] * rw[1]) + D[(6 + r) * L::GS] * rw[2];
    }
}

// ---------------------------------------------------------------------------- curvature block (Gap), out of line
// Classes that declare the curvature block (ItemC::sub = TerrainCurvatureDims of the class, structure.cc); compiled
// only into the CURV instantiations of the kernel (Lockstep<P, RT, true>).
// f . d(b_i)/d(dim) for the five force rows (force_constraint_discretized.cc:166-176): b_i = n, t1 -/+ mu n, t2 -/+ mu n
__device__ __forceinline__ void ForceBasisScales(const TerrainSample& ts, int dim, const double f[3], double s[5]) {
  const double mu = kFrictionCoeff;
  double dn[3], dt1[3], dt2[3], v[15];
  TerrainBasisDeriv(ts, 0, dim, dn);
  TerrainBasisDeriv(ts, 1, dim, dt1);
  TerrainBasisDeriv(ts, 2, dim, dt2);
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    v[d] = dn[d];
    v[3 + d] = dt1[d] - mu * dn[d];
    v[6 + d] = dt1[d] + mu * dn[d];
    v[9 + d] = dt2[d] - mu * dn[d];
    v[12 + d] = dt2[d] + mu * dn[d];
  }
#pragma unroll
  for (int i = 0; i < 5; ++i) s[i] = (f[0] * v[3 * i] + f[1] * v[3 * i + 1]) + f[2] * v[3 * i + 2];
}

// ROWS = 5: force constraint (scales f . d b_i); ROWS = 4: torque constraint (tau . dt1, tau . dt2, +-tau . dn - k mu f . dn,
// torque_constraint_discretized.cc:176-197).  Writes S[slot][ROWS] behind the item's other slots; with optimised
// durations also the ee-motion weights W_p and "+ scale * J_p,dT(row dim)" onto the finished ee-schedule block.
template <class L, int ROWS>
__device__ __forceinline__ void ItemCurvature(const ItemCtx<L>& cx, const ItemC& it) {
  const int dims = it.sub, nd = CurvDimCount(dims), E = cx.n_ee;
  if (nd == 0) return;                                  // a class that only admits table terrains declares no block
  Pt3 p, f, tau;
  RtPoint rp;
  if constexpr (L::RT) {
    rp = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0, p);
    EvalSplineRt<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 1, f);
    if (ROWS == 4) EvalSplineRt<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 2, tau);
  } else {
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, p);
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 1, f);
    if (ROWS == 4) EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 2, tau);
  }
  const TerrainSample ts = TerrainAt(cx, p.p[0], p.p[1]);
  double* G = cx.G + it.gslot * L::GS;
  int n_w = 0, n_sched = 0;
  if constexpr (L::RT) {
    n_w = cx.T.rt_splines[E + it.ee].n_cols + (ROWS == 4 ? cx.T.rt_splines[2 * E + it.ee].n_cols : 0);
    n_sched = cx.T.rt_feet[it.ee].n_phases - 1;
  }
  double* S = G + (ROWS == 5 ? ForceCurvOffset(n_w, n_sched) : TorqueCurvOffset(n_w, n_sched)) * L::GS;
  double sc[2][5];
  int slot = 0;
  for (int dim = 0; dim < 2; ++dim) {
    if (!((dims >> dim) & 1)) continue;
    if (ROWS == 5) {
      ForceBasisScales(ts, dim, f.p, sc[slot]);
    } else {
      double dn[3], dt1[3], dt2[3];
      TerrainBasisDeriv(ts, 0, dim, dn);
      TerrainBasisDeriv(ts, 1, dim, dt1);
      TerrainBasisDeriv(ts, 2, dim, dt2);
      const double kf = cx.T.B[it.b0], mu = kFrictionCoeff;
      const double s_tx = (tau.p[0] * dt1[0] + tau.p[1] * dt1[1]) + tau.p[2] * dt1[2];
      const double s_ty = (tau.p[0] * dt2[0] + tau.p[1] * dt2[1]) + tau.p[2] * dt2[2];
      const double s_tau_n = (tau.p[0] * dn[0] + tau.p[1] * dn[1]) + tau.p[2] * dn[2];
      const double s_f_n = (f.p[0] * dn[0] + f.p[1] * dn[1]) + f.p[2] * dn[2];
      const double s_lim = (kf * mu) * s_f_n;
      sc[slot][0] = s_tx; sc[slot][1] = s_ty; sc[slot][2] = s_tau_n - s_lim; sc[slot][3] = -s_tau_n - s_lim;
    }
    for (int i = 0; i < ROWS; ++i) S[(ROWS * slot + i) * L::GS] = sc[slot][i];
    ++slot;
  }
  if constexpr (L::RT) {
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, cx.T.rt_splines[it.ee], rp, S + ROWS * nd * L::GS);
    double* Sc = G + (15 + n_w) * L::GS;                  // the item's finished ee-schedule block [ROWS][P-1]
    const RtFoot ft = cx.T.rt_feet[it.ee];
    double dpT[3];
    const int cur_p = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rp, p.v, dpT);
    const int nc = ft.n_phases - 1;
    for (int j = 0; j < nc; ++j) {
      double cp[3];
      DurationColumn(j, cur_p, ft.n_phases, dpT, p.v, cp);
      slot = 0;
      for (int dim = 0; dim < 2; ++dim) {
        if (!((dims >> dim) & 1)) continue;
        for (int i = 0; i < ROWS; ++i) Sc[(i * nc + j) * L::GS] = Sc[(i * nc + j) * L::GS] + sc[slot][i] * cp[dim];
        ++slot;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------- force (discretised)
// g: force_constraint_discretized.cc:97-117.  G slots: b_i[3], i = 0..4: n, t1 -/+ mu n, t2 -/+ mu n.
template <class L>
__device__ __forceinline__ void ItemForce(const ItemCtx<L>& cx, const ItemC& it) {
  Pt3 p, f;
  RtPoint rf;
  if constexpr (L::RT) {
    EvalSplineRt<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0, p);
    rf = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 1, f);
  } else {
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, p);
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 1, f);
  }
  const TerrainSample ts = TerrainAt(cx, p.p[0], p.p[1]);
  double n[3], t1[3], t2[3];
  TerrainBasis(ts, n, t1, t2);
  const double mu = kFrictionCoeff;
  double* G = cx.G + it.gslot * L::GS;
  double b[15];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    b[d] = n[d];
    b[3 + d] = t1[d] - mu * n[d];
    b[6 + d] = t1[d] + mu * n[d];
    b[9 + d] = t2[d] - mu * n[d];
    b[12 + d] = t2[d] + mu * n[d];
  }
#pragma unroll
  for (int i = 0; i < 15; ++i) TB_G(i) = b[i];
  if (cx.g) {
    double* gr = cx.g + it.row * L::RS;
#pragma unroll
    for (int i = 0; i < 5; ++i) gr[(i) * L::RS] = (f.p[0] * b[3 * i] + f.p[1] * b[3 * i + 1]) + f.p[2] * b[3 * i + 2];
  }
  if constexpr (L::RT) {
    // G + 15 = W_f[n_cols] | schedule[5][P-1] = b_i . J_f,dT (force_constraint_discretized.cc:190-220; the terrain-basis
    // part is dropped where the terrain has no curvature, :58)
    const int E = cx.n_ee;
    const RtSpline sf = cx.T.rt_splines[E + it.ee];
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, sf, rf, G + 15 * L::GS);
    double* Sc = G + (15 + sf.n_cols) * L::GS;
    const RtFoot ft = cx.T.rt_feet[it.ee];
    double dfT[3];
    const int cur = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rf, f.v, dfT);
    const int nc = ft.n_phases - 1;
    for (int j = 0; j < nc; ++j) {
      double c[3];
      DurationColumn(j, cur, ft.n_phases, dfT, f.v, c);
#pragma unroll
      for (int i = 0; i < 5; ++i) Sc[(i * nc + j) * L::GS] = (b[3 * i] * c[0] + b[3 * i + 1] * c[1]) + b[3 * i + 2] * c[2];
    }
  }
  if constexpr (L::CURV) ItemCurvature<L, 5>(cx, it);
}

// ------------------------------------------------------------------------------------------- force (node-based)
// ForceConstraint (force_constraint.cc:62-171): basis at the stance foothold (ee-motion node at the start of the force
// node's phase, slot ItemC::aux), force = the force node's position value (slot ItemC::s0).
// G slots: b_i[3] (ee-force block), then f . d b_i / d(x|y) at 15 + 2 i + dim (ee-motion block, always stored).
template <class L>
__device__ __forceinline__ void ItemForceNode(const ItemCtx<L>& cx, const ItemC& it) {
  const double* fn = cx.nodes + it.s0 * 6 * L::NS;
  const double* pn = cx.nodes + it.aux * 6 * L::NS;
  const double f[3] = {fn[0], fn[L::NS], fn[2 * L::NS]};
  const TerrainSample ts = TerrainAt(cx, pn[0], pn[L::NS]);
  double n[3], t1[3], t2[3];
  TerrainBasis(ts, n, t1, t2);
  const double mu = kFrictionCoeff;
  double* G = cx.G + it.gslot * L::GS;
  double b[15];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    b[d] = n[d];
    b[3 + d] = t1[d] - mu * n[d];
    b[6 + d] = t1[d] + mu * n[d];
    b[9 + d] = t2[d] - mu * n[d];
    b[12 + d] = t2[d] + mu * n[d];
  }
#pragma unroll
  for (int i = 0; i < 15; ++i) TB_G(i) = b[i];
#pragma unroll
  for (int dim = 0; dim < 2; ++dim) {
    double dn[3], dt1[3], dt2[3], v[15];
    TerrainBasisDeriv(ts, 0, dim, dn);
    TerrainBasisDeriv(ts, 1, dim, dt1);
    TerrainBasisDeriv(ts, 2, dim, dt2);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      v[d] = dn[d];
      v[3 + d] = dt1[d] - mu * dn[d];
      v[6 + d] = dt1[d] + mu * dn[d];
      v[9 + d] = dt2[d] - mu * dn[d];
      v[12 + d] = dt2[d] + mu * dn[d];
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) TB_G(15 + 2 * i + dim) = (f[0] * v[3 * i] + f[1] * v[3 * i + 1]) + f[2] * v[3 * i + 2];
  }
  if (cx.g) {
    double* gr = cx.g + it.row * L::RS;
#pragma unroll
    for (int i = 0; i < 5; ++i) gr[(i) * L::RS] = (f[0] * b[3 * i] + f[1] * b[3 * i + 1]) + f[2] * b[3 * i + 2];
  }
}

// ------------------------------------------------------------------------------------------ torque (discretised)
// g: torque_constraint_discretized.cc:105-129.  G slots: t1[3], t2[3], n[3], -n[3], (-k mu) n[3]; B[b0] = k_friction.
template <class L>
__device__ __forceinline__ void ItemTorque(const ItemCtx<L>& cx, const ItemC& it) {
  Pt3 p, f, tau;
  RtPoint rf, rt;
  if constexpr (L::RT) {
    EvalSplineRt<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0, p);
    rf = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 1, f);
    rt = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + 2, tau);
  } else {
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, p);
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 1, f);
    EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 2, tau);
  }
  const TerrainSample ts = TerrainAt(cx, p.p[0], p.p[1]);
  double n[3], t1[3], t2[3];
  TerrainBasis(ts, n, t1, t2);
  const double mu = kFrictionCoeff, kf = cx.T.B[it.b0];
  double* G = cx.G + it.gslot * L::GS;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    TB_G(d) = t1[d];
    TB_G(3 + d) = t2[d];
    TB_G(6 + d) = n[d];
    TB_G(9 + d) = -n[d];
    TB_G(12 + d) = (-kf * mu) * n[d];
  }
  if (cx.g) {
    auto dot = [](const double* a, const double* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; };
    const double tau_t1 = dot(tau.p, t1), tau_t2 = dot(tau.p, t2), tau_n = dot(tau.p, n), f_n = dot(f.p, n);
    const double tz_lim = kf * mu * f_n;
    double* gr = cx.g + it.row * L::RS;
    gr[0 * L::RS] = tau_t1;
    gr[1 * L::RS] = tau_t2;
    gr[2 * L::RS] = tau_n - tz_lim;
    gr[3 * L::RS] = -tau_n - tz_lim;
  }
  if constexpr (L::RT) {
    // G + 15 = W_tau | W_f | schedule[4][P-1] (torque_constraint_discretized.cc:222-234)
    const int E = cx.n_ee;
    const RtSpline sf = cx.T.rt_splines[E + it.ee], st = cx.T.rt_splines[2 * E + it.ee];
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, st, rt, G + 15 * L::GS);
    WriteHermiteWeights<L::GS, L::NS>(cx.T, cx.rtd, sf, rf, G + (15 + st.n_cols) * L::GS);
    double* Sc = G + (15 + st.n_cols + sf.n_cols) * L::GS;
    const RtFoot ft = cx.T.rt_feet[it.ee];
    double dtT[3], dfT[3];
    const int cur_t = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rt, tau.v, dtT);
    const int cur_f = PosWrtPhaseDuration<L::NS>(cx.T, cx.nodes, cx.rtd, rf, f.v, dfT);
    const int nc = ft.n_phases - 1;
    for (int j = 0; j < nc; ++j) {
      double ct[3], cf[3];
      DurationColumn(j, cur_t, ft.n_phases, dtT, tau.v, ct);
      DurationColumn(j, cur_f, ft.n_phases, dfT, f.v, cf);
      const double bf[3] = {(-kf * mu) * n[0], (-kf * mu) * n[1], (-kf * mu) * n[2]};
      Sc[(0 * nc + j) * L::GS] = (t1[0] * ct[0] + t1[1] * ct[1]) + t1[2] * ct[2];
      Sc[(1 * nc + j) * L::GS] = (t2[0] * ct[0] + t2[1] * ct[1]) + t2[2] * ct[2];
      const double a2 = (n[0] * ct[0] + n[1] * ct[1]) + n[2] * ct[2];
      const double a3 = ((-n[0]) * ct[0] + (-n[1]) * ct[1]) + (-n[2]) * ct[2];
      Sc[(2 * nc + j) * L::GS] = ((a2 + bf[0] * cf[0]) + bf[1] * cf[1]) + bf[2] * cf[2];
      Sc[(3 * nc + j) * L::GS] = ((a3 + bf[0] * cf[0]) + bf[1] * cf[1]) + bf[2] * cf[2];
    }
  }
  if constexpr (L::CURV) ItemCurvature<L, 4>(cx, it);
}

// ------------------------------------------------------------------------------------------------ base motion
// BaseMotionConstraint (base_motion_constraint.cc:60-66): rows AX..AZ = base euler angles, LX..LZ = base position;
// the Jacobian is constant (Hermite basis values)
template <class L>
__device__ __forceinline__ void ItemBaseRom(const ItemCtx<L>& cx, const ItemC& it) {
  if (!cx.g) return;
  Pt3 lin, ang;
  EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, lin);
  EvalSpline<kNeedPos, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + 1, ang);
  double* gr = cx.g + it.row * L::RS;
#pragma unroll
  for (int d = 0; d < 3; ++d) { gr[d * L::RS] = ang.p[d]; gr[(3 + d) * L::RS] = lin.p[d]; }
}

// ------------------------------------------------------------------------------------------------ node-based
// TerrainConstraint (terrain_constraint.cc:61-111) and BaseHeightConstraint (base_height_constraint.cc:58-110)
template <class L>
__device__ __forceinline__ void ItemHeightNode(const ItemCtx<L>& cx, const ItemC& it, double offset) {
  const double* n = cx.nodes + it.s0 * 6 * L::NS;
  const double px = n[0], py = n[L::NS], pz = n[2 * L::NS];
  const TerrainSample ts = TerrainAt(cx, px, py);
  double* G = cx.G + it.gslot * L::GS;
  TB_G(0) = -ts.hx;
  TB_G(1) = -ts.hy;
  if (cx.g) cx.g[it.row * L::RS] = (it.type == kItemBaseHeightNode) ? (pz - ts.h) - offset : pz - ts.h;
}

// TerrainConstraintHard (terrain_constraint_hard.cc:50-132), one row per (sample, foot):
//   g = (p_z - h(p)) - min(k |v_tan|, k),   v_tan = (v.t1, v.t2) with the normalised tangents, k = 0.02
//   d g / d(column of dimension d) = c_d J_pos - (k td_d) J_vel:  c_Z = 1, c_X = -h_x, c_Y = -h_y; the velocity term only
//   while |v_tan| > 1e-6 and k |v_tan| < 0.05 - 1e-6 (:112-128), td = (v.t1 t1 + v.t2 t2) / |v_tan|.
// The reference forms the row as J_pos.row(Z), then -= h_x J_pos.row(X), -= h_y J_pos.row(Y), then -= (k td_d) J_vel.row(d):
// a column of dimension X holds 0 - h_x J_pos = -(h_x J_pos), then minus the velocity product.
// Fixed durations: B[b0 + 3 j ..] = J_pos, J_vel, dimension of stored column j (sub columns) -> G[j].
// Optimised durations: every column of the set; pos / vel weight sums of the active polynomial in scratch, G[c] finished.
template <class L>
__device__ __forceinline__ void ItemTerrainHard(const ItemCtx<L>& cx, const ItemC& it) {
  constexpr double k_coeff = 0.02;
  Pt3 p;
  RtPoint rp;
  if constexpr (L::RT) rp = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0, p);
  else EvalSpline<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, it.s0, p);
  const TerrainSample ts = TerrainAt(cx, p.p[0], p.p[1]);
  double n[3], t1[3], t2[3];
  TerrainBasis(ts, n, t1, t2);
  const double vt1 = (p.v[0] * t1[0] + p.v[1] * t1[1]) + p.v[2] * t1[2];
  const double vt2 = (p.v[0] * t2[0] + p.v[1] * t2[1]) + p.v[2] * t2[2];
  const double mag = sqrt(vt1 * vt1 + vt2 * vt2);
  if (cx.g) {
    const double delta_z = p.p[2] - ts.h;
    const double a = k_coeff * mag;
    cx.g[it.row * L::RS] = delta_z - (a < k_coeff ? a : k_coeff);       // std::min(k |v_tan|, k)
  }
  double kv0 = 0.0, kv1 = 0.0, kv2 = 0.0;
  bool vel_term = false;
  if (mag > 1e-6 && k_coeff * mag < 0.05 - 1e-6) {
    vel_term = true;
    kv0 = k_coeff * ((vt1 * t1[0] + vt2 * t2[0]) / mag);
    kv1 = k_coeff * ((vt1 * t1[1] + vt2 * t2[1]) / mag);
    kv2 = k_coeff * ((vt1 * t1[2] + vt2 * t2[2]) / mag);
  }
  auto finish = [&](int d, double jp, double jv) {       // selects, not indexed arrays: no local memory
    double v = d == 2 ? jp : -((d == 0 ? ts.hx : ts.hy) * jp);
    if (vel_term) v = v - (d == 0 ? kv0 : d == 1 ? kv1 : kv2) * jv;
    return v;
  };
  double* G = cx.G + it.gslot * L::GS;
  if constexpr (L::RT) {
    const RtSpline sm = cx.T.rt_splines[it.ee];
    double* Wp = G + sm.n_cols * L::GS;
    double* Wv = Wp + sm.n_cols * L::GS;
    if (!TB_PREZERO_G)
      for (int c = 0; c < sm.n_cols; ++c) { Wp[c * L::GS] = 0.0; Wv[c * L::GS] = 0.0; }
    const double* q = cx.rtd + rp.poly * kRtd * L::NS;
    const double t = rp.t, T = rp.T, T2 = q[kRtT2 * L::NS], T3 = q[kRtT3 * L::NS];
    const double r1 = q[kRtR1 * L::NS], r2 = q[kRtR2 * L::NS], r3 = q[kRtR3 * L::NS];
    const double tt2 = t * t, tt3 = Cube(t);
    double bp[4], bv[4];                                 // polynomial.cc:106-175, by side * 2 + node derivative
    bp[0] = ExactDiv(2 * tt3, T3, r3) - ExactDiv(3 * tt2, T2, r2) + 1;
    bp[1] = t - ExactDiv(2 * tt2, T, r1) + ExactDiv(tt3, T2, r2);
    bp[2] = ExactDiv(3 * tt2, T2, r2) - ExactDiv(2 * tt3, T3, r3);
    bp[3] = ExactDiv(tt3, T2, r2) - ExactDiv(tt2, T, r1);
    bv[0] = ExactDiv(6 * tt2, T3, r3) - ExactDiv(6 * t, T2, r2);
    bv[1] = ExactDiv(3 * tt2, T2, r2) - ExactDiv(4 * t, T, r1) + 1;
    bv[2] = ExactDiv(6 * t, T2, r2) - ExactDiv(6 * tt2, T3, r3);
    bv[3] = ExactDiv(3 * tt2, T2, r2) - ExactDiv(2 * t, T, r1);
    const int n0 = cx.T.polys[rp.poly] & 0xFFFF;
    unsigned touched = 0;                                // columns of the active polynomial (at most 12, by first hit)
    int col_of[12], dim_of[12];
#pragma unroll
    for (int side = 0; side < 2; ++side)
#pragma unroll
      for (int deriv = 0; deriv < 2; ++deriv)
#pragma unroll
        for (int d = 0; d < 3; ++d) {
          const unsigned src = cx.T.node_src[(n0 + side) * 6 + deriv * 3 + d];
          if (src != 0xFFFFu) {
            const int c = static_cast<int>(src) - sm.col0;
            Wp[c * L::GS] = Wp[c * L::GS] + bp[side * 2 + deriv];
            Wv[c * L::GS] = Wv[c * L::GS] + bv[side * 2 + deriv];
            col_of[touched] = c;
            dim_of[touched] = d;
            ++touched;
          }
        }
    if (!TB_PREZERO_G)
      for (int c = 0; c < sm.n_cols; ++c) G[c * L::GS] = 0.0;
    for (unsigned i = 0; i < touched; ++i) {
      const int c = col_of[i];
      G[c * L::GS] = finish(dim_of[i], Wp[c * L::GS], Wv[c * L::GS]);
    }
  } else {
    const double* Bc = cx.T.B + it.b0;
    for (int j = 0; j < it.sub; ++j)
      TB_G(j) = finish(static_cast<int>(Bc[3 * j + 2]), Bc[3 * j], Bc[3 * j + 1]);
  }
}

// EELinearConstraint (ee_linear_constraint.cc:20-48), one row per sample: g = sum over the terms of coeff * value of the
// term's foot spline in one dimension (position or velocity).  sub terms; samples s0 .. s0 + sub - 1 in term order;
// B[b0 + 4 j ..] = coeff, dimension, derivative, G offset of the foot's column block.
// Fixed durations: the Jacobian entries are class constants, the item only forms g.
// Optimised durations: G[gslot + offset + c] = sum of coeff * (run-time Hermite weight) over the terms that reach column c
// of the foot's set: the columns of the active polynomial's two nodes in the term's dimension; every other stored column
// stays 0 (`jac.row(k) += coeff * jac_nodes.row(dim)` on the pre-filled structure).
template <class L>
__device__ __forceinline__ void ItemEELinear(const ItemCtx<L>& cx, const ItemC& it) {
  const double* Bc = cx.T.B + it.b0;
  double* G = cx.G + it.gslot * L::GS;
  if constexpr (L::RT) {
    if (!TB_PREZERO_G)
      for (int j = 0; j < it.sub; ++j) {
        const RtSpline sm = cx.T.rt_splines[cx.T.samples[it.s0 + j] & 0xFFFF];
        double* Gf = G + static_cast<int>(Bc[4 * j + 3]) * L::GS;
        for (int c = 0; c < sm.n_cols; ++c) Gf[c * L::GS] = 0.0;
      }
  }
  double val = 0.0;
  for (int j = 0; j < it.sub; ++j) {
    const double coeff = Bc[4 * j];
    const int dim = static_cast<int>(Bc[4 * j + 1]);
    const bool vel = Bc[4 * j + 2] != 0.0;
    Pt3 p;
    if constexpr (L::RT) {
      const RtPoint rp = EvalSplineRt<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, cx.rtd, it.s0 + j, p);
      const RtSpline sm = cx.T.rt_splines[cx.T.samples[it.s0 + j] & 0xFFFF];
      double* Gf = G + static_cast<int>(Bc[4 * j + 3]) * L::GS;
      const double* q = cx.rtd + rp.poly * kRtd * L::NS;
      const double t = rp.t, T = rp.T, T2 = q[kRtT2 * L::NS], T3 = q[kRtT3 * L::NS];
      const double r1 = q[kRtR1 * L::NS], r2 = q[kRtR2 * L::NS], r3 = q[kRtR3 * L::NS];
      const double tt2 = t * t, tt3 = Cube(t);
      double w[4];                                       // polynomial.cc:106-175, by side * 2 + node derivative
      if (!vel) {
        w[0] = ExactDiv(2 * tt3, T3, r3) - ExactDiv(3 * tt2, T2, r2) + 1;
        w[1] = t - ExactDiv(2 * tt2, T, r1) + ExactDiv(tt3, T2, r2);
        w[2] = ExactDiv(3 * tt2, T2, r2) - ExactDiv(2 * tt3, T3, r3);
        w[3] = ExactDiv(tt3, T2, r2) - ExactDiv(tt2, T, r1);
      } else {
        w[0] = ExactDiv(6 * tt2, T3, r3) - ExactDiv(6 * t, T2, r2);
        w[1] = ExactDiv(3 * tt2, T2, r2) - ExactDiv(4 * t, T, r1) + 1;
        w[2] = ExactDiv(6 * t, T2, r2) - ExactDiv(6 * tt2, T3, r3);
        w[3] = ExactDiv(3 * tt2, T2, r2) - ExactDiv(2 * t, T, r1);
      }
      const int n0 = cx.T.polys[rp.poly] & 0xFFFF;
      // one value per column first (a stance variable writes both nodes of its polynomial: node_spline.cc:84-112 sums the
      // two weights into the Jacobian entry), then coeff * that value joins the row
      int cols[4];
      double sum[4];
      int n = 0;
#pragma unroll
      for (int sd = 0; sd < 4; ++sd) {
        const unsigned src = cx.T.node_src[(n0 + (sd >> 1)) * 6 + (sd & 1) * 3 + dim];
        if (src == 0xFFFFu) continue;
        const int c = static_cast<int>(src) - sm.col0;
        int hit = -1;
        for (int i = 0; i < n; ++i) hit = cols[i] == c ? i : hit;
        if (hit < 0) { cols[n] = c; sum[n] = 0.0 + w[sd]; ++n; }
        else sum[hit] = sum[hit] + w[sd];
      }
      for (int i = 0; i < n; ++i) Gf[cols[i] * L::GS] = Gf[cols[i] * L::GS] + coeff * sum[i];
    } else {
      EvalSpline<kNeedPos | kNeedVel, L::NS>(cx.T, cx.nodes, cx.coef, it.s0 + j, p);
    }
    const double pv = dim == 0 ? p.p[0] : dim == 1 ? p.p[1] : p.p[2];
    const double vv = dim == 0 ? p.v[0] : dim == 1 ? p.v[1] : p.v[2];
    val += coeff * (vel ? vv : pv);
  }
  if (cx.g) cx.g[it.row * L::RS] = val;
}

// Node-based TorqueConstraint (torque_constraint.cc:64-157): s0 = the stance torque node, aux = the ee-motion node at the
// start of its phase (the foothold), b0 = the torque node at the start of the phase.  G: t1[3], t2[3], n[3], then
// (tau_s . d b / d x, tau_s . d b / d y) for b = t1, t2, n with tau_s the start-of-phase torque (:141).
template <class L>
__device__ __forceinline__ void ItemTorqueNode(const ItemCtx<L>& cx, const ItemC& it) {
  const double* tn = cx.nodes + it.s0 * 6 * L::NS;
  const double* pn = cx.nodes + it.aux * 6 * L::NS;
  const double* sn = cx.nodes + it.b0 * 6 * L::NS;
  const double tau[3] = {tn[0], tn[L::NS], tn[2 * L::NS]};
  const double tau_s[3] = {sn[0], sn[L::NS], sn[2 * L::NS]};
  const TerrainSample ts = TerrainAt(cx, pn[0], pn[L::NS]);
  double n[3], t1[3], t2[3];
  TerrainBasis(ts, n, t1, t2);
  double* G = cx.G + it.gslot * L::GS;
#pragma unroll
  for (int d = 0; d < 3; ++d) { TB_G(d) = t1[d]; TB_G(3 + d) = t2[d]; TB_G(6 + d) = n[d]; }
#pragma unroll
  for (int dim = 0; dim < 2; ++dim) {
    double dn[3], dt1[3], dt2[3];
    TerrainBasisDeriv(ts, 0, dim, dn);
    TerrainBasisDeriv(ts, 1, dim, dt1);
    TerrainBasisDeriv(ts, 2, dim, dt2);
    TB_G(9 + 0 + dim) = (tau_s[0] * dt1[0] + tau_s[1] * dt1[1]) + tau_s[2] * dt1[2];
    TB_G(9 + 2 + dim) = (tau_s[0] * dt2[0] + tau_s[1] * dt2[1]) + tau_s[2] * dt2[2];
    TB_G(9 + 4 + dim) = (tau_s[0] * dn[0] + tau_s[1] * dn[1]) + tau_s[2] * dn[2];
  }
  if (cx.g) {
    double* gr = cx.g + it.row * L::RS;
    gr[0] = (tau[0] * t1[0] + tau[1] * t1[1]) + tau[2] * t1[2];
    gr[L::RS] = (tau[0] * t2[0] + tau[1] * t2[1]) + tau[2] * t2[2];
    gr[2 * L::RS] = (tau[0] * n[0] + tau[1] * n[1]) + tau[2] * n[2];
  }
}

// SwingConstraint (swing_constraint.cc:54-78)
template <class L>
__device__ __forceinline__ void ItemSwingNode(const ItemCtx<L>& cx, const ItemC& it) {
  if (!cx.g) return;
  const double* cur = cx.nodes + it.s0 * 6 * L::NS;
  const double* prev = cur - 6 * L::NS;
  const double* next = cur + 6 * L::NS;
  double* gr = cx.g + it.row * L::RS;
#pragma unroll
  for (int d = 0; d < 2; ++d) {
    const double distance = next[d * L::NS] - prev[d * L::NS];
    const double center = prev[d * L::NS] + 0.5 * distance;
    const double des_vel = distance / 0.3;
    gr[(2 * d) * L::RS] = cur[d * L::NS] - center;
    gr[(2 * d + 1) * L::RS] = cur[(3 + d) * L::NS] - des_vel;
  }
}

// SplineAccConstraint (spline_acc_constraint.cc:48-64): acc at the end of poly j minus acc at the start of poly j+1
template <class L>
__device__ __forceinline__ void ItemSplineAcc(const ItemCtx<L>& cx, const ItemC& it) {
  if (!cx.g) return;
  const int j = it.s0;
  const double T = cx.T.durs[3 * (cx.T.polys[j] >> 16)];
  double* gr = cx.g + it.row * L::RS;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double prev = 2.0 * cx.coef[(j * 6 + d) * L::NS] + (6 * T) * cx.coef[(j * 6 + 3 + d) * L::NS];
    const double next = 2.0 * cx.coef[((j + 1) * 6 + d) * L::NS] + (6 * 0.0) * cx.coef[((j + 1) * 6 + 3 + d) * L::NS];
    gr[(d) * L::RS] = prev - next;
  }
}

// TotalDurationConstraint (total_duration_constraint.cc:48-54): sum of the P-1 optimised durations of one foot
template <class L>
__device__ __forceinline__ void ItemTotalDuration(const ItemCtx<L>& cx, const ItemC& it) {
  if (!cx.g) return;
  if constexpr (L::RT) {
    const RtFoot ft = cx.T.rt_feet[it.ee];
    double s = 0.0;
    for (int i = 0; i < ft.n_phases - 1; ++i) s += cx.pdur[(ft.phase0 + i) * L::NS];
    cx.g[it.row * L::RS] = s;
  }
}

template <class L>
__device__ __forceinline__ void RunItem(const ItemCtx<L>& cx, const ItemC& it) {
  switch (it.type) {
    case kItemDynPre: ItemDynPre<L>(cx, it); break;
    case kItemDynFoot: ItemDynFoot<L>(cx, it); break;
    case kItemDynWrench: ItemDynWrench<L>(cx, it); break;
    case kItemDynJlin: ItemDynJlin<L>(cx, it); break;
    case kItemDynCol: ItemDynCol<L>(cx, it); break;
    case kItemRomPre: ItemRomPre<L>(cx, it); break;
    case kItemRomFoot: ItemRomFoot<L>(cx, it); break;
    case kItemForce: ItemForce<L>(cx, it); break;
    case kItemTerrainHard:                               // classes with this set run the CURV instantiations (kClassTerrainHard)
      if constexpr (L::CURV) ItemTerrainHard<L>(cx, it);
      break;
    case kItemTorqueNode:
      if constexpr (L::CURV) ItemTorqueNode<L>(cx, it);
      break;
    case kItemEELinear:
      if constexpr (L::CURV) ItemEELinear<L>(cx, it);
      break;
    case kItemTorque: ItemTorque<L>(cx, it); break;
    case kItemForceNode: ItemForceNode<L>(cx, it); break;
    case kItemBaseRom: ItemBaseRom<L>(cx, it); break;
    case kItemTerrainNode: ItemHeightNode<L>(cx, it, 0.0); break;
    case kItemBaseHeightNode: ItemHeightNode<L>(cx, it, 0.4); break;   // nlp_formulation.cc:597
    case kItemSwingNode: ItemSwingNode<L>(cx, it); break;
    case kItemSplineAcc: ItemSplineAcc<L>(cx, it); break;
    case kItemTotalDuration: ItemTotalDuration<L>(cx, it); break;
    default: break;
  }
}

#undef TB_G

}  // namespace tb

// items.cuh — stage A: one thread evaluates one item (a time sample or a node of one constraint family),
// writes the item's constraint rows into g and its partial-derivative slots into the shared G buffer.
#pragma once
#include "device_math.cuh"

namespace tb {

struct ItemCtx {
  Tables T;
  const double* nodes;
  const double* coef;
  double* G;          // segment-relative G buffer (shared memory)
  double* g;          // this problem's constraint values (global), or nullptr
  const ProblemConst* pc;
};

// ------------------------------------------------------------------------------------------------ dynamic
// Angular prelude shared by the four sub-items of a sample: euler point, trig, R, M, Mdot, omega, omegadot, I_w.
struct AngState {
  Pt3 ang;
  Trig tr;
  double R[9], RI[9], Iw[9];
  double M00, M01, M10, M11, M20, Md00, Md01, Md10, Md11, Md20;
  double w[3], wd[3], Iw_w[3];
};

__device__ __forceinline__ void AngPrelude(const ItemCtx& cx, const ItemC& it, AngState& a) {
  EvalSpline<kNeedPos | kNeedVel | kNeedAcc>(cx.T, cx.nodes, cx.coef, it.s0 + 1, a.ang);
  a.tr = MakeTrig(a.ang.p);
  const Trig& tr = a.tr;
  RotationMatrix(tr, a.R);
  const double yd = a.ang.v[1], zd = a.ang.v[2];
  // M and Mdot (euler_converter.cc:133-166); only columns X and Y of Mdot and X,Y,(2,Z)=1 of M are non-zero
  a.M00 = tr.cy * tr.cz; a.M01 = -tr.sz; a.M10 = tr.cy * tr.sz; a.M11 = tr.cz; a.M20 = -tr.sy;
  a.Md00 = -tr.cz * tr.sy * yd - tr.cy * tr.sz * zd; a.Md01 = -tr.cz * zd;
  a.Md10 = tr.cy * tr.cz * zd - tr.sy * tr.sz * yd;  a.Md11 = -tr.sz * zd;
  a.Md20 = -tr.cy * yd;
  const double* v = a.ang.v;
  const double* ac = a.ang.a;
  a.w[0] = a.M00 * v[0] + a.M01 * v[1];
  a.w[1] = a.M10 * v[0] + a.M11 * v[1];
  a.w[2] = a.M20 * v[0] + 1.0 * v[2];
  a.wd[0] = (a.Md00 * v[0] + a.Md01 * v[1]) + (a.M00 * ac[0] + a.M01 * ac[1]);
  a.wd[1] = (a.Md10 * v[0] + a.Md11 * v[1]) + (a.M10 * ac[0] + a.M11 * ac[1]);
  a.wd[2] = (a.Md20 * v[0]) + (a.M20 * ac[0] + 1.0 * ac[2]);
  // I_w = R I_b R^T, (R I_b) first, inner index ascending
  const double* Ib = cx.pc->I;
  const double* R = a.R;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) a.RI[3 * i + j] = (R[3 * i] * Ib[j] + R[3 * i + 1] * Ib[3 + j]) + R[3 * i + 2] * Ib[6 + j];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      a.Iw[3 * i + j] = (a.RI[3 * i] * R[3 * j] + a.RI[3 * i + 1] * R[3 * j + 1]) + a.RI[3 * i + 2] * R[3 * j + 2];
  MatVec3(a.Iw, a.w, a.Iw_w);
}

// sub-item 0: g rows (dynamic_constraint.cc:63-68,128-148 + single_rigid_body_dynamics.cc:76-102, same operation
// order), f_e and (c - p_e) slots, and the finished base-lin angular block -sum_e Cross(f_e) J_pos.
__device__ __noinline__ void ItemDynamicWrench(const ItemCtx& cx, const ItemC& it) {
  const int E = it.n_ee;
  AngState a;
  AngPrelude(cx, it, a);
  Pt3 lin;
  EvalSpline<kNeedPos | kNeedAcc>(cx.T, cx.nodes, cx.coef, it.s0, lin);
  double Iw_wd[3];
  MatVec3(a.Iw, a.wd, Iw_wd);
  double* G = cx.G + it.gslot;
  double* Jlin = G + 36;
  double* Gf = G + 72;
  double* Gr = G + 72 + 3 * E;
  double f_sum[3] = {0.0, 0.0, 0.0}, tau_sum[3] = {0.0, 0.0, 0.0};
  for (int e = 0; e < E; ++e) {
    Pt3 f, p, tq;
    EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0 + 2 + 3 * e, p);
    EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0 + 3 + 3 * e, f);
    EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0 + 4 + 3 * e, tq);
    double r[3] = {lin.p[0] - p.p[0], lin.p[1] - p.p[1], lin.p[2] - p.p[2]};
    double cr[3];
    Cross3(f.p, r, cr);
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      tau_sum[d] = tau_sum[d] + (cr[d] + tq.p[d]);
      f_sum[d] = f_sum[d] + f.p[d];
      Gf[3 * e + d] = f.p[d];
      Gr[3 * e + d] = r[d];
    }
  }
  if (cx.g) {
    const double* w = a.w;
    double gy[3];   // Cross(w) * (I_w w): off-diagonal entries in column order
    gy[0] = ((-w[2]) * a.Iw_w[1]) + w[1] * a.Iw_w[2];
    gy[1] = (w[2] * a.Iw_w[0]) + (-w[0]) * a.Iw_w[2];
    gy[2] = ((-w[1]) * a.Iw_w[0]) + w[0] * a.Iw_w[1];
    const double m = cx.pc->mass;
    const double grav = 9.80665;
    double* gr = cx.g + it.row;
#pragma unroll
    for (int d = 0; d < 3; ++d) gr[d] = (Iw_wd[d] + gy[d]) - tau_sum[d];
    gr[3] = (m * lin.a[0] - f_sum[0]) - 0.0;
    gr[4] = (m * lin.a[1] - f_sum[1]) - 0.0;
    gr[5] = (m * lin.a[2] - f_sum[2]) - (-m * grav);
  }
  // base-lin angular block: entry(r, (sd,d)) = -(sum_e Cross(f_e)(r,d) * Bpos[sd]), feet summed in order
  const double* Bl = cx.T.B + it.b0;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int d = 0; d < 3; ++d) {
      if (d == r) {            // no entry uses these slots; keep them finite for the status check
#pragma unroll
        for (int sd = 0; sd < 4; ++sd) Jlin[r * 12 + sd * 3 + d] = 0.0;
        continue;
      }
      const int k = 3 - r - d;
      const double sg = ((d - r + 3) % 3 == 2) ? 1.0 : -1.0;
#pragma unroll
      for (int sd = 0; sd < 4; ++sd) {
        const double b = Bl[sd];
        double acc = (sg * Gf[k]) * b;
        for (int e = 1; e < E; ++e) acc = acc + (sg * Gf[3 * e + k]) * b;
        Jlin[r * 12 + sd * 3 + d] = -acc;
      }
    }
}

// sub-items 1..3: euler column c of d(angular rows)/d(base-ang variables)
//   value(r, (sd,c)) = A_pos[r][c]*Bp[sd] + A_vel[r][c]*Bv[sd] + A_acc[r][c]*Ba[sd]
// with A_* the partial derivatives of I_w wd + w x (I_w w) w.r.t. the euler spline's pos, vel, acc
// (single_rigid_body_dynamics.cc:124-166 through euler_converter.cc:85-131,223-304, regrouped).
__device__ __noinline__ void ItemDynamicEulerColumn(const ItemCtx& cx, const ItemC& it, int c) {
  AngState a;
  AngPrelude(cx, it, a);
  const Trig& tr = a.tr;
  const double* v = a.ang.v;
  const double* ac = a.ang.a;
  const double yd = v[1], zd = v[2];
  double v11[3], v21[3], tmp[3];
  MatTVec3(a.R, a.wd, tmp);  MatVec3(cx.pc->I, tmp, v11);     // I_b R^T wd
  MatTVec3(a.R, a.w, tmp);   MatVec3(cx.pc->I, tmp, v21);     // I_b R^T w
  // d(omega)/d(theta_c) = dM_c thetadot ; d(omegadot)/d(theta_c) = dMdot_c thetadot + dM_c thetaddot
  double Wp[3], Wdp[3], Mc[3], Wdv[3];
  if (c == 0) {
    Wp[0] = Wp[1] = Wp[2] = 0.0;
    Wdp[0] = Wdp[1] = Wdp[2] = 0.0;
    Mc[0] = a.M00; Mc[1] = a.M10; Mc[2] = a.M20;
    Wdv[0] = a.Md00; Wdv[1] = a.Md10; Wdv[2] = a.Md20;
  } else if (c == 1) {
    const double dM00 = -tr.sy * tr.cz, dM10 = -tr.sy * tr.sz, dM20 = -tr.cy;
    Wp[0] = dM00 * v[0]; Wp[1] = dM10 * v[0]; Wp[2] = dM20 * v[0];
    const double dMd00 = -tr.cy * tr.cz * yd + tr.sy * tr.sz * zd;
    const double dMd10 = -tr.cy * tr.sz * yd - tr.cz * tr.sy * zd;
    const double dMd20 = tr.sy * yd;
    Wdp[0] = dMd00 * v[0] + dM00 * ac[0];
    Wdp[1] = dMd10 * v[0] + dM10 * ac[0];
    Wdp[2] = dMd20 * v[0] + dM20 * ac[0];
    Mc[0] = a.M01; Mc[1] = a.M11; Mc[2] = 0.0;
    Wdv[0] = (-tr.cz * tr.sy) * v[0] + a.Md01;
    Wdv[1] = (-tr.sy * tr.sz) * v[0] + a.Md11;
    Wdv[2] = (-tr.cy) * v[0];
  } else {
    const double dM00 = -tr.cy * tr.sz, dM01 = -tr.cz, dM10 = tr.cy * tr.cz, dM11 = -tr.sz;
    Wp[0] = dM00 * v[0] + dM01 * v[1];
    Wp[1] = dM10 * v[0] + dM11 * v[1];
    Wp[2] = 0.0;
    const double dMd00 = tr.sy * tr.sz * yd - tr.cy * tr.cz * zd, dMd01 = tr.sz * zd;
    const double dMd10 = -tr.cz * tr.sy * yd - tr.cy * tr.sz * zd, dMd11 = -tr.cz * zd;
    Wdp[0] = (dMd00 * v[0] + dMd01 * v[1]) + (dM00 * ac[0] + dM01 * ac[1]);
    Wdp[1] = (dMd10 * v[0] + dMd11 * v[1]) + (dM10 * ac[0] + dM11 * ac[1]);
    Wdp[2] = 0.0;
    Mc[0] = 0.0; Mc[1] = 0.0; Mc[2] = 1.0;
    Wdv[0] = (-tr.cy * tr.sz) * v[0] + (-tr.cz) * v[1];
    Wdv[1] = (tr.cy * tr.cz) * v[0] + (-tr.sz) * v[1];
    Wdv[2] = 0.0;
  }
  double dR[9];
  RotationMatrixDeriv(tr, c, dR);
  double t1[3], t2[3], t3[3], u1[3], u2[3], u3[3], q[3];
  MatVec3(dR, v11, t1);                                // dR I_b R^T wd
  MatTVec3(dR, a.wd, q);  MatVec3(a.RI, q, t2);        // R I_b dR^T wd
  MatVec3(a.Iw, Wdp, t3);                              // I_w d(wd)
  MatVec3(dR, v21, u1);
  MatTVec3(dR, a.w, q);   MatVec3(a.RI, q, u2);
  MatVec3(a.Iw, Wp, u3);
  double inner[3] = {u1[0] + u2[0] + u3[0], u1[1] + u2[1] + u3[1], u1[2] + u2[2] + u3[2]};
  double c1[3], c2[3], IwM[3], a1[3], b1[3], b2[3];
  Cross3(a.w, inner, c1);                              // w x d(I_w w)
  Cross3(a.Iw_w, Wp, c2);                              // (I_w w) x dw
  MatVec3(a.Iw, Mc, IwM);
  MatVec3(a.Iw, Wdv, a1);
  Cross3(a.w, IwM, b1);
  Cross3(a.Iw_w, Mc, b2);
  const double* Ba = cx.T.B + it.b0 + 4;                 // pos[4], vel[4], acc[4] basis by sd
  double* Jang = cx.G + it.gslot;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double Ap = (t1[r] + t2[r] + t3[r]) + (c1[r] - c2[r]);
    const double Av = a1[r] + (b1[r] - b2[r]);
    const double Aa = IwM[r];
#pragma unroll
    for (int sd = 0; sd < 4; ++sd) Jang[r * 12 + sd * 3 + c] = (Ap * Ba[sd] + Av * Ba[4 + sd]) + Aa * Ba[8 + sd];
  }
}

// ------------------------------------------------------------------------------------------ range of motion
// g: range_of_motion_constraint.cc:72-83.  G slots: R^T[9], then per foot d(R^T r_e)/d(euler)[9].
__device__ __noinline__ void ItemRom(const ItemCtx& cx, const ItemC& it) {
  Pt3 lin, ang;
  EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0, lin);
  EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0 + 1, ang);
  const Trig tr = MakeTrig(ang.p);
  double R[9], dRx[9], dRy[9], dRz[9];
  RotationMatrix(tr, R);
  RotationMatrixDeriv(tr, 0, dRx);
  RotationMatrixDeriv(tr, 1, dRy);
  RotationMatrixDeriv(tr, 2, dRz);
  double* G = cx.G + it.gslot;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int d = 0; d < 3; ++d) G[3 * r + d] = R[3 * d + r];
  for (int e = 0; e < it.n_ee; ++e) {
    Pt3 p;
    EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0 + 2 + e, p);
    const double rw[3] = {p.p[0] - lin.p[0], p.p[1] - lin.p[1], p.p[2] - lin.p[2]};
    if (cx.g) {
      double* gr = cx.g + it.row + e * it.aux;
#pragma unroll
      for (int r = 0; r < 3; ++r) gr[r] = (R[r] * rw[0] + R[3 + r] * rw[1]) + R[6 + r] * rw[2];
    }
    double* D = G + 9 + 9 * e;
    double dx[3], dy[3], dz[3];
    MatTVec3(dRx, rw, dx);
    MatTVec3(dRy, rw, dy);
    MatTVec3(dRz, rw, dz);
#pragma unroll
    for (int r = 0; r < 3; ++r) { D[3 * r] = dx[r]; D[3 * r + 1] = dy[r]; D[3 * r + 2] = dz[r]; }
  }
}

// ------------------------------------------------------------------------------------------- force (discretised)
// g: force_constraint_discretized.cc:97-117.  G slots: b_i[3], i = 0..4: n, t1 -/+ mu n, t2 -/+ mu n.
__device__ __noinline__ void ItemForce(const ItemCtx& cx, const ItemC& it) {
  Pt3 p, f;
  EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0, p);
  EvalSpline<kNeedPos>(cx.T, cx.nodes, cx.coef, it.s0 + 1, f);
  const TerrainSample ts = EvalTerrain(cx.pc->terrain_id, p.p[0], p.p[1]);
  double n[3], t1[3], t2[3];
  TerrainBasis(ts, n, t1, t2);
  const double mu = kFrictionCoeff;
  double* G = cx.G + it.gslot;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    G[d] = n[d];
    G[3 + d] = t1[d] - mu * n[d];
    G[6 + d] = t1[d] + mu * n[d];
    G[9 + d] = t2[d] - mu * n[d];
    G[12 + d] = t2[d] + mu * n[d];
  }
  if (cx.g) {
    double* gr = cx.g + it.row;
#pragma unroll
    for (int i = 0; i < 5; ++i) gr[i] = (f.p[0] * G[3 * i] + f.p[1] * G[3 * i + 1]) + f.p[2] * G[3 * i + 2];
  }
}

// ------------------------------------------------------------------------------------------------ node-based
// TerrainConstraint (terrain_constraint.cc:61-111) and BaseHeightConstraint (base_height_constraint.cc:58-110)
__device__ __forceinline__ void ItemHeightNode(const ItemCtx& cx, const ItemC& it, double offset) {
  const double* n = cx.nodes + it.s0 * 6;
  const TerrainSample ts = EvalTerrain(cx.pc->terrain_id, n[0], n[1]);
  double* G = cx.G + it.gslot;
  G[0] = -ts.hx;
  G[1] = -ts.hy;
  if (cx.g) cx.g[it.row] = (it.type == kItemBaseHeightNode) ? (n[2] - ts.h) - offset : n[2] - ts.h;
}

// SwingConstraint (swing_constraint.cc:54-78)
__device__ __forceinline__ void ItemSwingNode(const ItemCtx& cx, const ItemC& it) {
  if (!cx.g) return;
  const double* cur = cx.nodes + it.s0 * 6;
  const double* prev = cur - 6;
  const double* next = cur + 6;
  double* gr = cx.g + it.row;
#pragma unroll
  for (int d = 0; d < 2; ++d) {
    const double distance = next[d] - prev[d];
    const double center = prev[d] + 0.5 * distance;
    const double des_vel = distance / 0.3;
    gr[2 * d] = cur[d] - center;
    gr[2 * d + 1] = cur[3 + d] - des_vel;
  }
}

// SplineAccConstraint (spline_acc_constraint.cc:48-64): acc at the end of poly j minus acc at the start of poly j+1
__device__ __forceinline__ void ItemSplineAcc(const ItemCtx& cx, const ItemC& it) {
  if (!cx.g) return;
  const int j = it.s0;
  const double T = cx.T.durs[3 * (cx.T.polys[j] >> 16)];
  double* gr = cx.g + it.row;
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double prev = 2.0 * cx.coef[j * 6 + d] + (6 * T) * cx.coef[j * 6 + 3 + d];
    const double next = 2.0 * cx.coef[(j + 1) * 6 + d] + (6 * 0.0) * cx.coef[(j + 1) * 6 + 3 + d];
    gr[d] = prev - next;
  }
}

__device__ __forceinline__ void RunItem(const ItemCtx& cx, const ItemC& it) {
  switch (it.type) {
    case kItemDynamic:
      if (it.sub == 0) ItemDynamicWrench(cx, it);
      else ItemDynamicEulerColumn(cx, it, it.sub - 1);
      break;
    case kItemRom: ItemRom(cx, it); break;
    case kItemForce: ItemForce(cx, it); break;
    case kItemTerrainNode: ItemHeightNode(cx, it, 0.0); break;
    case kItemBaseHeightNode: ItemHeightNode(cx, it, 0.4); break;   // nlp_formulation.cc:597
    case kItemSwingNode: ItemSwingNode(cx, it); break;
    case kItemSplineAcc: ItemSplineAcc(cx, it); break;
    default: break;
  }
}

}  // namespace tb

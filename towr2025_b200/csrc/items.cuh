// items.cuh — stage A: one thread evaluates one item (a time sample or a node of one constraint family),
// writes the item's constraint rows into g and its partial-derivative slots into the shared G buffer.
#pragma once
#include "device_math.cuh"

namespace tb {

struct ItemCtx {
  const DeviceProgram* P;
  const double* nodes;
  const double* coef;
  double* G;          // segment-relative G buffer (shared memory)
  double* g;          // this problem's constraint values (global), or nullptr
  const ProblemConst* pc;
};

// ------------------------------------------------------------------------------------------------ dynamic
// g rows: dynamic_constraint.cc:63-68,128-148 + single_rigid_body_dynamics.cc:76-102 (same operation order).
// G slots: A_pos[9], A_vel[9], A_acc[9] = d(angular rows)/d(euler pos, vel, acc); f_e[3] per foot; (c - p_e)[3] per foot.
__device__ __noinline__ void ItemDynamic(const ItemCtx& cx, const ItemRec& it) {
  const DeviceProgram& P = *cx.P;
  const int E = it.n_ee;
  Pt3 lin, ang;
  EvalSpline<kNeedPos | kNeedAcc>(P, cx.nodes, cx.coef, it.s[kSBaseLin], lin);
  EvalSpline<kNeedPos | kNeedVel | kNeedAcc>(P, cx.nodes, cx.coef, it.s[kSBaseAng], ang);
  const Trig tr = MakeTrig(ang.p);
  double R[9];
  RotationMatrix(tr, R);
  const double yd = ang.v[1], zd = ang.v[2];
  // M and Mdot (euler_converter.cc:133-166); only columns X and Y of Mdot and X,Y,(2,Z)=1 of M are non-zero
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

  // I_w = R I_b R^T, (R I_b) first, inner index ascending
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

  double Iw_w[3], Iw_wd[3];
  MatVec3(Iw, w, Iw_w);
  MatVec3(Iw, wd, Iw_wd);

  double* G = cx.G + it.gslot;
  double* Gf = G + 27;
  double* Gr = G + 27 + 3 * E;
  double f_sum[3] = {0.0, 0.0, 0.0}, tau_sum[3] = {0.0, 0.0, 0.0};
  for (int e = 0; e < E; ++e) {
    Pt3 f, p, tq;
    EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSForce + e], f);
    EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSMotion + e], p);
    EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSTorque + e], tq);
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
    // Cross(w) * (I_w w): off-diagonal entries in column order
    double gy[3];
    gy[0] = ((-w[2]) * Iw_w[1]) + w[1] * Iw_w[2];
    gy[1] = (w[2] * Iw_w[0]) + (-w[0]) * Iw_w[2];
    gy[2] = ((-w[1]) * Iw_w[0]) + w[0] * Iw_w[1];
    const double m = cx.pc->mass;
    const double grav = 9.80665;
    double* gr = cx.g + it.row[0];
#pragma unroll
    for (int d = 0; d < 3; ++d) gr[d] = (Iw_wd[d] + gy[d]) - tau_sum[d];
    gr[3] = (m * lin.a[0] - f_sum[0]) - 0.0;
    gr[4] = (m * lin.a[1] - f_sum[1]) - 0.0;
    gr[5] = (m * lin.a[2] - f_sum[2]) - (-m * grav);
  }

  // ---- partial derivatives of the angular rows w.r.t. euler (pos, vel, acc)
  double v11[3], v21[3], tmp[3];
  MatTVec3(R, wd, tmp);  MatVec3(Ib, tmp, v11);     // I_b R^T wd
  MatTVec3(R, w, tmp);   MatVec3(Ib, tmp, v21);     // I_b R^T w
  double* Ap = G;
  double* Av = G + 9;
  double* Aa = G + 18;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    // d(omega)/d(theta_c) = dM_c thetadot ; d(omegadot)/d(theta_c) = dMdot_c thetadot + dM_c thetaddot
    double Wp[3], Wdp[3];
    if (c == 0) {
      Wp[0] = Wp[1] = Wp[2] = 0.0;
      Wdp[0] = Wdp[1] = Wdp[2] = 0.0;
    } else if (c == 1) {
      const double dM00 = -tr.sy * tr.cz, dM10 = -tr.sy * tr.sz, dM20 = -tr.cy;
      Wp[0] = dM00 * ang.v[0]; Wp[1] = dM10 * ang.v[0]; Wp[2] = dM20 * ang.v[0];
      const double dMd00 = -tr.cy * tr.cz * yd + tr.sy * tr.sz * zd;
      const double dMd10 = -tr.cy * tr.sz * yd - tr.cz * tr.sy * zd;
      const double dMd20 = tr.sy * yd;
      Wdp[0] = dMd00 * ang.v[0] + dM00 * ang.a[0];
      Wdp[1] = dMd10 * ang.v[0] + dM10 * ang.a[0];
      Wdp[2] = dMd20 * ang.v[0] + dM20 * ang.a[0];
    } else {
      const double dM00 = -tr.cy * tr.sz, dM01 = -tr.cz, dM10 = tr.cy * tr.cz, dM11 = -tr.sz;
      Wp[0] = dM00 * ang.v[0] + dM01 * ang.v[1];
      Wp[1] = dM10 * ang.v[0] + dM11 * ang.v[1];
      Wp[2] = 0.0;
      const double dMd00 = tr.sy * tr.sz * yd - tr.cy * tr.cz * zd, dMd01 = tr.sz * zd;
      const double dMd10 = -tr.cz * tr.sy * yd - tr.cy * tr.sz * zd, dMd11 = -tr.cz * zd;
      Wdp[0] = (dMd00 * ang.v[0] + dMd01 * ang.v[1]) + (dM00 * ang.a[0] + dM01 * ang.a[1]);
      Wdp[1] = (dMd10 * ang.v[0] + dMd11 * ang.v[1]) + (dM10 * ang.a[0] + dM11 * ang.a[1]);
      Wdp[2] = 0.0;
    }
    double dR[9];
    RotationMatrixDeriv(tr, c, dR);
    double t1[3], t2[3], t3[3], u1[3], u2[3], u3[3], q[3];
    MatVec3(dR, v11, t1);                              // dR I_b R^T wd
    MatTVec3(dR, wd, q);  MatVec3(RI, q, t2);          // R I_b dR^T wd
    MatVec3(Iw, Wdp, t3);                              // I_w d(wd)
    MatVec3(dR, v21, u1);
    MatTVec3(dR, w, q);   MatVec3(RI, q, u2);
    MatVec3(Iw, Wp, u3);
    double inner[3] = {u1[0] + u2[0] + u3[0], u1[1] + u2[1] + u3[1], u1[2] + u2[2] + u3[2]};
    double c1[3], c2[3];
    Cross3(w, inner, c1);                              // w x d(I_w w)
    Cross3(Iw_w, Wp, c2);                              // (I_w w) x dw
#pragma unroll
    for (int r = 0; r < 3; ++r) Ap[3 * r + c] = (t1[r] + t2[r] + t3[r]) + (c1[r] - c2[r]);

    // velocity: d(omega)/d(thetadot_c) = M(:,c) ; d(omegadot)/d(thetadot_c) = dMdot/d(thetadot_c) thetadot + Mdot(:,c)
    double Mc[3], Wdv[3];
    if (c == 0) {
      Mc[0] = M00; Mc[1] = M10; Mc[2] = M20;
      Wdv[0] = Md00; Wdv[1] = Md10; Wdv[2] = Md20;
    } else if (c == 1) {
      Mc[0] = M01; Mc[1] = M11; Mc[2] = 0.0;
      Wdv[0] = (-tr.cz * tr.sy) * ang.v[0] + Md01;
      Wdv[1] = (-tr.sy * tr.sz) * ang.v[0] + Md11;
      Wdv[2] = (-tr.cy) * ang.v[0];
    } else {
      Mc[0] = 0.0; Mc[1] = 0.0; Mc[2] = 1.0;
      Wdv[0] = (-tr.cy * tr.sz) * ang.v[0] + (-tr.cz) * ang.v[1];
      Wdv[1] = (tr.cy * tr.cz) * ang.v[0] + (-tr.sz) * ang.v[1];
      Wdv[2] = 0.0;
    }
    double IwM[3], a1[3], b1[3], b2[3];
    MatVec3(Iw, Mc, IwM);
    MatVec3(Iw, Wdv, a1);
    Cross3(w, IwM, b1);
    Cross3(Iw_w, Mc, b2);
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      Av[3 * r + c] = a1[r] + (b1[r] - b2[r]);
      Aa[3 * r + c] = IwM[r];
    }
  }
}

// ------------------------------------------------------------------------------------------ range of motion
// g: range_of_motion_constraint.cc:72-83.  G slots: R^T[9], then per foot d(R^T r_e)/d(euler)[9].
__device__ __noinline__ void ItemRom(const ItemCtx& cx, const ItemRec& it) {
  const DeviceProgram& P = *cx.P;
  Pt3 lin, ang;
  EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSBaseLin], lin);
  EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSBaseAng], ang);
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
    EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSMotion + e], p);
    const double rw[3] = {p.p[0] - lin.p[0], p.p[1] - lin.p[1], p.p[2] - lin.p[2]};
    if (cx.g) {
      double* gr = cx.g + it.row[e];
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
__device__ __noinline__ void ItemForce(const ItemCtx& cx, const ItemRec& it) {
  const DeviceProgram& P = *cx.P;
  Pt3 p, f;
  EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSMotion + it.ee], p);
  EvalSpline<kNeedPos>(P, cx.nodes, cx.coef, it.s[kSForce + it.ee], f);
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
    double* gr = cx.g + it.row[0];
#pragma unroll
    for (int i = 0; i < 5; ++i) gr[i] = (f.p[0] * G[3 * i] + f.p[1] * G[3 * i + 1]) + f.p[2] * G[3 * i + 2];
  }
}

// ------------------------------------------------------------------------------------------------ node-based
// TerrainConstraint (terrain_constraint.cc:61-111) and BaseHeightConstraint (base_height_constraint.cc:58-110)
__device__ __forceinline__ void ItemHeightNode(const ItemCtx& cx, const ItemRec& it, double offset) {
  const double* n = cx.nodes + it.s[0] * 6;
  const TerrainSample ts = EvalTerrain(cx.pc->terrain_id, n[0], n[1]);
  double* G = cx.G + it.gslot;
  G[0] = -ts.hx;
  G[1] = -ts.hy;
  if (cx.g) cx.g[it.row[0]] = (it.type == kItemBaseHeightNode) ? (n[2] - ts.h) - offset : n[2] - ts.h;
}

// SwingConstraint (swing_constraint.cc:54-78)
__device__ __forceinline__ void ItemSwingNode(const ItemCtx& cx, const ItemRec& it) {
  if (!cx.g) return;
  const double* cur = cx.nodes + it.s[0] * 6;
  const double* prev = cur - 6;
  const double* next = cur + 6;
  double* gr = cx.g + it.row[0];
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
__device__ __forceinline__ void ItemSplineAcc(const ItemCtx& cx, const ItemRec& it) {
  if (!cx.g) return;
  const int j = it.s[0];
  const double T = __ldg(&cx.P->polys[j].T);
  double* gr = cx.g + it.row[0];
#pragma unroll
  for (int d = 0; d < 3; ++d) {
    const double prev = 2.0 * cx.coef[j * 6 + d] + (6 * T) * cx.coef[j * 6 + 3 + d];
    const double next = 2.0 * cx.coef[(j + 1) * 6 + d] + (6 * 0.0) * cx.coef[(j + 1) * 6 + 3 + d];
    gr[d] = prev - next;
  }
}

__device__ __forceinline__ void RunItem(const ItemCtx& cx, const ItemRec& it) {
  switch (it.type) {
    case kItemDynamic: ItemDynamic(cx, it); break;
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

// terrain.h — analytic height maps as a switch on the terrain id, usable from host and device code.
//
// Follows towr/src/terrain/height_map_examples.cc:35-211 with the constants of
// towr/include/towr/terrain/examples/height_map_examples.h:45-166 (closed-interval comparisons kept
// exactly), FiveStepStairs from towr/test/hopper_example.cc:53-86, and the basis/derivative formulas of
// towr/src/terrain/height_map.cc:62-163.  Friction coefficient 0.5 (height_map.h:136).
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define TB_HD __host__ __device__ __forceinline__
#else
#define TB_HD inline
#endif

namespace tb {

constexpr double kFrictionCoeff = 0.5;

struct TerrainSample {
  double h;          // height
  double hx, hy;     // first derivatives
  double hxx, hxy, hyx, hyy;
};

// A height map given as data (towr_b200_spec::terrain_table_*, terrain id 8): a profile along x, constant in y.
constexpr int kTerrainTableId = 8, kMaxTerrainKnots = 16;
struct TerrainTable {
  int n, mode;               // mode 0: steps, 1: ramps
  double x[kMaxTerrainKnots], h[kMaxTerrainKnots];
};
// steps: h = 0 before x[0], h[i] on [x[i], x[i+1]), h[n-1] from x[n-1] on.  ramps: h[0] before x[0], linear on
// [x[i], x[i+1]) with h_x = (h[i+1] - h[i]) / (x[i+1] - x[i]), h[n-1] from x[n-1] on.  Second derivatives 0.
TB_HD TerrainSample EvalTerrainTable(const TerrainTable& t, double x) {
  TerrainSample s;
  s.h = 0.0; s.hx = 0.0; s.hy = 0.0; s.hxx = 0.0; s.hxy = 0.0; s.hyx = 0.0; s.hyy = 0.0;
  if (t.n < 1) return s;
  if (x < t.x[0]) {
    if (t.mode == 1) s.h = t.h[0];
    return s;
  }
  int i = 0;
  while (i + 1 < t.n && x >= t.x[i + 1]) ++i;       // x in [x[i], x[i+1]) or beyond the last knot
  if (t.mode == 1 && i + 1 < t.n) {
    const double slope = (t.h[i + 1] - t.h[i]) / (t.x[i + 1] - t.x[i]);
    s.h = t.h[i] + slope * (x - t.x[i]);
    s.hx = slope;
  } else {
    s.h = t.h[i];
  }
  return s;
}

TB_HD TerrainSample EvalTerrain(int id, double x, double y) {
  TerrainSample s;
  s.h = 0.0; s.hx = 0.0; s.hy = 0.0; s.hxx = 0.0; s.hxy = 0.0; s.hyx = 0.0; s.hyy = 0.0;
  switch (id) {
    default:
    case 0: break;                                   // FlatGround
    case 1: {                                        // Block
      const double start = 0.7, length = 3.5, height = 0.5, eps = 0.03;
      const double slope = height / eps;
      if (start <= x && x <= start + eps) { s.h = slope * (x - start); s.hx = slope; }
      if (start + eps <= x && x <= start + length) s.h = height;
      break;
    }
    case 2: {                                        // Stairs
      const double start = 1.0, width = 0.4, h1 = 0.2, h2 = 0.4, top = 1.0;
      if (x >= start) s.h = h1;
      if (x >= start + width) s.h = h2;
      if (x >= start + width + top) s.h = 0.0;
      break;
    }
    case 3: {                                        // Gap (parabola)
      const double gap_start = 1.0, w = 0.5, hh = 1.5;
      const double dx = w / 2.0, xc = gap_start + dx, gap_end = gap_start + w;
      const double a = (4 * hh) / (w * w);
      const double b = -(8 * hh * xc) / (w * w);
      const double c = -(hh * (w - 2 * xc) * (w + 2 * xc)) / (w * w);
      if (gap_start <= x && x <= gap_end) {
        s.h = a * x * x + b * x + c;
        s.hx = 2 * a * x + b;
        s.hxx = 2 * a;
      }
      break;
    }
    case 4: {                                        // Slope
      const double start = 1.0, up = 1.0, down = 1.0, centre = 0.7;
      const double x_down = start + up, x_flat = x_down + down, slope = centre / up;
      if (x >= start) { s.h = slope * (x - start); s.hx = slope; }
      if (x >= x_down) { s.h = centre - slope * (x - x_down); s.hx = -slope; }
      if (x >= x_flat) { s.h = 0.0; s.hx = 0.0; }
      break;
    }
    case 5: {                                        // Chimney
      const double x_start = 1.0, length = 1.5, y_start = 0.5, slope = 3.0;
      const double x_end = x_start + length;
      if (x_start <= x && x <= x_end) { s.h = slope * (y - y_start); s.hy = slope; }
      break;
    }
    case 6: {                                        // ChimneyLR
      const double x_start = 0.5, length = 1.0, y_start = 0.5, slope = 2;
      const double x_end1 = x_start + length, x_end2 = x_start + 2 * length;
      if (x_start <= x && x <= x_end1) { s.h = slope * (y - y_start); s.hy = slope; }
      if (x_end1 <= x && x <= x_end2) { s.h = -slope * (y + y_start); s.hy = -slope; }
      break;
    }
    case 7: {                                        // FiveStepStairs
      const double depth = 0.3, height = 0.15, start = 0.5;
      const int steps = 5;
      if (!(x < start)) {
        int n = static_cast<int>((x - start) / depth);
        s.h = n >= steps ? steps * height : (n + 1) * height;
      }
      break;
    }
  }
  return s;
}

// Dimensions (bit 0: x, bit 1: y) in which the terrain's SECOND derivatives can be non-zero.  Only then does the reference
// store the force / torque constraint's ee-motion block (force_constraint_discretized.cc:52-67 skips scale == 0.0): a
// structure class built for such a terrain declares that block (superset pattern, structure.cc).
TB_HD int TerrainCurvatureDims(int id) { return id == 3 ? 1 : 0; }   // Gap: h_xx != 0 over the parabola

// false: h_x = h_y = 0 everywhere (piecewise constant height): normal and tangents are the same constants at every (x, y)
TB_HD bool TerrainHasSlope(int id) { return !(id == 0 || id == 2 || id == 7); }   // Flat, Stairs, FiveStepStairs
TB_HD bool TerrainHasSlope(int id, const TerrainTable& t) { return id == kTerrainTableId ? t.mode == 1 : TerrainHasSlope(id); }
// bit 2 of a class's curvature_dims: the class admits table terrains (its kernels are the CURV instantiations, the only
// ones that read the table); bits 0, 1 as TerrainCurvatureDims
constexpr int kClassTableTerrain = 4;
// bit 3: the class has a TerrainConstraintHard or a node-based TorqueConstraint set; their items are compiled into the
// CURV instantiations only, so that the kernels of the default constraint list do not carry their code (TerrainHard alone
// was 0.6 % of their run time, measured)
constexpr int kClassTerrainHard = 8;
TB_HD int TerrainClassBits(int id) { return id == kTerrainTableId ? kClassTableTerrain : TerrainCurvatureDims(id); }

// normalised normal / tangents at (x,y): n = (-hx,-hy,1)/|.|, t1 = (1,0,hx)/|.|, t2 = (0,1,hy)/|.|
// v.normalized() = v / sqrt(v.v) with the squared norm summed left to right
TB_HD void Normalize3(const double v[3], double out[3]) {
  double z = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  if (z == 1.0) {            // unit already (flat ground): sqrt(1) = 1 and v/1 = v exactly, skip sqrt + 3 divisions
    out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
  } else if (z > 0.0) {
    double n = sqrt(z);
    out[0] = v[0] / n; out[1] = v[1] / n; out[2] = v[2] / n;
  } else {
    out[0] = v[0]; out[1] = v[1]; out[2] = v[2];
  }
}

TB_HD void TerrainBasis(const TerrainSample& s, double n[3], double t1[3], double t2[3]) {
  double vn[3] = {-s.hx, -s.hy, 1.0};
  double v1[3] = {1.0, 0.0, s.hx};
  double v2[3] = {0.0, 1.0, s.hy};
  Normalize3(vn, n);
  Normalize3(v1, t1);
  Normalize3(v2, t2);
}

// d(normalised basis)/d(dim): cwiseProduct( (|v| e_dim - v_dim v^) / |v|^2 , dv/d(dim) )   height_map.cc:76-90,136-143
// which: 0 normal, 1 tangent1, 2 tangent2
TB_HD void TerrainBasisDeriv(const TerrainSample& s, int which, int dim, double out[3]) {
  double v[3], dv[3];
  const double d2_x = dim == 0 ? s.hxx : s.hxy;   // d(hx)/d(dim)
  const double d2_y = dim == 0 ? s.hyx : s.hyy;   // d(hy)/d(dim)
  if (which == 0) { v[0] = -s.hx; v[1] = -s.hy; v[2] = 1.0; dv[0] = -d2_x; dv[1] = -d2_y; dv[2] = 0.0; }
  else if (which == 1) { v[0] = 1.0; v[1] = 0.0; v[2] = s.hx; dv[0] = 0.0; dv[1] = 0.0; dv[2] = d2_x; }
  else { v[0] = 0.0; v[1] = 1.0; v[2] = s.hy; dv[0] = 0.0; dv[1] = 0.0; dv[2] = d2_y; }
  double sq = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  double nrm = sqrt(sq);
  double vh[3];
  Normalize3(v, vh);
  double inv = 1 / sq;
  for (int i = 0; i < 3; ++i) {
    double unit = (i == dim) ? 1.0 : 0.0;
    double dn = inv * (nrm * unit - v[dim] * vh[i]);
    out[i] = dn * dv[i];
  }
}

}  // namespace tb

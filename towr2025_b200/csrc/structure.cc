// structure.cc — builds the evaluation program (pattern + item/entry tables) of a structure class.
//
// The stored-entry rules restate, per constraint family, which coefficients the reference's Eigen
// code stores: coeffRef inserts, sparse products keep structural unions, Cross() has
// six off-diagonal entries (single_rigid_body_dynamics.cc:47-57), DerivOfRotVecMult(inverse) has no
// roll column in its first row (euler_converter.cc:223-268).  Row order = constraint-set order of
// nlp_formulation.cc:365-378, column order = variable-set order of nlp_formulation.cc:81-101; inside a
// row entries are sorted by column (ifopt's setFromTriplets on a row-major matrix).
#include "structure.h"

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <sstream>
#include <stdexcept>

namespace tb {

std::vector<double> SampleTimes(double T, double dt) {
  std::vector<double> ts;
  double t = 0.0;
  ts.push_back(t);
  const int n = static_cast<int>(std::floor(T / dt));
  for (int i = 0; i < n; ++i) {
    t += dt;                 // repeated addition, not i*dt
    ts.push_back(t);
  }
  ts.push_back(T);           // may duplicate the last sample
  return ts;
}

double RuntimePow(double x, int n) {
  volatile int e = n;          // not a compile-time constant: no folding into x * x
  return std::pow(x, static_cast<double>(e));
}

void HermiteBasis(double T, double t, double o[12]) {
  const double t2 = RuntimePow(t, 2), t3 = RuntimePow(t, 3);
  const double T2 = RuntimePow(T, 2), T3 = RuntimePow(T, 3);
  // position
  o[0] = (2 * t3) / T3 - (3 * t2) / T2 + 1;   o[1] = t - (2 * t2) / T + t3 / T2;
  o[2] = (3 * t2) / T2 - (2 * t3) / T3;       o[3] = t3 / T2 - t2 / T;
  // velocity
  o[4] = (6 * t2) / T3 - (6 * t) / T2;        o[5] = (3 * t2) / T2 - (4 * t) / T + 1;
  o[6] = (6 * t) / T2 - (6 * t2) / T3;        o[7] = (3 * t2) / T2 - (2 * t) / T;
  // acceleration
  o[8] = (12 * t) / T3 - 6 / T2;              o[9] = (6 * t) / T2 - 4 / T;
  o[10] = 6 / T2 - (12 * t) / T3;             o[11] = (6 * t) / T2 - 2 / T;
}

int SegmentId(double t_global, const std::vector<double>& durations) {
  double acc = 0.0;
  for (size_t i = 0; i < durations.size(); ++i) {
    acc += durations[i];
    if (acc >= t_global - 1e-10) return static_cast<int>(i);
  }
  return static_cast<int>(durations.size()) - 1;
}

double LocalTime(double t_global, int id, const std::vector<double>& durations) {
  double t = t_global;
  for (int i = 0; i < id; ++i) t -= durations[i];
  return t;
}

namespace {

// value = G[g] * B[b], or G[g] * G[b] when gg (runtime weights).  code >= 0: G[g] is a component of the terrain basis
// (BasisCode, program.h), a constant of the problem on a terrain without slope
struct Term { int g = 0, b = 0; bool gg = false; int code = -1; };
struct ColBasis { int idx, dim; double v[3]; int sd; };   // v[dx]: d{pos,vel,acc}/d(variable idx); sd = side*2+deriv of the last hit
struct SplineAt { int sample; int poly; double t_local; };

class Builder {
 public:
  Builder(const towr_b200_spec& s, HostProgram& hp, const std::vector<char>* new_segment_before_group)
      : s_(s), hp_(hp), plan_(new_segment_before_group) {}

  void Run();
  std::vector<int> group_g_size;    // filled by Run(): G slots each group allocated

 private:
  // ---- tables
  int AddB(double v) {
    uint64_t bits;
    std::memcpy(&bits, &v, 8);
    auto it = b1_.find(bits);
    if (it != b1_.end()) return it->second;
    int i = static_cast<int>(hp_.B.size());
    hp_.B.push_back(v);
    b1_[bits] = i;
    return i;
  }
  // a run of consecutive constants (items read them by position); identical runs are shared
  int AddBRun(const std::vector<double>& v) {
    std::string key(reinterpret_cast<const char*>(v.data()), v.size() * sizeof(double));
    auto it = brun_.find(key);
    if (it != brun_.end()) return it->second;
    int i = static_cast<int>(hp_.B.size());
    hp_.B.insert(hp_.B.end(), v.begin(), v.end());
    brun_[key] = i;
    return i;
  }
  SplineAt At(const VarSet& set, double t_global) {
    SplineAt a;
    a.poly = SegmentId(t_global, set.poly_T);
    a.t_local = LocalTime(t_global, a.poly, set.poly_T);
    SampleRec r;
    r.poly = set.poly0 + a.poly;
    r.pad = 0;
    r.t = a.t_local;
    r.t2 = RuntimePow(a.t_local, 2);
    r.t3 = RuntimePow(a.t_local, 3);
    a.sample = static_cast<int>(hp_.samples.size());
    hp_.samples.push_back(r);
    return a;
  }
  // optimised durations: a PhaseSpline sample is resolved per problem on the device; kind_idx 0 motion, 1 force, 2 torque
  bool rt() const { return hp_.optimize_durations; }
  int AtRt(int kind_idx, int ee, double t_global) {
    SampleRec r;
    r.poly = -(1 + kind_idx * s_.n_ee + ee);
    r.pad = 0;
    r.t = t_global;
    r.t2 = 0.0;
    r.t3 = 0.0;
    hp_.samples.push_back(r);
    return static_cast<int>(hp_.samples.size()) - 1;
  }
  // PhaseSpline's pre-filled structure (phase_spline.cc:45-51): every variable of the set, with its dimension
  static std::vector<std::pair<int, int>> AllCols(const VarSet& set) {
    std::vector<std::pair<int, int>> c;
    for (int idx = 0; idx < set.n_vars; ++idx) c.emplace_back(idx, set.refs[idx].front().dim);
    return c;
  }
  Term GG(int g, int w) const { Term t; t.g = g; t.b = w; t.gg = true; return t; }
  // the finished ee-schedule block of one item: rows x (P-1) values at slot0 + r * (P-1) + j
  void AddScheduleBlock(int row0_local, int n_rows, int ee, int slot0) {
    const VarSet& sc = Set(kEESchedule, ee);
    const int one = AddB(1.0);
    for (int r = 0; r < n_rows; ++r)
      for (int j = 0; j < sc.n_vars; ++j) {
        Term tm; tm.g = slot0 + r * sc.n_vars + j; tm.b = one;
        Add(row0_local + r, sc.col0 + j, tm);
      }
  }
  // stored columns of NodeSpline::GetJacobianWrtNodes(poly, t_local, *) with their values for all three dx
  // (node_spline.cc:84-112): ascending idx, start side before end side, contributions added from 0.0
  std::vector<ColBasis> Columns(const VarSet& set, int poly, double t_local) const {
    double hb[12];
    HermiteBasis(set.poly_T[poly], t_local, hb);
    std::vector<ColBasis> cols;
    for (int idx = 0; idx < set.n_vars; ++idx) {
      ColBasis cb{idx, -1, {0.0, 0.0, 0.0}, 0};
      for (const NodeRef& r : set.refs[idx])
        for (int side = 0; side < 2; ++side)
          if (poly + side == r.node) {
            cb.dim = r.dim;
            cb.sd = side * 2 + r.deriv;
            for (int dx = 0; dx < 3; ++dx) cb.v[dx] += hb[dx * 4 + side * 2 + r.deriv];
          }
      if (cb.dim >= 0) cols.push_back(cb);
    }
    return cols;
  }

  // ---- segments / groups
  // A group is the unit of segment packing.  When a group starts a new segment in the middle of a constraint set
  // (time-discretised sets are split by sample ranges), the rows finished so far are flushed first so that every
  // segment owns a contiguous range of entries.  next_row_local = first row of the set the new group writes.
  void BeginGroup(int next_row_local = 0) {
    int gi = static_cast<int>(group_g_size.size());
    bool new_seg = hp_.segments.empty() || (plan_ ? (*plan_)[gi] != 0 : true);
    if (new_seg) {
      FlushRows(next_row_local);
      CloseSegment();
      SegmentRec sg{};
      sg.item0 = static_cast<int>(hp_.items.size());
      sg.entry0 = static_cast<int>(hp_.entries.size());
      sg.g_size = kGReserved;
      hp_.segments.push_back(sg);
    }
    group_g_size.push_back(0);
  }
  void CloseSegment() {
    if (hp_.segments.empty()) return;
    SegmentRec& sg = hp_.segments.back();
    sg.n_items = static_cast<int>(hp_.items.size()) - sg.item0;
    sg.n_entries = static_cast<int>(hp_.entries.size()) - sg.entry0;
    hp_.g_capacity = std::max(hp_.g_capacity, sg.g_size);
  }
  int AllocG(int n) {
    SegmentRec& sg = hp_.segments.back();
    int g = sg.g_size;
    sg.g_size += n;
    group_g_size.back() += n;
    if (sg.g_size > 8191) throw std::runtime_error("unsupported: a constraint set needs more than 8191 G slots in one segment");
    return g;
  }

  // ---- rows of the constraint set being emitted
  void BeginSet(const std::string& name, int n_rows) {
    set_row0_ = hp_.n_rows;
    hp_.con_sets.push_back(ConSet{name, hp_.n_rows, n_rows});
    hp_.n_rows += n_rows;
    rows_.assign(n_rows, {});
    rows_flushed_ = 0;
    hp_.row_bounds.resize(hp_.n_rows, RowBound{kBoundZero, 0, 0});
  }
  void Bound(int row_local, int kind, int ee = 0, int dim = 0) { hp_.row_bounds.at(set_row0_ + row_local) = RowBound{kind, ee, dim}; }
  void Add(int row_local, int col, const Term& t, bool overwrite = false) {
    auto& r = rows_.at(row_local);
    for (auto& e : r)
      if (e.first == col) {
        if (!overwrite) throw std::runtime_error("duplicate Jacobian entry emitted");
        e.second = t;
        return;
      }
    r.emplace_back(col, t);
  }
  void FlushRows(int upto) {            // rows [rows_flushed_, upto) of the open set are complete
    for (int r = rows_flushed_; r < upto && r < static_cast<int>(rows_.size()); ++r) {
      auto& row = rows_[r];
      std::stable_sort(row.begin(), row.end(), [](const auto& a, const auto& b) { return a.first < b.first; });
      for (auto& e : row) {
        const Term& t = e.second;
        if (t.g < 0 || t.g > 65535 || t.b < 0 || t.b > 32767) throw std::runtime_error("entry descriptor out of range");
        EntryRec er = static_cast<uint32_t>(t.g) | (static_cast<uint32_t>(t.b) << 16) | (t.gg ? kEntryGG : 0u);
        hp_.entries.push_back(er);
        hp_.entry_code.push_back(static_cast<signed char>(t.gg ? -1 : t.code));
        hp_.pat_rows.push_back(set_row0_ + r);
        hp_.pat_cols.push_back(e.first);
      }
      row.clear();
    }
    if (upto > rows_flushed_) rows_flushed_ = std::min(upto, static_cast<int>(rows_.size()));
  }
  void EndSet() {
    FlushRows(static_cast<int>(rows_.size()));
    rows_.clear();
    rows_flushed_ = 0;
  }
  ItemRec NewItem(int type) const {
    ItemRec it;
    std::memset(&it, 0, sizeof(it));
    it.type = type;
    it.n_ee = static_cast<short>(s_.n_ee);
    return it;
  }
  const VarSet& Set(int kind, int ee = -1) const {
    const VarSet* v = hp_.Find(kind, ee);
    if (!v) throw std::runtime_error("variable set missing");
    return *v;
  }

  void EmitTerrain(int ee);
  void EmitDynamic();
  void EmitSplineAcc();
  void EmitRom();
  void EmitForceDisc(int ee);
  void EmitForceNodes(int ee);
  void EmitTorqueDisc(int ee);
  void EmitBaseRom();
  void EmitTerrainHard(int ee);
  void EmitTorqueNodes(int ee);
  void EmitEELinear(int def);
  void EmitRomSplit();
  void EmitTotalDuration(int ee);
  void EmitSwing(int ee);
  void EmitBaseHeight();
  void EmitHeightRows(const VarSet& set, int item_type, int ee);

  const towr_b200_spec& s_;
  HostProgram& hp_;
  const std::vector<char>* plan_;
 public:
  int rom_feet_per_group = 0;   // 0: all feet of a sample share one RomPre item (one group); else groups of this many feet
 private:
  std::map<uint64_t, int> b1_;
  std::map<std::string, int> brun_;
  std::vector<std::vector<std::pair<int, Term>>> rows_;
  int set_row0_ = 0;
  int rows_flushed_ = 0;
};

// Cross(v)(r,d) = sign * v[k] for r != d (single_rigid_body_dynamics.cc:47-57)
inline void CrossEntry(int r, int d, int* k, double* sign) {
  *k = 3 - r - d;
  *sign = ((d - r + 3) % 3 == 2) ? 1.0 : -1.0;
}

// TerrainConstraint / BaseHeightConstraint: one row per node id >= 1: z - h(x,y) [- offset];
// entries (row, z) = 1, (row, x) = -h_x, (row, y) = -h_y  (terrain_constraint.cc:93-111, base_height_constraint.cc:90-110)
void Builder::EmitHeightRows(const VarSet& set, int item_type, int ee) {
  const int one = AddB(1.0);
  for (int node = 1; node < set.n_nodes; ++node) {
    ItemRec it = NewItem(item_type);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + node - 1;
    it.s[0] = set.node_slot0 + node;
    it.gslot = AllocG(2);
    hp_.items.push_back(it);
    if (item_type == kItemBaseHeightNode) Bound(node - 1, kBoundBaseHeight);
    else Bound(node - 1, set.IsConstantNode(node) ? kBoundZero : kBoundTerrainSwing, ee);
    Term tz; tz.g = kGOne; tz.b = one;
    Add(node - 1, set.col0 + set.OptIndex(node, 0, 2), tz, true);
    for (int dim = 0; dim < 2; ++dim) {
      Term t; t.g = it.gslot + dim; t.b = one; t.code = kCodeHeight + dim;
      Add(node - 1, set.col0 + set.OptIndex(node, 0, dim), t, true);
    }
  }
}

void Builder::EmitTerrain(int ee) {
  const VarSet& m = Set(kEEMotion, ee);
  BeginGroup();
  BeginSet("terrain-" + m.name, m.n_nodes - 1);
  EmitHeightRows(m, kItemTerrainNode, ee);
  EndSet();
}

void Builder::EmitBaseHeight() {
  const VarSet& lin = Set(kBaseLin);
  BeginGroup();
  BeginSet("base-height-" + lin.name, lin.n_nodes - 1);
  EmitHeightRows(lin, kItemBaseHeightNode, 0);
  EndSet();
}

// DynamicConstraint, 6 rows per sample (dynamic_constraint.cc:57-126 with single_rigid_body_dynamics.cc:104-204).
// Per sample: level-0 items DynPre (everything that depends on the euler spline only) and DynFoot x E (forces, lever
// arms, the foot's torque term); level-1 items DynWrench (the g rows), DynJlin x3 (one row of the base-lin angular
// block) and DynCol x3 (one euler column of the base-ang block).
// G layout per sample: Jang[3][12] at r*12 + sd*3 + c | Jlin[3][12] at r*12 + sd*3 + d | f_e[3] | (c - p_e)[3] |
// scratch: foot torque terms [3] per foot, DynPre block.
// B run of the sample (ItemRec::b0): base-lin pos basis [4], then base-ang pos/vel/acc basis [4] each, by sd.
void Builder::EmitDynamic() {
  const int E = s_.n_ee;
  const VarSet& lin = Set(kBaseLin);
  const VarSet& ang = Set(kBaseAng);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_dynamic);
  const int one = AddB(1.0);
  // odd, so that the same slot of neighbouring samples (threads of one warp) falls into different shared-memory banks
  int per_sample = (72 + 9 * E + kDynPreSize) | 1;
  std::vector<int> foot_block(E, 0);                     // optimised durations: offset of each foot's block in the sample
  if (rt()) {
    per_sample = DynNegOffset(E) + 6 * E;
    for (int e = 0; e < E; ++e) {
      foot_block[e] = per_sample;
      per_sample += Set(kEEMotion, e).n_vars + Set(kEEForce, e).n_vars + Set(kEETorque, e).n_vars +
                    6 * Set(kEESchedule, e).n_vars;
    }
    per_sample |= 1;
  }
  const int chunk = std::max(1, 1024 / per_sample);      // samples per group
  for (size_t k = 0; k < ts.size(); ++k) {
    const double t = ts[k];
    const int r0 = static_cast<int>(k) * 6;
    if (k % chunk == 0) {
      if (k == 0) { BeginGroup(); BeginSet("dynamic", static_cast<int>(ts.size()) * 6); }
      else BeginGroup(r0);
    }
    ItemRec it = NewItem(kItemDynWrench);
    it.row[0] = set_row0_ + r0;
    const int g0 = AllocG(per_sample);
    it.gslot = g0;
    it.aux = g0 + 72 + 9 * E;
    const int gJang = g0 + kDynJang, gJlin = g0 + kDynJlin, gF = g0 + kDynF, gR = g0 + kDynF + 3 * E;

    SplineAt al = At(lin, t), aa = At(ang, t);
    it.s[kSBaseLin] = al.sample;
    it.s[kSBaseAng] = aa.sample;
    double hbl[12], hba[12];
    HermiteBasis(lin.poly_T[al.poly], al.t_local, hbl);
    HermiteBasis(ang.poly_T[aa.poly], aa.t_local, hba);
    std::vector<double> run(hbl, hbl + 4);
    run.insert(run.end(), hba, hba + 12);
    it.b0 = AddBRun(run);
    // base-lin: angular rows -sum_e Cross(f_e) J_pos (finished by the item), linear rows m J_acc
    for (const ColBasis& c : Columns(lin, al.poly, al.t_local)) {
      for (int r = 0; r < 3; ++r) {
        if (r == c.dim) continue;
        Term tm; tm.g = gJlin + r * 12 + c.sd * 3 + c.dim; tm.b = one;
        Add(r0 + r, lin.col0 + c.idx, tm);
      }
      Term tm; tm.g = kGMass; tm.b = AddB(c.v[2]);
      Add(r0 + 3 + c.dim, lin.col0 + c.idx, tm);
    }
    // base-ang: full 3 x 12 block, A_pos*Bp + A_vel*Bv + A_acc*Ba finished by the items
    for (const ColBasis& c : Columns(ang, aa.poly, aa.t_local))
      for (int r = 0; r < 3; ++r) {
        Term tm; tm.g = gJang + r * 12 + c.sd * 3 + c.dim; tm.b = one;
        Add(r0 + r, ang.col0 + c.idx, tm);
      }
    for (int e = 0; e < E; ++e) {
      const VarSet& mo = Set(kEEMotion, e);
      const VarSet& fo = Set(kEEForce, e);
      const VarSet& to = Set(kEETorque, e);
      if (rt()) {
        // every column of the three PhaseSplines (pre-filled structure), weights written by the foot's DynFoot item
        it.s[kSMotion + e] = AtRt(0, e, t);
        it.s[kSForce + e] = AtRt(1, e, t);
        it.s[kSTorque + e] = AtRt(2, e, t);
        const int gNegF = g0 + DynNegOffset(E), gNegR = gNegF + 3 * E;
        const int wP = g0 + foot_block[e], wF = wP + mo.n_vars, wT = wF + fo.n_vars, sched = wT + to.n_vars;
        for (auto& c : AllCols(mo))
          for (int r = 0; r < 3; ++r) {
            if (r == c.second) continue;
            int kk; double sg;
            CrossEntry(r, c.second, &kk, &sg);
            Add(r0 + r, mo.col0 + c.first, GG((sg > 0 ? gF : gNegF) + 3 * e + kk, wP + c.first));
          }
        for (auto& c : AllCols(fo)) {
          for (int r = 0; r < 3; ++r) {
            if (r == c.second) continue;
            int kk; double sg;
            CrossEntry(r, c.second, &kk, &sg);
            Add(r0 + r, fo.col0 + c.first, GG((sg > 0 ? gR : gNegR) + 3 * e + kk, wF + c.first));
          }
          Add(r0 + 3 + c.second, fo.col0 + c.first, GG(kGMinusOne, wF + c.first));
        }
        for (auto& c : AllCols(to)) Add(r0 + c.second, to.col0 + c.first, GG(kGMinusOne, wT + c.first));
        AddScheduleBlock(r0, 6, e, sched);
        continue;
      }
      SplineAt am = At(mo, t), af = At(fo, t), at = At(to, t);
      it.s[kSMotion + e] = am.sample;
      it.s[kSForce + e] = af.sample;
      it.s[kSTorque + e] = at.sample;
      // ee-motion: Cross(f_e) J_p on the angular rows
      for (const ColBasis& c : Columns(mo, am.poly, am.t_local))
        for (int r = 0; r < 3; ++r) {
          if (r == c.dim) continue;
          int kk; double sg;
          CrossEntry(r, c.dim, &kk, &sg);
          Term tm; tm.g = gF + 3 * e + kk; tm.b = AddB(sg * c.v[0]);
          Add(r0 + r, mo.col0 + c.idx, tm);
        }
      // ee-force: Cross(c - p_e) J_f on the angular rows, -J_f on the linear rows
      for (const ColBasis& c : Columns(fo, af.poly, af.t_local)) {
        for (int r = 0; r < 3; ++r) {
          if (r == c.dim) continue;
          int kk; double sg;
          CrossEntry(r, c.dim, &kk, &sg);
          Term tm; tm.g = gR + 3 * e + kk; tm.b = AddB(sg * c.v[0]);
          Add(r0 + r, fo.col0 + c.idx, tm);
        }
        Term tm; tm.g = kGOne; tm.b = AddB(-c.v[0]);
        Add(r0 + 3 + c.dim, fo.col0 + c.idx, tm);
      }
      // ee-torque: -J_tau on the angular rows
      for (const ColBasis& c : Columns(to, at.poly, at.t_local)) {
        Term tm; tm.g = kGOne; tm.b = AddB(-c.v[0]);
        Add(r0 + c.dim, to.col0 + c.idx, tm);
      }
    }
    ItemRec pre = it;
    pre.type = kItemDynPre;
    pre.gslot = it.aux;
    hp_.items.push_back(pre);
    for (int e = 0; e < E; ++e) {
      ItemRec foot = it;
      foot.type = kItemDynFoot;
      foot.ee = static_cast<short>(e);
      if (rt()) foot.aux = g0 + foot_block[e];
      hp_.items.push_back(foot);
    }
    hp_.items.push_back(it);
    for (int c = 0; c < 3; ++c) {
      ItemRec col = it;
      col.type = kItemDynCol;
      col.sub = static_cast<short>(c);
      hp_.items.push_back(col);
    }
    {
      ItemRec jl = it;                                    // one item finishes all three rows of the base-lin angular block
      jl.type = kItemDynJlin;
      jl.sub = 3;
      hp_.items.push_back(jl);
    }
  }
  EndSet();
}

// SplineAccConstraint for base-lin and base-ang: acc at the end of poly j minus acc at the start of poly j+1
// (spline_acc_constraint.cc:48-80); the Jacobian is constant
void Builder::EmitSplineAcc() {
  BeginGroup();
  for (int which = 0; which < 2; ++which) {
    const VarSet& set = Set(which == 0 ? kBaseLin : kBaseAng);
    const int nj = set.n_polys() - 1;
    BeginSet("splineacc-" + set.name, 3 * nj);
    for (int j = 0; j < nj; ++j) {
      ItemRec it = NewItem(kItemSplineAcc);
      it.row[0] = set_row0_ + 3 * j;
      it.s[0] = set.poly0 + j;
      it.gslot = 0;
      hp_.items.push_back(it);
      std::vector<ColBasis> prev = Columns(set, j, set.poly_T[j]);
      std::vector<ColBasis> next = Columns(set, j + 1, 0.0);
      struct Pair { int dim = 0; double prev = 0.0, next = 0.0; };   // missing side contributes 0.0
      std::map<int, Pair> merged;
      for (const ColBasis& c : prev) { merged[c.idx].dim = c.dim; merged[c.idx].prev = c.v[2]; }
      for (const ColBasis& c : next) { merged[c.idx].dim = c.dim; merged[c.idx].next = c.v[2]; }
      for (auto& m : merged) {
        Term tm; tm.g = kGOne; tm.b = AddB(m.second.prev - m.second.next);
        Add(3 * j + m.second.dim, set.col0 + m.first, tm);
      }
    }
    EndSet();
  }
}

// RangeOfMotionConstraint, one set per foot, 3 rows per sample (range_of_motion_constraint.cc:66-131).
// Per sample one level-0 item (RomPre: base position, R, dR/d(euler) -> 39 scratch slots) shared by the feet, and
// per (sample, foot) one level-1 item (RomFoot: 3 g rows, d(R^T r_e)/d(euler) -> 9 slots).
void Builder::EmitRom() {
  const int E = s_.n_ee;
  const VarSet& lin = Set(kBaseLin);
  const VarSet& ang = Set(kBaseAng);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_range_of_motion);
  const int K = static_cast<int>(ts.size());
  struct PerSample { int pre; SplineAt al, aa; };
  std::vector<PerSample> ps(K);
  const int fpg = rom_feet_per_group > 0 ? rom_feet_per_group : E;
  for (int e = 0; e < E; ++e) {
    if (e % fpg == 0) {                                    // a new group with its own RomPre items
      BeginGroup();
      for (int k = 0; k < K; ++k) {
        ItemRec it = NewItem(kItemRomPre);
        ps[k].pre = AllocG(kRomPreSize);
        it.gslot = ps[k].pre;
        ps[k].al = At(lin, ts[k]);
        ps[k].aa = At(ang, ts[k]);
        it.s[kSBaseLin] = ps[k].al.sample;
        it.s[kSBaseAng] = ps[k].aa.sample;
        hp_.items.push_back(it);
      }
    }
    const VarSet& mo = Set(kEEMotion, e);
    BeginSet("rangeofmotion-" + std::to_string(e), 3 * K);
    for (int k = 0; k < K; ++k) {
      ItemRec it = NewItem(kItemRomFoot);
      it.ee = static_cast<short>(e);
      it.row[0] = set_row0_ + 3 * k;
      it.aux = ps[k].pre;
      it.gslot = AllocG(9);
      SplineAt am = At(mo, ts[k]);
      it.s[kSMotion + e] = am.sample;
      hp_.items.push_back(it);
      for (int r = 0; r < 3; ++r) Bound(3 * k + r, kBoundRom, e, r);
      const int gR = ps[k].pre + kRpR, gD = it.gslot;      // R^T(r,d) = R[d][r] lives at gR + d*3 + r
      for (const ColBasis& c : Columns(lin, ps[k].al.poly, ps[k].al.t_local))
        for (int r = 0; r < 3; ++r) {            // (-1 * R^T) J_c
          Term tm; tm.g = gR + c.dim * 3 + r; tm.b = AddB(-c.v[0]);
          Add(3 * k + r, lin.col0 + c.idx, tm);
        }
      for (const ColBasis& c : Columns(ang, ps[k].aa.poly, ps[k].aa.t_local))
        for (int r = 0; r < 3; ++r) {            // DerivOfRotVecMult(t, r_W, inverse): no roll column in row X
          if (r == 0 && c.dim == 0) continue;
          Term tm; tm.g = gD + r * 3 + c.dim; tm.b = AddB(c.v[0]);
          Add(3 * k + r, ang.col0 + c.idx, tm);
        }
      for (const ColBasis& c : Columns(mo, am.poly, am.t_local))
        for (int r = 0; r < 3; ++r) {            // R^T J_p
          Term tm; tm.g = gR + c.dim * 3 + r; tm.b = AddB(c.v[0]);
          Add(3 * k + r, mo.col0 + c.idx, tm);
        }
    }
    EndSet();
  }
}

// ForceConstraintDiscretized, 5 rows per sample (force_constraint_discretized.cc:97-221)
void Builder::EmitForceDisc(int ee) {
  const VarSet& mo = Set(kEEMotion, ee);
  const VarSet& fo = Set(kEEForce, ee);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_force);
  // G slots per sample: 15, plus with optimised durations W_f[n_cols] and the finished ee-schedule block [5][P-1];
  // classes that declare the curvature block (hp_.curvature_dims): S[dim][5] = f . d b_i / d(dim) after those, and with
  // optimised durations the ee-motion spline's weights W_p[n_cols] (ForceCurvOffset, program.h)
  const int nd = CurvDimCount(hp_.curvature_dims);
  const int off_S = ForceCurvOffset(rt() ? fo.n_vars : 0, rt() ? Set(kEESchedule, ee).n_vars : 0);
  const int per_item = (off_S + 5 * nd + ((rt() && nd) ? mo.n_vars : 0)) | 1;
  const int chunk = std::max(1, 960 / per_item);    // samples per group
  for (size_t k = 0; k < ts.size(); ++k) {
    if (k % chunk == 0) {
      if (k == 0) { BeginGroup(); BeginSet("force-disc-" + fo.name, static_cast<int>(ts.size()) * 5); }
      else BeginGroup(static_cast<int>(k) * 5);
    }
    ItemRec it = NewItem(kItemForce);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + static_cast<int>(k) * 5;
    it.gslot = AllocG(per_item);
    it.sub = static_cast<short>(hp_.curvature_dims);
    const int bk[5] = {kBoundForceNormal, kBoundLE0, kBoundGE0, kBoundLE0, kBoundGE0};
    for (int i = 0; i < 5; ++i) Bound(static_cast<int>(k) * 5 + i, bk[i], ee);
    // ee-motion block f . d b_i / d(dim) * J_p(row dim): the reference stores it only where the scale is non-zero
    // (force_constraint_discretized.cc:52-67,163-183); a class with curvature declares it for every sample (superset),
    // entries the reference would skip are exact zeros
    auto dim_slot = [&](int dim) { return dim == 0 ? 0 : (hp_.curvature_dims & 1 ? 1 : 0); };
    if (rt()) {
      it.s[kSMotion + ee] = AtRt(0, ee, ts[k]);
      it.s[kSForce + ee] = AtRt(1, ee, ts[k]);
      hp_.items.push_back(it);
      for (auto& c : AllCols(mo))
        if (c.second < 2 && (hp_.curvature_dims >> c.second) & 1)
          for (int i = 0; i < 5; ++i)
            Add(static_cast<int>(k) * 5 + i, mo.col0 + c.first,
                GG(it.gslot + off_S + 5 * dim_slot(c.second) + i, it.gslot + off_S + 5 * nd + c.first));
      for (auto& c : AllCols(fo))
        for (int i = 0; i < 5; ++i)
          Add(static_cast<int>(k) * 5 + i, fo.col0 + c.first, GG(it.gslot + 3 * i + c.second, it.gslot + 15 + c.first));
      AddScheduleBlock(static_cast<int>(k) * 5, 5, ee, it.gslot + 15 + fo.n_vars);
      continue;
    }
    SplineAt am = At(mo, ts[k]), af = At(fo, ts[k]);
    it.s[kSMotion + ee] = am.sample;
    it.s[kSForce + ee] = af.sample;
    hp_.items.push_back(it);
    for (const ColBasis& c : Columns(fo, af.poly, af.t_local))
      for (int i = 0; i < 5; ++i) {              // b_i . J_f
        Term tm; tm.g = it.gslot + 3 * i + c.dim; tm.b = AddB(c.v[0]); tm.code = kCodeForce + 3 * i + c.dim;
        Add(static_cast<int>(k) * 5 + i, fo.col0 + c.idx, tm);
      }
    for (const ColBasis& c : Columns(mo, am.poly, am.t_local))
      if (c.dim < 2 && (hp_.curvature_dims >> c.dim) & 1)
        for (int i = 0; i < 5; ++i) {
          Term tm; tm.g = it.gslot + off_S + 5 * dim_slot(c.dim) + i; tm.b = AddB(c.v[0]);
          Add(static_cast<int>(k) * 5 + i, mo.col0 + c.idx, tm);
        }
  }
  EndSet();
}

// RangeOfMotionConstraint split into sample groups (every foot's set carries its own RomPre items), used when one
// group for all samples would not fit a segment, and with optimised durations: there the ee-motion block stores every
// column of the set (weights written by the RomFoot item) and the ee-schedule block is finished by the item.
void Builder::EmitRomSplit() {
  const int E = s_.n_ee;
  const VarSet& lin = Set(kBaseLin);
  const VarSet& ang = Set(kBaseAng);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_range_of_motion);
  const int K = static_cast<int>(ts.size());
  for (int e = 0; e < E; ++e) {
    const VarSet& mo = Set(kEEMotion, e);
    const int nsched = rt() ? Set(kEESchedule, e).n_vars : 0;
    const int foot_size = (9 + (rt() ? mo.n_vars + 3 * nsched : 0)) | 1;
    const int chunk = std::max(1, 960 / (kRomPreSize + foot_size));
    for (int k = 0; k < K; ++k) {
      if (k % chunk == 0) {
        if (k == 0) { BeginGroup(); BeginSet("rangeofmotion-" + std::to_string(e), 3 * K); }
        else BeginGroup(3 * k);
      }
      ItemRec pre = NewItem(kItemRomPre);
      pre.gslot = AllocG(kRomPreSize);
      SplineAt al = At(lin, ts[k]), aa = At(ang, ts[k]);
      pre.s[kSBaseLin] = al.sample;
      pre.s[kSBaseAng] = aa.sample;
      hp_.items.push_back(pre);
      ItemRec it = NewItem(kItemRomFoot);
      it.ee = static_cast<short>(e);
      it.row[0] = set_row0_ + 3 * k;
      it.aux = pre.gslot;
      it.gslot = AllocG(foot_size);
      SplineAt am{0, 0, 0.0};
      if (rt()) it.s[kSMotion + e] = AtRt(0, e, ts[k]);
      else { am = At(mo, ts[k]); it.s[kSMotion + e] = am.sample; }
      hp_.items.push_back(it);
      for (int r = 0; r < 3; ++r) Bound(3 * k + r, kBoundRom, e, r);
      const int gR = pre.gslot + kRpR, gD = it.gslot;
      for (const ColBasis& c : Columns(lin, al.poly, al.t_local))
        for (int r = 0; r < 3; ++r) {
          Term tm; tm.g = gR + c.dim * 3 + r; tm.b = AddB(-c.v[0]);
          Add(3 * k + r, lin.col0 + c.idx, tm);
        }
      for (const ColBasis& c : Columns(ang, aa.poly, aa.t_local))
        for (int r = 0; r < 3; ++r) {
          if (r == 0 && c.dim == 0) continue;
          Term tm; tm.g = gD + 3 * r + c.dim; tm.b = AddB(c.v[0]);
          Add(3 * k + r, ang.col0 + c.idx, tm);
        }
      if (rt()) {
        for (auto& c : AllCols(mo))
          for (int r = 0; r < 3; ++r) Add(3 * k + r, mo.col0 + c.first, GG(gR + c.second * 3 + r, it.gslot + 9 + c.first));
        AddScheduleBlock(3 * k, 3, e, it.gslot + 9 + mo.n_vars);
      } else {
        for (const ColBasis& c : Columns(mo, am.poly, am.t_local))
          for (int r = 0; r < 3; ++r) {            // R^T J_p
            Term tm; tm.g = gR + c.dim * 3 + r; tm.b = AddB(c.v[0]);
            Add(3 * k + r, mo.col0 + c.idx, tm);
          }
      }
    }
    EndSet();
  }
}

// TotalDurationConstraint (total_duration_constraint.cc:36-72): one row per foot, 1.0 in every ee-schedule column
void Builder::EmitTotalDuration(int ee) {
  const VarSet& sc = Set(kEESchedule, ee);
  BeginGroup();
  BeginSet("totalduration-" + std::to_string(ee), 1);
  ItemRec it = NewItem(kItemTotalDuration);
  it.ee = static_cast<short>(ee);
  it.row[0] = set_row0_;
  hp_.items.push_back(it);
  Bound(0, kBoundTotalDuration, ee);
  const int one = AddB(1.0);
  for (int j = 0; j < sc.n_vars; ++j) {
    Term tm; tm.g = kGOne; tm.b = one;
    Add(0, sc.col0 + j, tm);
  }
  EndSet();
}

// ForceConstraint, node-based (force_constraint.cc:37-171; used when dt_constraint_force <= 0): 5 rows per non-constant
// force node; the ee-force block holds b_i, the ee-motion block (x, y of the stance foothold) f . d b_i - always stored.
void Builder::EmitForceNodes(int ee) {
  const VarSet& mo = Set(kEEMotion, ee);
  const VarSet& fo = Set(kEEForce, ee);
  const std::vector<int> nodes = fo.NonConstantNodes();
  const int one = AddB(1.0);
  BeginGroup();
  BeginSet("force-" + fo.name, static_cast<int>(nodes.size()) * 5);
  int row = 0;
  for (int node : nodes) {
    const int phase = fo.PhaseOfNode(node);
    const int ee_node = mo.PolyAtStartOfPhase(phase);      // GetNodeIDAtStartOfPhase: start node of that polynomial
    ItemRec it = NewItem(kItemForceNode);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + row;
    it.s[0] = fo.node_slot0 + node;
    it.aux = mo.node_slot0 + ee_node;
    it.gslot = AllocG(25);
    hp_.items.push_back(it);
    const int bk[5] = {kBoundForceNormal, kBoundLE0, kBoundGE0, kBoundLE0, kBoundGE0};
    for (int i = 0; i < 5; ++i) Bound(row + i, bk[i], ee);
    for (int dim = 0; dim < 3; ++dim) {
      const int idx = fo.OptIndex(node, 0, dim);
      if (idx < 0) throw std::runtime_error("force constraint touches a node value that is not optimised");
      for (int i = 0; i < 5; ++i) { Term tm; tm.g = it.gslot + 3 * i + dim; tm.b = one; tm.code = kCodeForce + 3 * i + dim; Add(row + i, fo.col0 + idx, tm, true); }
    }
    for (int dim = 0; dim < 2; ++dim) {
      const int idx = mo.OptIndex(ee_node, 0, dim);
      if (idx < 0) throw std::runtime_error("force constraint touches a foothold that is not optimised");
      for (int i = 0; i < 5; ++i) { Term tm; tm.g = it.gslot + 15 + 2 * i + dim; tm.b = one; Add(row + i, mo.col0 + idx, tm, true); }
    }
    row += 5;
  }
  EndSet();
}

// TorqueConstraintDiscretized (torque_constraint_discretized.cc:103-235): 4 rows per sample; G slots t1, t2, n, -n,
// (-k mu) n.  The ee-motion block exists only where the terrain has curvature (same rule as the force constraint).
void Builder::EmitTorqueDisc(int ee) {
  const VarSet& mo = Set(kEEMotion, ee);
  const VarSet& fo = Set(kEEForce, ee);
  const VarSet& to = Set(kEETorque, ee);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_torque);
  // curvature classes: S[dim][4] and (optimised durations) W_p[n_cols] after the existing slots, as in EmitForceDisc
  const int nd = CurvDimCount(hp_.curvature_dims);
  const int off_S = TorqueCurvOffset(rt() ? to.n_vars + fo.n_vars : 0, rt() ? Set(kEESchedule, ee).n_vars : 0);
  const int per_item = (off_S + 4 * nd + ((rt() && nd) ? mo.n_vars : 0)) | 1;
  const int chunk = std::max(1, 960 / per_item);
  const int bk_friction = AddB(s_.torque_k_friction);
  auto dim_slot = [&](int dim) { return dim == 0 ? 0 : (hp_.curvature_dims & 1 ? 1 : 0); };
  for (size_t k = 0; k < ts.size(); ++k) {
    const int r0 = static_cast<int>(k) * 4;
    if (k % chunk == 0) {
      if (k == 0) { BeginGroup(); BeginSet("torque-disc-" + to.name, static_cast<int>(ts.size()) * 4); }
      else BeginGroup(r0);
    }
    ItemRec it = NewItem(kItemTorque);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + r0;
    it.gslot = AllocG(per_item);
    it.b0 = bk_friction;
    it.sub = static_cast<short>(hp_.curvature_dims);
    Bound(r0 + 0, kBoundTorqueX, ee);
    Bound(r0 + 1, kBoundTorqueY, ee);
    Bound(r0 + 2, kBoundLE0, ee);
    Bound(r0 + 3, kBoundLE0, ee);
    if (rt()) {
      it.s[kSMotion + ee] = AtRt(0, ee, ts[k]);
      it.s[kSForce + ee] = AtRt(1, ee, ts[k]);
      it.s[kSTorque + ee] = AtRt(2, ee, ts[k]);
      hp_.items.push_back(it);
      const int wT = it.gslot + 15, wF = wT + to.n_vars;
      for (auto& c : AllCols(to))
        for (int i = 0; i < 4; ++i) Add(r0 + i, to.col0 + c.first, GG(it.gslot + 3 * i + c.second, wT + c.first));
      for (auto& c : AllCols(fo))
        for (int i = 2; i < 4; ++i) Add(r0 + i, fo.col0 + c.first, GG(it.gslot + 12 + c.second, wF + c.first));
      AddScheduleBlock(r0, 4, ee, wF + fo.n_vars);
      for (auto& c : AllCols(mo))       // torque_constraint_discretized.cc:176-197
        if (c.second < 2 && (hp_.curvature_dims >> c.second) & 1)
          for (int i = 0; i < 4; ++i)
            Add(r0 + i, mo.col0 + c.first, GG(it.gslot + off_S + 4 * dim_slot(c.second) + i, it.gslot + off_S + 4 * nd + c.first));
      continue;
    }
    SplineAt am = At(mo, ts[k]), af = At(fo, ts[k]), at = At(to, ts[k]);
    it.s[kSMotion + ee] = am.sample;
    it.s[kSForce + ee] = af.sample;
    it.s[kSTorque + ee] = at.sample;
    hp_.items.push_back(it);
    for (const ColBasis& c : Columns(to, at.poly, at.t_local))
      for (int i = 0; i < 4; ++i) {              // t1, t2, n, -n . J_tau
        Term tm; tm.g = it.gslot + 3 * i + c.dim; tm.b = AddB(c.v[0]); tm.code = kCodeTorque + 3 * i + c.dim;
        Add(r0 + i, to.col0 + c.idx, tm);
      }
    for (const ColBasis& c : Columns(fo, af.poly, af.t_local))
      for (int i = 2; i < 4; ++i) {              // (-k mu) n . J_f on the two yaw-friction rows
        Term tm; tm.g = it.gslot + 12 + c.dim; tm.b = AddB(c.v[0]); tm.code = kCodeTorque + 12 + c.dim;
        Add(r0 + i, fo.col0 + c.idx, tm);
      }
    for (const ColBasis& c : Columns(mo, am.poly, am.t_local))
      if (c.dim < 2 && (hp_.curvature_dims >> c.dim) & 1)
        for (int i = 0; i < 4; ++i) {
          Term tm; tm.g = it.gslot + off_S + 4 * dim_slot(c.dim) + i; tm.b = AddB(c.v[0]);
          Add(r0 + i, mo.col0 + c.idx, tm);
        }
  }
  EndSet();
}

// BaseMotionConstraint (base_motion_constraint.cc:38-85): rows AX..AZ <- base-ang position, LX..LZ <- base-lin position;
// bounds are NoBound unless Parameters sets them (nlp_formulation.cc, not in the default set-up)
void Builder::EmitBaseRom() {
  const VarSet& lin = Set(kBaseLin);
  const VarSet& ang = Set(kBaseAng);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_base_motion);
  BeginGroup();
  BeginSet("baseMotion", static_cast<int>(ts.size()) * 6);
  for (size_t k = 0; k < ts.size(); ++k) {
    const int r0 = static_cast<int>(k) * 6;
    ItemRec it = NewItem(kItemBaseRom);
    it.row[0] = set_row0_ + r0;
    SplineAt al = At(lin, ts[k]), aa = At(ang, ts[k]);
    it.s[kSBaseLin] = al.sample;
    it.s[kSBaseAng] = aa.sample;
    hp_.items.push_back(it);
    for (int d = 0; d < 6; ++d) Bound(r0 + d, kBoundNone);
    for (const ColBasis& c : Columns(ang, aa.poly, aa.t_local)) {
      Term tm; tm.g = kGOne; tm.b = AddB(c.v[0]);
      Add(r0 + c.dim, ang.col0 + c.idx, tm);
    }
    for (const ColBasis& c : Columns(lin, al.poly, al.t_local)) {
      Term tm; tm.g = kGOne; tm.b = AddB(c.v[0]);
      Add(r0 + 3 + c.dim, lin.col0 + c.idx, tm);
    }
  }
  EndSet();
}

// EELinearConstraint (ee_linear_constraint.cc:5-48, MakeEELinearConstraint nlp_formulation.cc:682-704), definition `def`:
// one row per sample of its own dt, g = sum_j coeff_j * spline_{ee_j}(t).at(deriv)(dim_j), bounds +- tolerance; the row of
// the Jacobian is sum_j coeff_j * (row dim_j of the foot's GetJacobianWrtNodes(t, deriv)), accumulated term by term
// (`jac.row(k) += ...`): with class-constant durations every entry is a class constant; with optimised durations the item
// finishes the columns of the active polynomials from run-time Hermite weights (ItemEELinear).
void Builder::EmitEELinear(int def) {
  const int n_terms = s_.ee_linear_n_terms[def], deriv = s_.ee_linear_deriv[def];
  const bool angle = s_.ee_linear_target[def] == 1;                   // s.ee_ang_ instead of s.ee_motion_ (nlp_formulation.cc:692-699)
  const int set_kind = angle ? kEEAng : kEEMotion;
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.ee_linear_dt[def]);
  const int one = AddB(1.0);
  // optimised durations: one block of G slots per foot the definition names (every column of the foot's set), offsets
  // within the item; the row stores, per term, the set's columns of the term's dimension (the pre-filled structure of
  // PhaseSpline, phase_spline.cc:45-51, row dim of GetJacobianWrtNodes)
  std::vector<int> foot_off(kMaxEE, -1);
  int per_item_rt = 0;
  for (int j = 0; j < n_terms; ++j) {
    const int ee = s_.ee_linear_term_ee[def][j];
    if (foot_off[ee] < 0) { foot_off[ee] = per_item_rt; per_item_rt += Set(set_kind, ee).n_vars; }
  }
  per_item_rt |= 1;
  const int chunk = std::max(1, 960 / (rt() ? per_item_rt : 1));
  for (size_t k = 0; k < ts.size(); ++k) {
    if (k == 0) { BeginGroup(); BeginSet("ee-linear", static_cast<int>(ts.size())); }
    else if (rt() && k % chunk == 0) BeginGroup(static_cast<int>(k));   // a new group of items: bounded G use per group
    ItemRec it = NewItem(kItemEELinear);
    it.row[0] = set_row0_ + static_cast<int>(k);
    it.sub = static_cast<short>(n_terms);
    Bound(static_cast<int>(k), kBoundEELinear, 0, def);
    std::vector<double> brun(4 * static_cast<size_t>(n_terms));
    std::vector<std::pair<int, double>> row;               // column -> value, in order of first appearance
    if (rt()) it.gslot = AllocG(per_item_rt);
    for (int j = 0; j < n_terms; ++j) {
      const int ee = s_.ee_linear_term_ee[def][j], dim = s_.ee_linear_term_dim[def][j];
      const double coeff = s_.ee_linear_term_coeff[def][j];
      const VarSet& mo = Set(set_kind, ee);
      brun[4 * j] = coeff;
      brun[4 * j + 1] = dim;
      brun[4 * j + 2] = deriv;
      brun[4 * j + 3] = rt() ? foot_off[ee] : 0;
      if (rt()) {
        it.s[j] = AtRt(angle ? 3 : 0, ee, ts[k]);
        for (auto& c : AllCols(mo)) {
          if (c.second != dim) continue;
          bool found = false;
          for (auto& e : row) found = found || e.first == mo.col0 + c.first;
          if (!found) row.emplace_back(mo.col0 + c.first, static_cast<double>(it.gslot + foot_off[ee] + c.first));
        }
        continue;
      }
      const SplineAt am = At(mo, ts[k]);
      it.s[j] = am.sample;
      for (const ColBasis& c : Columns(mo, am.poly, am.t_local)) {
        if (c.dim != dim) continue;
        const int col = mo.col0 + c.idx;
        const double v = coeff * c.v[deriv];
        bool found = false;
        for (auto& e : row)
          if (e.first == col) { e.second = e.second + v; found = true; }
        if (!found) row.emplace_back(col, 0.0 + v);
      }
    }
    it.b0 = AddBRun(brun);
    hp_.items.push_back(it);
    for (const auto& e : row) {
      Term tm;
      if (rt()) { tm.g = static_cast<int>(e.second); tm.b = one; }      // the item's finished value of that column
      else { tm.g = kGOne; tm.b = AddB(e.second); }
      Add(static_cast<int>(k), e.first, tm);
    }
  }
  EndSet();
}

// SwingConstraint (swing_constraint.cc:54-108): per swing node and dim in {x,y} a position and a velocity row
void Builder::EmitSwing(int ee) {
  const VarSet& mo = Set(kEEMotion, ee);
  const std::vector<int> nodes = mo.NonConstantNodes();
  const double t_swing_avg = 0.3;                // swing_constraint.h:68
  BeginGroup();
  BeginSet("swing-" + mo.name, static_cast<int>(nodes.size()) * 4);
  int row = 0;
  for (int node : nodes) {
    ItemRec it = NewItem(kItemSwingNode);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + row;
    it.s[0] = mo.node_slot0 + node;
    hp_.items.push_back(it);
    for (int dim = 0; dim < 2; ++dim) {
      auto put = [&](int r, int n, int deriv, double v) {
        int idx = mo.OptIndex(n, deriv, dim);
        if (idx < 0) throw std::runtime_error("swing constraint touches a node value that is not optimised");
        Term tm; tm.g = kGOne; tm.b = AddB(v);
        Add(r, mo.col0 + idx, tm, true);
      };
      put(row, node, 0, 1.0);
      put(row, node + 1, 0, -0.5);
      put(row, node - 1, 0, -0.5);
      row++;
      put(row, node, 1, 1.0);
      put(row, node + 1, 0, -1.0 / t_swing_avg);
      put(row, node - 1, 0, +1.0 / t_swing_avg);
      row++;
    }
  }
  EndSet();
}

// TerrainConstraintHard (terrain_constraint_hard.cc:34-132), one row per sample of dt_constraint_range_of_motion_
// (nlp_formulation.cc:476-488): g = p_z - h(p) - min(k |v_tangent|, k).  Only the ee-motion block exists (the reference
// gives no ee-schedule block).  A row stores the union of J_pos rows Z, X, Y and - structurally, whether or not the
// velocity term is active - the same columns of J_vel: every variable of the active polynomial's two nodes (all variables
// of the set with optimised durations).  The value of a column of dimension d is
//   c_d J_pos - (k td_d) J_vel,  c_Z = 1, c_X = -h_x, c_Y = -h_y,  td = (v.t1 t1 + v.t2 t2) / |v_tangent| or 0,
// two products, so the item finishes it: fixed durations read (J_pos, J_vel, d) per column from a B run, optimised
// durations accumulate the run-time Hermite weights (ItemTerrainHard, items.cuh).  Entries are G[slot] * 1.
void Builder::EmitTerrainHard(int ee) {
  const VarSet& mo = Set(kEEMotion, ee);
  const std::vector<double> ts = SampleTimes(hp_.total_time, s_.dt_constraint_range_of_motion);
  const int one = AddB(1.0);
  const int per_item_rt = (3 * mo.n_vars) | 1;            // finished values, then scratch: pos and vel weight sums
  const int chunk = std::max(1, 960 / (rt() ? per_item_rt : 12));
  for (size_t k = 0; k < ts.size(); ++k) {
    if (k % chunk == 0) {
      if (k == 0) { BeginGroup(); BeginSet("terrainhard-" + std::to_string(ee), static_cast<int>(ts.size())); }
      else BeginGroup(static_cast<int>(k));
    }
    ItemRec it = NewItem(kItemTerrainHard);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + static_cast<int>(k);
    Bound(static_cast<int>(k), kBoundBaseHeight);          // (0, 1e20), terrain_constraint_hard.cc:72-77
    if (rt()) {
      it.gslot = AllocG(per_item_rt);
      it.s[kSMotion + ee] = AtRt(0, ee, ts[k]);
      it.sub = 0;
      hp_.items.push_back(it);
      for (auto& c : AllCols(mo)) {
        Term tm; tm.g = it.gslot + c.first; tm.b = one;
        Add(static_cast<int>(k), mo.col0 + c.first, tm);
      }
      continue;
    }
    SplineAt am = At(mo, ts[k]);
    it.s[kSMotion + ee] = am.sample;
    const std::vector<ColBasis> cols = Columns(mo, am.poly, am.t_local);
    if (cols.size() > 12) throw std::runtime_error("unsupported: more than 12 ee-motion columns at one sample");
    std::vector<double> run;
    for (const ColBasis& c : cols) { run.push_back(c.v[0]); run.push_back(c.v[1]); run.push_back(static_cast<double>(c.dim)); }
    it.b0 = AddBRun(run);
    it.sub = static_cast<short>(cols.size());
    it.gslot = AllocG(static_cast<int>(cols.size()) | 1);
    hp_.items.push_back(it);
    for (size_t j = 0; j < cols.size(); ++j) {
      Term tm; tm.g = it.gslot + static_cast<int>(j); tm.b = one;
      Add(static_cast<int>(k), mo.col0 + cols[j].idx, tm);
    }
  }
  EndSet();
}

// Node-based TorqueConstraint (torque_constraint.cc:34-159), three rows per stance torque node: tau . t1, tau . t2, tau . n
// with the basis at the foothold of the node's phase.  ee-torque block: the basis components (:121-133); ee-motion block:
// tau_s . d(t1, t2, n)/d(x, y) with tau_s the torque at the START of the phase (:141), stored whatever its value; no
// ee-force block (:155-157).  Rows 0, 1 bounded by the torque limits, row 2 by +- k mu f_n (kBoundTorqueNodeZ).
void Builder::EmitTorqueNodes(int ee) {
  const VarSet& mo = Set(kEEMotion, ee);
  const VarSet& to = Set(kEETorque, ee);
  const std::vector<int> nodes = to.NonConstantNodes();
  const int one = AddB(1.0);
  BeginGroup();
  BeginSet("torque-" + to.name, static_cast<int>(nodes.size()) * 3);
  int row = 0;
  for (int node : nodes) {
    const int phase = to.PhaseOfNode(node);
    const int ee_node = mo.PolyAtStartOfPhase(phase);
    ItemRec it = NewItem(kItemTorqueNode);
    it.ee = static_cast<short>(ee);
    it.row[0] = set_row0_ + row;
    it.s[0] = to.node_slot0 + node;
    it.aux = mo.node_slot0 + ee_node;
    it.b0 = to.node_slot0 + to.PolyAtStartOfPhase(phase);  // node slot, not a B run (pack.cc passes it through)
    it.gslot = AllocG(15);
    hp_.items.push_back(it);
    Bound(row + 0, kBoundTorqueX, ee);
    Bound(row + 1, kBoundTorqueY, ee);
    Bound(row + 2, kBoundTorqueNodeZ, ee, phase);
    for (int dim = 0; dim < 3; ++dim) {
      const int idx = to.OptIndex(node, 0, dim);
      if (idx < 0) throw std::runtime_error("torque constraint touches a node value that is not optimised");
      for (int i = 0; i < 3; ++i) { Term tm; tm.g = it.gslot + 3 * i + dim; tm.b = one; tm.code = kCodeTorque + (i == 2 ? 6 : 3 * i) + dim; Add(row + i, to.col0 + idx, tm, true); }
    }
    for (int dim = 0; dim < 2; ++dim) {
      const int idx = mo.OptIndex(ee_node, 0, dim);
      if (idx < 0) throw std::runtime_error("torque constraint touches a foothold that is not optimised");
      for (int i = 0; i < 3; ++i) { Term tm; tm.g = it.gslot + 9 + 2 * i + dim; tm.b = one; Add(row + i, mo.col0 + idx, tm, true); }
    }
    row += 3;
  }
  EndSet();
}

void Builder::Run() {
  hp_.n_rows = 0;
  for (int i = 0; i < s_.n_constraints; ++i) {
    switch (s_.constraints[i]) {
      case TOWR_B200_C_TERRAIN:
        for (int e = 0; e < s_.n_ee; ++e) EmitTerrain(e);
        break;
      case TOWR_B200_C_DYNAMIC: EmitDynamic(); break;
      case TOWR_B200_C_BASE_ACC: EmitSplineAcc(); break;
      case TOWR_B200_C_ENDEFFECTOR_ROM: {
        const int K = static_cast<int>(SampleTimes(hp_.total_time, s_.dt_constraint_range_of_motion).size());
        if (rt() || K * (kRomPreSize + 9 * s_.n_ee) > 6000) EmitRomSplit();
        else EmitRom();
        break;
      }
      case TOWR_B200_C_FORCE:
        for (int e = 0; e < s_.n_ee; ++e) {
          if (s_.dt_constraint_force > 0.0) EmitForceDisc(e);
          else EmitForceNodes(e);
        }
        break;
      case TOWR_B200_C_SWING:
        for (int e = 0; e < s_.n_ee; ++e) EmitSwing(e);
        break;
      case TOWR_B200_C_BASE_HEIGHT: EmitBaseHeight(); break;
      case TOWR_B200_C_TORQUE:
        for (int e = 0; e < s_.n_ee; ++e) {
          if (s_.dt_constraint_torque > 0.0) EmitTorqueDisc(e);
          else EmitTorqueNodes(e);                          // nlp_formulation.cc:543-567
        }
        break;
      case TOWR_B200_C_BASE_ROM: EmitBaseRom(); break;
      case TOWR_B200_C_TOTAL_TIME:
        if (!rt()) throw std::runtime_error("TotalTime constraint without ee-schedule variables");
        for (int e = 0; e < s_.n_ee; ++e) EmitTotalDuration(e);
        break;
      case TOWR_B200_C_TERRAIN_HARD:
        for (int e = 0; e < s_.n_ee; ++e) EmitTerrainHard(e);
        break;
      default: throw std::runtime_error("constraint not defined!");   // nlp_formulation.cc:396
    }
  }
  for (int i = 0; i < s_.n_ee_linear; ++i) EmitEELinear(i);          // nlp_formulation.cc:373-375
  CloseSegment();
  hp_.nnz = static_cast<int>(hp_.entries.size());
}

void ResetTables(HostProgram& hp) {
  hp.con_sets.clear(); hp.pat_rows.clear(); hp.pat_cols.clear(); hp.row_bounds.clear();
  hp.samples.clear(); hp.items.clear(); hp.entries.clear(); hp.entry_code.clear(); hp.B.clear(); hp.segments.clear();
  hp.g_capacity = 0; hp.n_rows = 0; hp.nnz = 0;
}

}  // namespace

std::string StructureSignature(const towr_b200_spec& s, int curvature_dims) {
  // every field the tables above depend on, printed exactly
  std::ostringstream o;
  o.precision(17);
  o << "E" << s.n_ee << "|curv" << curvature_dims;
  for (int e = 0; e < s.n_ee; ++e) {
    o << "|f" << e << ":" << s.in_contact_at_start[e] << ":";
    for (int i = 0; i < s.n_phases[e]; ++i) o << s.phase_durations[e][i] << ",";
  }
  o << "|c";
  for (int i = 0; i < s.n_constraints; ++i) o << s.constraints[i] << ",";
  o << "|dt" << s.dt_constraint_dynamic << "," << s.dt_constraint_range_of_motion << "," << s.dt_constraint_base_motion
    << "," << s.dt_constraint_force << "," << s.dt_constraint_torque << "|bp" << s.duration_base_polynomial
    << "|pp" << s.ee_polynomials_per_swing_phase << "," << s.force_polynomials_per_stance_phase << ","
    << s.torque_polynomials_per_stance_phase;
  for (int i = 0; i < s.n_constraints; ++i)
    if (s.constraints[i] == TOWR_B200_C_TORQUE) o << "|kf" << s.torque_k_friction;   // baked into the class tables
  for (int c = 0; c < s.n_costs; ++c) {                                                            // class-level cost tables
    o << "|cost" << s.cost_ids[c] << ":" << s.cost_weights[c];
    if (s.cost_ids[c] == TOWR_B200_COST_ENERGY) o << ":" << s.dt_cost_energy << ":" << s.energy_cost_torque_weight;
    if (s.cost_ids[c] == TOWR_B200_COST_ANG_MOM) o << ":" << s.dt_cost_ang_mom;
  }
  for (int i = 0; i < s.n_ee_linear; ++i) {
    o << "|eelin" << s.ee_linear_target[i] << ":" << s.ee_linear_deriv[i] << ":" << s.ee_linear_tolerance[i] << ":" << s.ee_linear_dt[i];
    for (int j = 0; j < s.ee_linear_n_terms[i]; ++j)
      o << ":" << s.ee_linear_term_ee[i][j] << "," << s.ee_linear_term_dim[i][j] << "," << s.ee_linear_term_coeff[i][j];
  }
  for (int c = 0; c < s.n_node_costs; ++c)
    o << "|nc" << s.node_cost_set[c] << ":" << (s.node_cost_set[c] < 2 ? -1 : s.node_cost_ee[c]) << ":" << s.node_cost_deriv[c] << ":"
      << s.node_cost_dim[c] << ":" << s.node_cost_weight[c];
  if (s.enable_swing_ee_base_pos_tracking && s.swing_ee_base_pos_tracking_weight > 0.0)
    o << "|swingpos" << s.swing_ee_base_pos_tracking_weight << ":" << s.dt_cost_swing_ee_base_pos_tracking;
  return o.str();
}

void BuildProgram(const towr_b200_spec& s, HostProgram& hp, int g_cap) {
  BuildLayout(s, hp);
  hp.signature = StructureSignature(s, hp.curvature_dims);
  // pass 1: one segment per group, to learn each group's G footprint
  ResetTables(hp);
  const int rom_fpg = 0;
  Builder sizing(s, hp, nullptr);
  sizing.rom_feet_per_group = rom_fpg;
  sizing.Run();
  std::vector<int> group_entries;                          // one segment per group in this pass
  for (const SegmentRec& sg : hp.segments) group_entries.push_back(sg.n_entries);
  group_entries.resize(sizing.group_g_size.size(), 0);
  // G slots per segment.  Every segment costs the kernel a fixed stage-A latency, so use as few segments as the
  // 16-bit byte offsets of the entry descriptors allow (8191 slots) and balance them.  TOWR_B200_GCAP overrides.
  const int cap_max = 8000;
  int total = 0, largest = 0;
  for (int g : sizing.group_g_size) { total += g; largest = std::max(largest, g); }
  auto plan_with = [&](int cap, std::vector<char>& plan) {
    int used = kGReserved, n = 0;
    for (size_t g = 0; g < plan.size(); ++g) {
      plan[g] = 0;
      if (g == 0 || used + sizing.group_g_size[g] > cap) {
        plan[g] = 1;
        used = kGReserved;
        ++n;
      }
      used += sizing.group_g_size[g];
    }
    return n;
  };
  std::vector<char> plan(sizing.group_g_size.size(), 0);
  if (g_cap > 0) {
    plan_with(std::max(largest + kGReserved, g_cap), plan);
  } else if (const char* env = std::getenv("TOWR_B200_GCAP")) {
    plan_with(std::max(64, std::atoi(env)), plan);
  } else {
    const int n_min = plan_with(cap_max, plan);            // fewest segments possible
    int cap = std::max(largest + kGReserved, total / std::max(1, n_min));
    while (cap < cap_max && plan_with(cap, plan) > n_min) cap += 16;   // smallest cap that still gives n_min: balanced
    if (cap >= cap_max) plan_with(cap_max, plan);
  }
  // pass 2: the real tables
  ResetTables(hp);
  Builder real(s, hp, &plan);
  real.rom_feet_per_group = rom_fpg;
  real.Run();
  if (std::getenv("TOWR_B200_DEBUG")) {
    std::fprintf(stderr, "groups:");
    for (size_t g = 0; g < sizing.group_g_size.size(); ++g) std::fprintf(stderr, " %d/%d", sizing.group_g_size[g], group_entries[g]);
    std::fprintf(stderr, "\n");
  }
  PackProgram(hp);
  if (std::getenv("TOWR_B200_DEBUG"))
    for (const SegmentRec& sg : hp.segments)
      std::fprintf(stderr, "segment items %d (level 0: %d, padded) entries %d g %d\n", sg.n_items, sg.n_items0, sg.n_entries, sg.g_size);
  if (std::getenv("TOWR_B200_DEBUG"))
    std::fprintf(stderr, "blob %d B (B %d, times %d, items %d, samples %d, polys %d), G capacity %d, %zu segments\n", hp.blob.bytes,
                 hp.blob.n_B, hp.blob.n_times, hp.blob.n_items, hp.blob.n_samples, hp.blob.n_polys, hp.g_capacity, hp.segments.size());
}

}  // namespace tb

// cost.cc — host tables of the two cost kernels (cost.h): the NodeCost weight table and the blob of the time-discretised terms.
#include "cost.h"
#include "structure.h"

#include <algorithm>
#include <cstring>
#include <numeric>
#include <stdexcept>

namespace tb {

void BuildCostTables(const HostProgram& hp, const towr_b200_spec& s, CostTables& o) {
  o = CostTables{};
  o.n_vars = hp.n_vars;
  struct Term { int kind, ee, deriv, dim; double w; };
  std::vector<Term> terms;
  for (int c = 0; c < s.n_costs; ++c) {
    const double weight = s.cost_weights[c];
    if (s.cost_ids[c] == TOWR_B200_COST_FORCES) {            // MakeForcesCost, nlp_formulation.cc:648-666
      for (int ee = 0; ee < hp.n_ee; ++ee)
        for (int dim = 0; dim < 3; ++dim) {
          terms.push_back({kEEForce, ee, 0, dim, weight});
          terms.push_back({kEETorque, ee, 0, dim, weight});
          terms.push_back({kEEForce, ee, 1, dim, 0.1 * weight});
          terms.push_back({kEETorque, ee, 1, dim, 0.1 * weight});
        }
    } else if (s.cost_ids[c] == TOWR_B200_COST_EE_MOTION) {  // MakeEEMotionCost, nlp_formulation.cc:668-681
      for (int ee = 0; ee < hp.n_ee; ++ee) {
        terms.push_back({kEEMotion, ee, 1, 0, weight});
        terms.push_back({kEEMotion, ee, 1, 1, weight});
        terms.push_back({kEEMotion, ee, 1, 2, 0.5 * weight});
      }
    } else if (s.cost_ids[c] == TOWR_B200_COST_ENERGY || s.cost_ids[c] == TOWR_B200_COST_ANG_MOM) {
      // time-discretised terms: BuildSampledCost
    } else {
      throw std::runtime_error("cost not defined!");         // nlp_formulation.cc:644
    }
  }
  for (int c = 0; c < s.n_node_costs; ++c)                    // individual NodeCost(nodes_id, deriv, dim, weight) terms
    terms.push_back({s.node_cost_set[c], s.node_cost_set[c] < kEEMotion ? -1 : s.node_cost_ee[c], s.node_cost_deriv[c],
                     s.node_cost_dim[c], s.node_cost_weight[c]});
  o.n_terms = static_cast<int>(terms.size());
  std::vector<std::vector<double>> per_col(static_cast<size_t>(hp.n_vars));
  for (const Term& t : terms) {
    const VarSet* v = hp.Find(t.kind, t.ee);
    if (!v) continue;
    for (int idx = 0; idx < v->n_vars; ++idx)
      for (const NodeRef& r : v->refs[idx])                 // NodeCost::FillJacobianBlock, node_cost.cc:68-74
        if (r.deriv == t.deriv && r.dim == t.dim) per_col[v->col0 + idx].push_back(t.w);
  }
  o.col_ptr.assign(1, 0);
  for (const auto& ws : per_col) {
    o.weights.insert(o.weights.end(), ws.begin(), ws.end());
    o.col_ptr.push_back(static_cast<int>(o.weights.size()));
  }
}

namespace {
template <class T>
int Append(std::vector<unsigned char>& blob, const std::vector<T>& v) {
  while (blob.size() % 8) blob.push_back(0);
  const int off = static_cast<int>(blob.size());
  const unsigned char* p = reinterpret_cast<const unsigned char*>(v.data());
  blob.insert(blob.end(), p, p + sizeof(T) * v.size());
  return off;
}
}  // namespace

std::vector<SampledTermDef> SampledTermDefs(const HostProgram& hp, const towr_b200_spec& s) {
  std::vector<SampledTermDef> defs;
  for (int c = 0; c < s.n_costs; ++c) {
    if (s.cost_ids[c] == TOWR_B200_COST_ENERGY) defs.push_back({kSampledEnergy, s.cost_weights[c], s.dt_cost_energy, -1});
    if (s.cost_ids[c] == TOWR_B200_COST_ANG_MOM) defs.push_back({kSampledAngMom, s.cost_weights[c], s.dt_cost_ang_mom, -1});
  }
  if (HasSwingTracking(s))                                     // after the listed costs (nlp_formulation.cc:612-626)
    for (int e = 0; e < hp.n_ee; ++e)
      defs.push_back({kSampledEEBasePos, s.swing_ee_base_pos_tracking_weight, s.dt_cost_swing_ee_base_pos_tracking, e});
  std::vector<SampledTermDef> live;
  for (const SampledTermDef& d : defs)
    if (d.weight > 0.0) live.push_back(d);                     // GetCost / FillJacobianBlock return at once otherwise (:70, :94)
  return live;
}

void BuildSampledCostParts(const HostProgram& hp, const towr_b200_spec& s, std::vector<SampledCost>& out) {
  out.clear();
  const std::vector<SampledTermDef> defs = SampledTermDefs(hp, s);
  const int n = static_cast<int>(defs.size());
  int first = 0;
  while (first < n) {
    SampledCost part;
    int count = n - first;
    BuildSampledCost(hp, s, defs, first, count, part, count > 1);
    while (count > 1 && part.shared_bytes > kSampledTwoCtaShared) {   // the longest run of terms that leaves room for two CTAs
      --count;
      BuildSampledCost(hp, s, defs, first, count, part, count > 1);
    }
    out.push_back(std::move(part));
    first += count;
  }
}

void BuildSampledCost(const HostProgram& hp, const towr_b200_spec& s, const std::vector<SampledTermDef>& all_defs, int first,
                      int count, SampledCost& out, bool size_only_ok) {
  out = SampledCost{};
  std::vector<SampledTerm> terms;
  std::vector<SampledSpline> splines;
  std::vector<SampledFoot> feet;
  std::vector<SampledJob> jobs;
  std::vector<uint16_t> node_src, poly_spline, fix_seg;
  std::vector<double> poly_T, times, fix_t;
  std::vector<uint32_t> poly_meta;
  int n_nodes = 0, n_polys = 0, n_phases = 0, n_phase_slots = 0;
  const bool rt = hp.optimize_durations;
  const VarSet* lin = hp.Find(kBaseLin);
  // Spline::GetTotalTime of base_linear_ (helpers/spline.cc:119-123)
  const double T_total = std::accumulate(lin->poly_T.begin(), lin->poly_T.end(), 0.0);
  auto sample_times = [&](double dt) {      // GetSampleTimes (energy_cost.cc:52-66, angular_momentum_cost.cc:51-65)
    std::vector<double> ts;
    if (dt <= 0.0) return std::vector<double>{0.0, T_total};
    for (double t = 0.0; t <= T_total + 1e-9; t += dt) {
      ts.push_back(t);
      if (ts.size() > 100000) throw std::runtime_error("unsupported: cost sampling period too small");
    }
    return ts;
  };
  auto add_spline = [&](const VarSet& v, int foot, bool is_torque, double scale, double mscale) {
    SampledSpline r{};
    r.node0 = n_nodes;
    r.poly0 = n_polys;
    r.n_polys = v.n_polys();
    r.foot = foot;
    r.rt = (foot >= 0 && rt) ? 1 : 0;
    r.is_torque = is_torque ? 1 : 0;
    r.ps0 = n_phase_slots;
    if (r.rt) n_phase_slots += s.n_phases[foot];
    r.scale = scale;
    r.mscale = mscale;
    std::vector<int> src(static_cast<size_t>(v.n_nodes) * 6, 0xFFFF);
    for (int idx = 0; idx < v.n_vars; ++idx)
      for (const NodeRef& nr : v.refs[idx]) src[nr.node * 6 + nr.deriv * 3 + nr.dim] = v.col0 + idx;
    for (int x : src) node_src.push_back(static_cast<uint16_t>(x));
    for (int p = 0; p < v.n_polys(); ++p) {
      poly_T.push_back(v.poly_T[p]);
      poly_spline.push_back(static_cast<uint16_t>(splines.size()));
      poly_meta.push_back(r.rt ? (static_cast<uint32_t>(v.poly_phase[p]) | (static_cast<uint32_t>(v.polys_in_phase[p]) << 8) |
                                  (static_cast<uint32_t>(v.poly_in_phase[p]) << 16))
                               : kPolyFixed);
    }
    n_nodes += v.n_nodes;
    n_polys += v.n_polys();
    splines.push_back(r);
  };
  auto node_jobs = [&](const VarSet& v, int spline_in_term, int term) {
    // one job per (node, dim): the variables that hold the node's position and velocity value there
    std::vector<int> col(static_cast<size_t>(v.n_nodes) * 6, -1);
    for (int idx = 0; idx < v.n_vars; ++idx) {
      if (v.refs[idx].size() != 1) throw std::runtime_error("unsupported: a cost variable that writes several node values");
      const NodeRef& nr = v.refs[idx][0];
      col[nr.node * 6 + nr.deriv * 3 + nr.dim] = v.col0 + idx;
    }
    for (int node = 0; node < v.n_nodes; ++node) {
      SampledJob j{};
      bool any = false;
      for (int i = 0; i < 6; ++i) { j.col[i] = col[node * 6 + i]; any = any || j.col[i] >= 0; }
      if (!any) continue;
      j.kind = static_cast<uint16_t>(kJobNode | (term << 8));
      j.spline = static_cast<uint16_t>(spline_in_term);
      j.node = static_cast<uint16_t>(node);
      jobs.push_back(j);
    }
  };
  // variables that write several node values (the stance variables of ee-motion): one job per column
  auto column_jobs = [&](const VarSet& v, int spline_in_term, int term) {
    for (int idx = 0; idx < v.n_vars; ++idx) {
      if (v.refs[idx].empty() || v.refs[idx].size() > 5) throw std::runtime_error("unsupported: cost variable with 0 or > 5 node values");
      SampledJob j{};
      j.col[0] = v.col0 + idx;
      j.kind = static_cast<uint16_t>(kJobColumn | (term << 8));
      j.spline = static_cast<uint16_t>(spline_in_term);
      j.aux = static_cast<uint16_t>(v.refs[idx].size());
      for (size_t r = 0; r < v.refs[idx].size(); ++r)
        j.col[1 + r] = v.refs[idx][r].node | (v.refs[idx][r].deriv << 16) | (v.refs[idx][r].dim << 20);
      jobs.push_back(j);
    }
  };
  const std::vector<SampledTermDef> defs(all_defs.begin() + first, all_defs.begin() + first + count);
  int n_items = 0, item_doubles = 0;
  for (const SampledTermDef& def : defs) {
    const double weight = def.weight;
    if (weight <= 0.0) continue;                               // GetCost / FillJacobianBlock return at once (:70, :94)
    const bool energy = def.kind == kSampledEnergy;
    const double dt = def.dt;
    SampledTerm t{};
    t.kind = def.kind;
    t.foot = def.foot;
    t.spline0 = static_cast<int>(splines.size());
    t.time0 = static_cast<int>(times.size());
    t.job0 = static_cast<int>(jobs.size());
    t.item0 = n_items;
    t.scratch0 = item_doubles;
    t.rec = energy ? (rt ? 10 : 4) : def.kind == kSampledAngMom ? 8 : 9;
    const int term = static_cast<int>(terms.size());
    t.weight = weight;
    t.dtf = dt > 0.0 ? dt : 1.0;
    const std::vector<double> ts = sample_times(dt);
    times.insert(times.end(), ts.begin(), ts.end());
    t.n_samples = static_cast<int>(ts.size());
    const double wdt = weight * t.dtf;
    if (energy) {
      const double tw = s.energy_cost_torque_weight;
      for (int e = 0; e < hp.n_ee; ++e) {
        const VarSet* vf = hp.Find(kEEForce, e);
        const VarSet* vt = hp.Find(kEETorque, e);
        add_spline(*vf, e, false, 1.0, 2.0 * wdt);
        add_spline(*vt, e, true, tw, 2.0 * wdt * tw);
        node_jobs(*vf, 2 * e, term);
        if (tw != 0.0) node_jobs(*vt, 2 * e + 1, term);             // energy_cost.cc:119
        if (rt) {                                              // id::EESchedule(ee) exists only then (:129-157)
          const VarSet* vs = hp.Find(kEESchedule, e);
          for (int ph = 0; ph < vs->n_vars; ++ph) {
            SampledJob j{};
            j.col[0] = vs->col0 + ph;
            j.kind = static_cast<uint16_t>(kJobSchedule | (term << 8));
            j.spline = static_cast<uint16_t>(2 * e);
            j.node = static_cast<uint16_t>(ph);
            j.aux = tw != 0.0 ? 1 : 0;
            jobs.push_back(j);
          }
        }
      }
    } else if (def.kind == kSampledAngMom) {
      const VarSet* va = hp.Find(kBaseAng);
      add_spline(*va, -1, false, 1.0, 2.0 * wdt);
      node_jobs(*va, 0, term);
    } else {
      // EEBasePosCost (ee_base_pos_cost.cc): the foot's motion spline, base-lin, base-ang; the gradient reaches their
      // node variables (:123-152), not the ee-schedule (:154-157)
      const VarSet* vm = hp.Find(kEEMotion, def.foot);
      const VarSet* vl = hp.Find(kBaseLin);
      const VarSet* va = hp.Find(kBaseAng);
      add_spline(*vm, def.foot, false, 1.0, 2.0 * weight);     // m = (2.0*weight_) * e  (:121)
      add_spline(*vl, -1, false, 1.0, 2.0 * weight);
      add_spline(*va, -1, false, 1.0, 2.0 * weight);
      if (vm->n_polys() >= 4096 || vl->n_polys() >= 4096) throw std::runtime_error("unsupported: more than 4095 polynomials in a spline");
      column_jobs(*vm, 0, term);
      node_jobs(*vl, 1, term);
      node_jobs(*va, 2, term);
    }
    t.n_splines = static_cast<int>(splines.size()) - t.spline0;
    t.item_splines = energy ? t.n_splines : 1;
    t.n_jobs = static_cast<int>(jobs.size()) - t.job0;
    // fixed durations: segment and local time of every (spline, sample) are class constants (spline.cc:48-78)
    for (int sidx = t.spline0; sidx < static_cast<int>(splines.size()); ++sidx) {
      SampledSpline& sp = splines[sidx];
      sp.fix0 = -1;
      if (sp.rt) continue;
      sp.fix0 = static_cast<int>(fix_t.size());
      const std::vector<double> d(poly_T.begin() + sp.poly0, poly_T.begin() + sp.poly0 + sp.n_polys);
      std::vector<double> pd;
      if (sp.foot >= 0) pd.assign(s.phase_durations[sp.foot], s.phase_durations[sp.foot] + s.n_phases[sp.foot]);
      for (double tg : ts) {
        const int id = SegmentId(tg, d);
        int contact = 0;
        if (sp.foot >= 0) {                                    // PhaseDurations::IsContactPhase (phase_durations.cc:113-118)
          const int ph = SegmentId(tg, pd);
          contact = ((ph % 2 == 0) ? s.in_contact_at_start[sp.foot] != 0 : s.in_contact_at_start[sp.foot] == 0) ? 1 : 0;
        }
        if (id >= 0x8000) throw std::runtime_error("unsupported: more than 32767 polynomials in a spline");
        fix_seg.push_back(static_cast<uint16_t>(id | (contact << 15)));
        fix_t.push_back(LocalTime(tg, id, d));
      }
    }
    n_items += t.item_splines * t.n_samples;
    item_doubles += t.rec * t.item_splines * t.n_samples;
    terms.push_back(t);
  }
  if (terms.empty()) return;
  // Columns that several jobs write (base-ang: AngularMomentumCost and every foot's EEBasePosCost; base-lin: every
  // foot's): those jobs park their sums and one thread per column adds them in job = term order, which is also the order
  // in which ifopt adds the terms' gradients.
  std::vector<int> comb_col, comb_ptr(1, 0), comb_src;
  int n_scratch = 0;
  {
    auto n_cols = [](const SampledJob& j) { return (j.kind & 0xFF) == kJobNode ? 6 : 1; };
    std::vector<int> writers(static_cast<size_t>(hp.n_vars), 0);
    for (const SampledJob& j : jobs)
      for (int i = 0; i < n_cols(j); ++i)
        if (j.col[i] >= 0) ++writers[j.col[i]];
    std::vector<std::vector<int>> src(static_cast<size_t>(hp.n_vars));
    for (SampledJob& j : jobs) {
      j.slot = -1;
      bool shared = false;
      for (int i = 0; i < n_cols(j); ++i) shared = shared || (j.col[i] >= 0 && writers[j.col[i]] > 1);
      if (!shared) continue;
      j.slot = n_scratch;
      for (int i = 0; i < n_cols(j); ++i)
        if (j.col[i] >= 0) src[j.col[i]].push_back(n_scratch + i);
      n_scratch += n_cols(j);
    }
    for (int c = 0; c < hp.n_vars; ++c) {
      if (src[c].empty()) continue;
      comb_col.push_back(c);
      comb_src.insert(comb_src.end(), src[c].begin(), src[c].end());
      comb_ptr.push_back(static_cast<int>(comb_src.size()));
    }
  }
  if (terms.size() > 8) throw std::runtime_error("unsupported: more than 8 time-discretised cost terms");   // one warp each (cost_kernel.cu)
  for (int e = 0; e < hp.n_ee; ++e) {
    SampledFoot f{};
    f.phase0 = n_phases;
    f.n_phases = s.n_phases[e];
    f.col0 = rt ? hp.Find(kEESchedule, e)->col0 : 0;
    f.contact_at_start = s.in_contact_at_start[e] != 0;
    for (int i = 0; i < s.n_phases[e]; ++i) f.total += s.phase_durations[e][i];   // phase_durations.cc:40-50
    feet.push_back(f);
    n_phases += s.n_phases[e];
  }
  SampledHeader h{};
  h.n_vars = hp.n_vars;
  h.n_ee = hp.n_ee;
  h.rt = rt ? 1 : 0;
  h.n_terms = static_cast<int>(terms.size());
  h.n_splines = static_cast<int>(splines.size());
  h.n_nodes = n_nodes;
  h.n_polys = n_polys;
  h.n_phases = n_phases;
  h.n_times = static_cast<int>(times.size());
  h.n_jobs = static_cast<int>(jobs.size());
  h.n_phase_slots = n_phase_slots;
  h.n_scratch = n_scratch;
  h.n_comb = static_cast<int>(comb_col.size());
  h.n_items = n_items;
  h.item_doubles = item_doubles;
  // body inertia as SingleRigidBodyDynamics builds it (single_rigid_body_dynamics.cc:36-44): off-diagonals negated
  const double* in = s.inertia;   // Ixx, Iyy, Izz, Ixy, Ixz, Iyz
  const double I[9] = {in[0], -in[3], -in[4], -in[3], in[1], -in[5], -in[4], -in[5], in[2]};
  std::memcpy(h.I, I, sizeof(I));
  SwingReferenceB(s, h.p_ref_B);
  std::vector<unsigned char>& blob = out.blob;
  blob.assign(sizeof(SampledHeader), 0);
  h.off_terms = Append(blob, terms);
  h.off_splines = Append(blob, splines);
  h.off_feet = Append(blob, feet);
  h.off_poly_T = Append(blob, poly_T);
  h.off_times = Append(blob, times);
  h.off_jobs = Append(blob, jobs);
  h.off_poly_meta = Append(blob, poly_meta);
  h.off_node_src = Append(blob, node_src);
  h.off_poly_spline = Append(blob, poly_spline);
  h.off_comb_col = Append(blob, comb_col);
  h.off_comb_ptr = Append(blob, comb_ptr);
  h.off_comb_src = Append(blob, comb_src);
  h.off_fix_t = Append(blob, fix_t);
  h.off_fix_seg = Append(blob, fix_seg);
  while (blob.size() % 16) blob.push_back(0);
  h.bytes = static_cast<int>(blob.size());
  std::memcpy(blob.data(), &h, sizeof(h));
  out.shared_bytes = SampledSharedBytes(h);
  if (out.shared_bytes > kSampledMaxShared && !size_only_ok)
    throw std::runtime_error("unsupported: the cost samples of this class exceed one SM's shared memory (larger dt_cost_*)");
}

}  // namespace tb

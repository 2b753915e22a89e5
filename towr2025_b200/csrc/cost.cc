// cost.cc — host tables of the cost kernel (cost.h).
#include "cost.h"

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
    } else {
      throw std::runtime_error("unsupported: EnergyCost / AngularMomentumCost are not built");
    }
  }
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

}  // namespace tb

// cost.h — cost terms of a structure class (NodeCost, costs/node_cost.cc) as a column-compressed weight table.
#pragma once
#include <vector>

#ifdef __CUDACC__
#include <cuda_runtime.h>
#endif

#include "program.h"
#include "../../include/towr_b200.h"

namespace tb {

struct CostTables {              // host side
  int n_vars = 0, n_terms = 0;
  std::vector<int> col_ptr;      // n_vars + 1
  std::vector<double> weights;   // one per (variable, node value) pair of a cost term, grouped by column, term order kept
};
struct CostDevice {
  int n_vars;
  const int* col_ptr;
  const double* weights;
};

// NlpFormulation::GetCosts over params_.costs_ (nlp_formulation.cc:605-681); throws "unsupported" for EnergyCost /
// AngularMomentumCost
void BuildCostTables(const HostProgram& hp, const towr_b200_spec& s, CostTables& out);
#ifdef __CUDACC__
cudaError_t LaunchCost(const CostDevice& c, int what, const double* x, double* f, double* grad, int n_problems, int max_ctas,
                       cudaStream_t stream);
// allocation-free variant: the tables go into a file-scope device array of cost_kernel.cu (no cudaMalloc / cudaFree);
// *why names the limit or call that failed
cudaError_t LaunchCostStatic(const CostTables& h, int device, int what, const double* x, double* f, double* grad,
                             int n_problems, cudaStream_t stream, const char** why);
#endif

}  // namespace tb

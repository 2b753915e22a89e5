// Exercises include/towr_b200_coalescer.hpp with a stub evaluator (TEST INFRASTRUCTURE: the rendezvous policy needs
// no device; `gpu` binds the real batch).  Solver threads run different numbers of "iterations" with different
// numbers of line-search trials, as Ipopt instances do.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>

#include "towr_b200_coalescer.hpp"
#include "towr_b200_ifopt.hpp"

using namespace towr_b200;

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  try {
    const int N = 24;
    NlpFormulation f;
    f.SetRobot(TOWR_B200_ANYMAL);
    f.SetGait(1, 2.0);
    f.SetNominalStates(1.0, 0.0, 0.0);
    towr_b200_structure* st = nullptr;
    Check(towr_b200_structure_create(&f.params(), &st));
    const int nv = towr_b200_structure_n_vars(st), nr = towr_b200_structure_n_rows(st), nnz = towr_b200_structure_nnz(st);
    std::vector<double> x0(nv);
    Check(towr_b200_problem_x0(st, &f.params(), x0.data()));
    towr_b200_batch* batch = nullptr;
    std::atomic<long long> stub_calls{0};
    EvalCoalescer::Evaluator ev;
    if (gpu) {
      Check(towr_b200_batch_create(st, N, 0, &batch));
      Check(towr_b200_batch_set_all_problems(batch, &f.params()));
      ev = EvalCoalescer::BindBatch(batch);
    } else {
      ev = [&](const int* ids, int n, int what, const double* x, double* g, double* jac) {
        ++stub_calls;
        std::this_thread::sleep_for(std::chrono::microseconds(300));     // a launch takes time
        for (int i = 0; i < n; ++i) {
          double s = 0.0;
          for (int k = 0; k < nv; ++k) s += x[static_cast<size_t>(i) * nv + k];
          if (what & TOWR_B200_EVAL_G) for (int r = 0; r < nr; ++r) g[static_cast<size_t>(i) * nr + r] = s + ids[i] + r;
          if (what & TOWR_B200_EVAL_JAC) for (int e = 0; e < nnz; ++e) jac[static_cast<size_t>(i) * nnz + e] = s - ids[i] + e;
        }
        return 0;
      };
    }
    EvalCoalescer co(N, nv, nr, nnz, ev, std::chrono::microseconds(2000));
    std::atomic<int> errors{0};
    std::vector<double> ref_g0(N, 0.0);
    std::vector<std::thread> th;
    for (int p = 0; p < N; ++p)
      th.emplace_back([&, p] {
        try {
          std::vector<double> x(x0), g(nr), jac(nnz);
          const int iters = 3 + (p * 7) % 11;                         // solvers converge at different times
          for (int it = 0; it < iters; ++it) {
            const int trials = 1 + (p + it) % 3;                     // line-search trials: eval_g only
            for (int t = 0; t < trials; ++t) {
              for (int k = 0; k < nv; ++k) x[k] = x0[k] + 1e-3 * std::sin(0.1 * k + p + it + t);
              co.Eval(p, TOWR_B200_EVAL_G, x.data(), g.data(), nullptr);
              if (!gpu) {
                double s = 0.0;
                for (int k = 0; k < nv; ++k) s += x[k];
                if (g[5] != s + p + 5) ++errors;
              } else if (!std::isfinite(g[5])) ++errors;
            }
            co.Eval(p, TOWR_B200_EVAL_BOTH, x.data(), g.data(), jac.data());
            if (!gpu) {
              double s = 0.0;
              for (int k = 0; k < nv; ++k) s += x[k];
              if (jac[7] != s - p + 7) ++errors;
            } else if (!std::isfinite(jac[7])) ++errors;
          }
          ref_g0[p] = g[0];
          co.Deactivate(p);
        } catch (const std::exception& e) {
          std::printf("thread %d: %s\n", p, e.what());
          ++errors;
          co.Deactivate(p);
        }
      });
    for (auto& t : th) t.join();
    std::printf("requests %lld launches %lld active %d errors %d\n", co.requests(), co.launches(), co.active(), errors.load());
    if (batch) towr_b200_batch_destroy(batch);
    towr_b200_structure_destroy(st);
    if (errors.load() != 0 || co.active() != 0) { std::printf("FAILED\n"); return 1; }
    if (co.launches() * 2 > co.requests()) { std::printf("FAILED: no coalescing\n"); return 1; }
    std::printf("ok\n");
    return 0;
  } catch (const std::exception& e) {
    std::printf("FAILED: %s\n", e.what());
    return 1;
  }
}

// Exercises include/towr_b200_coalescer.hpp with a stub evaluator (TEST INFRASTRUCTURE: the rendezvous policy needs
// no device; `gpu` binds the real batch).  Solver threads run different numbers of "iterations" with different
// numbers of line-search trials, as Ipopt instances do.
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>

#include "towr_b200_coalescer.hpp"
#include "towr_b200_ifopt.hpp"

using namespace towr_b200;

// TEST INFRASTRUCTURE: the oracle's C surface (oracle/oracle_capi.h), linked only into this test.  towr_b200_spec and
// orc_spec have the same fields in the same order (tests/util.py: to_oracle_spec).
#include "../../oracle/oracle_capi.h"
static_assert(sizeof(orc_spec) == sizeof(towr_b200_spec), "spec layouts must match");

static double BandRatio(const double* a, const double* b, int n) {    // max |a-b| / (1e-14 + 1e-12 |b|)
  double worst = 0.0;
  for (int i = 0; i < n; ++i) {
    const double r = std::fabs(a[i] - b[i]) / (1e-14 + 1e-12 * std::fabs(b[i]));
    if (!(r <= worst)) worst = r;      // NaN propagates
  }
  return worst;
}

// policy test: an evaluator that throws, and a staging allocator that fails, must reach every waiting solver thread as
// an exception - nobody may be left blocked in Eval()
static int ThrowMode() {
  const int N = 12, nv = 8, nr = 4, nnz = 6;
  std::atomic<int> calls{0}, caught{0}, ok{0};
  EvalCoalescer::Evaluator ev = [&](const int*, int n, int what, const double*, double* g, double* jac) -> int {
    if (++calls == 3) throw std::runtime_error("injected evaluator failure");
    for (int i = 0; i < n * nr && g; ++i) g[i] = 1.0;
    for (int i = 0; i < n * nnz && jac && (what & TOWR_B200_EVAL_JAC); ++i) jac[i] = 2.0;
    return 0;
  };
  std::atomic<int> allocs{0};
  EvalCoalescer::Staging st;
  st.alloc = [&](size_t b) -> void* { return ++allocs == 5 ? nullptr : std::malloc(b); };   // the 5th staging allocation fails
  st.free = [](void* p) { std::free(p); };
  EvalCoalescer co(N, nv, nr, nnz, ev, std::chrono::microseconds(500), st);
  std::vector<std::thread> th;
  for (int p = 0; p < N; ++p)
    th.emplace_back([&, p] {
      std::vector<double> x(nv, 0.5), g(nr), jac(nnz);
      for (int it = 0; it < 6; ++it) {
        try {
          co.Eval(p, it % 2 ? TOWR_B200_EVAL_BOTH : TOWR_B200_EVAL_G, x.data(), g.data(), jac.data());
          ++ok;
        } catch (const std::exception&) {
          ++caught;
        }
      }
      co.Deactivate(p);
    });
  for (auto& t : th) t.join();           // a deadlock would hang here (the Python test has a timeout)
  std::printf("throw_mode ok %d caught %d calls %d\n", ok.load(), caught.load(), calls.load());
  if (caught.load() < 1 || ok.load() < 1 || ok.load() + caught.load() != N * 6) { std::printf("FAILED\n"); return 1; }
  std::printf("ok\n");
  return 0;
}

// bench N iters [dynamic]: N solver threads (one per problem, as one Ipopt instance per host thread would run:
// towr_ros/src/towr_ros_app.cc:95-117 is the solver set-up they stand in for; Ipopt itself is absent), every iteration
// = one eval_g + one eval_jac_g round trip through the coalescer on the GPU.  Prints one JSON line.
static int BenchMode(int N, int iters) {
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
  Check(towr_b200_batch_create(st, N, 0, &batch));
  Check(towr_b200_batch_set_all_problems(batch, &f.params()));
  EvalCoalescer co(N, nv, nr, nnz, EvalCoalescer::BindBatch(batch), std::chrono::microseconds(500), EvalCoalescer::PinnedStaging());
  std::atomic<int> errors{0};
  std::vector<std::thread> th;
  std::atomic<int> ready{0};
  std::atomic<bool> go{false};
  std::chrono::steady_clock::time_point t0;
  for (int p = 0; p < N; ++p)
    th.emplace_back([&, p] {
      try {
        std::vector<double> x(x0), g(nr), jac(nnz);
        for (int k = 0; k < nv; ++k) x[k] = x0[k] + 1e-3 * std::sin(0.1 * k + p);
        co.Eval(p, TOWR_B200_EVAL_BOTH, x.data(), g.data(), jac.data());        // warm-up: buffers, first launch
        ++ready;
        while (!go.load()) std::this_thread::yield();
        for (int it = 0; it < iters; ++it) {
          x[it % nv] += 1e-6;
          co.Eval(p, TOWR_B200_EVAL_G, x.data(), g.data(), nullptr);            // eval_g
          co.Eval(p, TOWR_B200_EVAL_JAC, x.data(), nullptr, jac.data());        // eval_jac_g
        }
        if (!std::isfinite(g[0]) || !std::isfinite(jac[0])) ++errors;
        co.Deactivate(p);
      } catch (const std::exception& e) {
        std::printf("thread %d: %s\n", p, e.what());
        ++errors;
        co.Deactivate(p);
      }
    });
  while (ready.load() < N) std::this_thread::yield();
  const long long l0 = co.launches();
  t0 = std::chrono::steady_clock::now();
  go = true;
  for (auto& t : th) t.join();
  const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
  std::printf("{\"solver_threads\": %d, \"iterations\": %d, \"ms_per_iteration\": %.4f, \"evaluator_calls\": %lld, "
              "\"host_cores\": %u, \"errors\": %d, \"iteration\": \"one eval_g + one eval_jac_g round trip per solver thread through "
              "EvalCoalescer (pinned staging) and towr_b200_batch_eval; stub solver threads, Ipopt absent\"}\n",
              N, iters, ms / iters, co.launches() - l0, std::thread::hardware_concurrency(), errors.load());
  towr_b200_batch_destroy(batch);
  towr_b200_structure_destroy(st);
  return errors.load() ? 1 : 0;
}

int main(int argc, char** argv) {
  const bool gpu = argc > 1 && std::strcmp(argv[1], "gpu") == 0;
  if (argc > 1 && std::strcmp(argv[1], "throw") == 0) return ThrowMode();
  if (argc > 3 && std::strcmp(argv[1], "bench") == 0) {
    try {
      return BenchMode(std::atoi(argv[2]), std::atoi(argv[3]));
    } catch (const std::exception& e) {
      std::printf("FAILED: %s\n", e.what());
      return 1;
    }
  }
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
    EvalCoalescer co(N, nv, nr, nnz, ev, std::chrono::microseconds(2000),
                     gpu ? EvalCoalescer::PinnedStaging() : EvalCoalescer::Staging());
    std::atomic<int> compared{0};
    double worst_g = 0.0, worst_j = 0.0;
    std::mutex worst_mu;
    orc_spec ospec;
    std::memcpy(&ospec, &f.params(), sizeof(ospec));
    std::atomic<int> errors{0};
    std::vector<double> ref_g0(N, 0.0);
    std::vector<std::thread> th;
    for (int p = 0; p < N; ++p)
      th.emplace_back([&, p] {
        try {
          std::vector<double> x(x0), g(nr), jac(nnz), go(nr), vo(nnz);
          orc_problem* op = gpu ? orc_create(&ospec) : nullptr;       // the checker: one oracle object graph per thread
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
              } else {                                               // every g row against the oracle, strict band
                orc_eval_g(op, x.data(), go.data());
                const double r = BandRatio(g.data(), go.data(), nr);
                if (!(r <= 1.0)) ++errors;
                std::lock_guard<std::mutex> lk(worst_mu);
                if (r > worst_g) worst_g = r;
                ++compared;
              }
            }
            co.Eval(p, TOWR_B200_EVAL_BOTH, x.data(), g.data(), jac.data());
            if (!gpu) {
              double s = 0.0;
              for (int k = 0; k < nv; ++k) s += x[k];
              if (jac[7] != s - p + 7) ++errors;
            } else {
              // every stored value against the oracle: 4 x the band covers the dynamic constraint's base-ang block
              // (regrouped product rule, tests/util.py: REGROUPED_BAND), everything else is within 1 x
              orc_eval_g(op, x.data(), go.data());
              if (orc_eval_jac(op, x.data(), nullptr, nullptr, vo.data()) != nnz) ++errors;
              const double rg = BandRatio(g.data(), go.data(), nr), rj = BandRatio(jac.data(), vo.data(), nnz);
              if (!(rg <= 1.0) || !(rj <= 4.0)) ++errors;
              std::lock_guard<std::mutex> lk(worst_mu);
              if (rg > worst_g) worst_g = rg;
              if (rj > worst_j) worst_j = rj;
              ++compared;
            }
          }
          ref_g0[p] = g[0];
          if (op) orc_destroy(op);
          co.Deactivate(p);
        } catch (const std::exception& e) {
          std::printf("thread %d: %s\n", p, e.what());
          ++errors;
          co.Deactivate(p);
        }
      });
    for (auto& t : th) t.join();
    std::printf("requests %lld launches %lld active %d errors %d\n", co.requests(), co.launches(), co.active(), errors.load());
    if (gpu) std::printf("oracle_checks %d worst_g %.3g worst_jac %.3g\n", compared.load(), worst_g, worst_j);
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

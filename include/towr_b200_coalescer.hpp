// towr_b200_coalescer.hpp — coalesces the eval callbacks of many host solver threads into batched launches.
//
// In the reference every problem is solved by its own ifopt::IpoptSolver (towr_ros_app.cc, towr/test/*_example.cc):
// Ipopt calls eval_g / eval_jac_g once per iteration and per line-search trial, each walking the whole component tree
// on one core.  Here one solver instance per host thread keeps running unchanged, but its
// callbacks block in EvalCoalescer::Eval() until a launch that contains its problem has finished.  A launch goes out
// as soon as every still-active solver has a request pending, or when the oldest pending request has waited
// `max_wait` (solvers diverge: different iteration counts, line-search trials, early convergence - SURVEY.md H5).
// A solver that is done calls Deactivate(); the rendezvous never waits for it again, so there is no deadlock when
// problems converge at different times.
//
// The evaluator is a callable so that the policy can be tested without a device; BindBatch() binds the product path
// (towr_b200_batch_eval, one H2D + launch + D2H for all pending problems).  There is no CPU fallback in the product.
#pragma once
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <functional>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "towr_b200.h"

namespace towr_b200 {

class EvalCoalescer {
 public:
  // evaluator(problem_ids, n, what, x[n*n_vars], g[n*n_rows], jac[n*nnz]) -> 0 on success (the towr_b200_batch_eval contract)
  using Evaluator = std::function<int(const int*, int, int, const double*, double*, double*)>;

  EvalCoalescer(int n_problems, int n_vars, int n_rows, int nnz, Evaluator eval,
                std::chrono::microseconds max_wait = std::chrono::microseconds(200))
      : n_(n_problems), n_vars_(n_vars), n_rows_(n_rows), nnz_(nnz), eval_(std::move(eval)), max_wait_(max_wait),
        active_(n_problems, 1), n_active_(n_problems), req_(n_problems) {}

  static Evaluator BindBatch(towr_b200_batch* batch) {
    return [batch](const int* ids, int n, int what, const double* x, double* g, double* jac) {
      return towr_b200_batch_eval(batch, ids, n, what, x, g, jac);
    };
  }

  // Called from solver thread `problem` (an Ipopt eval_g / eval_jac_g callback).  Blocks until the values are there.
  // what: TOWR_B200_EVAL_G / _JAC / _BOTH; g_out / jac_out may be null when not asked for.
  void Eval(int problem, int what, const double* x, double* g_out, double* jac_out) {
    std::unique_lock<std::mutex> lk(mu_);
    if (problem < 0 || problem >= n_ || !active_[problem]) throw std::runtime_error("EvalCoalescer: inactive problem");
    Request& r = req_[problem];
    if (r.pending) throw std::runtime_error("EvalCoalescer: one request per problem at a time");
    r = Request{true, false, what, x, g_out, jac_out, 0, {}};
    if (n_pending_ == 0) oldest_ = std::chrono::steady_clock::now();
    ++n_pending_;
    cv_.notify_all();         // a waiting thread may now see "everyone active is pending"
    while (true) {
      if (r.done) break;
      if (!launching_ && n_pending_ > 0 && (n_pending_ >= n_active_ || std::chrono::steady_clock::now() - oldest_ >= max_wait_)) {
        Launch(lk);           // this thread becomes the launcher for everything pending
        continue;
      }
      // wait for the running launch to finish, or for more requests / the timeout of the oldest pending request
      if (launching_) cv_.wait(lk);
      else cv_.wait_until(lk, oldest_ + max_wait_);
    }
    const int rc = r.rc;
    const std::string err = r.error;
    r.pending = false;
    if (rc != 0) throw std::runtime_error("EvalCoalescer: batched evaluation failed: " + err);
  }

  // The solver of `problem` has finished (converged, failed or gave up): never wait for it again.
  void Deactivate(int problem) {
    std::lock_guard<std::mutex> lk(mu_);
    if (problem >= 0 && problem < n_ && active_[problem]) {
      active_[problem] = 0;
      --n_active_;
    }
    cv_.notify_all();         // the others may now be "everyone who is still active"
  }

  long long launches() const { std::lock_guard<std::mutex> lk(mu_); return launches_; }
  long long requests() const { std::lock_guard<std::mutex> lk(mu_); return requests_; }
  int active() const { std::lock_guard<std::mutex> lk(mu_); return n_active_; }

 private:
  struct Request {
    bool pending = false, done = false;
    int what = 0;
    const double* x = nullptr;
    double* g = nullptr;
    double* jac = nullptr;
    int rc = 0;
    std::string error;
  };

  // mu_ held on entry and exit; released around the evaluator call
  void Launch(std::unique_lock<std::mutex>& lk) {
    launching_ = true;
    std::vector<int> ids;
    int what = 0;
    for (int p = 0; p < n_; ++p)
      if (req_[p].pending && !req_[p].done) { ids.push_back(p); what |= req_[p].what; }
    const int n = static_cast<int>(ids.size());
    xs_.resize(static_cast<size_t>(n) * n_vars_);
    gs_.resize((what & TOWR_B200_EVAL_G) ? static_cast<size_t>(n) * n_rows_ : 0);
    js_.resize((what & TOWR_B200_EVAL_JAC) ? static_cast<size_t>(n) * nnz_ : 0);
    for (int i = 0; i < n; ++i) std::memcpy(xs_.data() + static_cast<size_t>(i) * n_vars_, req_[ids[i]].x, sizeof(double) * n_vars_);
    n_pending_ = 0;           // later arrivals form the next launch
    ++round_;
    lk.unlock();
    int rc = 0;
    std::string err;
    try {
      rc = eval_(ids.data(), n, what, xs_.data(), gs_.empty() ? nullptr : gs_.data(), js_.empty() ? nullptr : js_.data());
      if (rc != 0) err = towr_b200_last_error();
    } catch (const std::exception& e) {
      rc = -1;
      err = e.what();
    }
    lk.lock();
    for (int i = 0; i < n; ++i) {
      Request& r = req_[ids[i]];
      if (rc == 0) {
        if ((r.what & TOWR_B200_EVAL_G) && r.g) std::memcpy(r.g, gs_.data() + static_cast<size_t>(i) * n_rows_, sizeof(double) * n_rows_);
        if ((r.what & TOWR_B200_EVAL_JAC) && r.jac) std::memcpy(r.jac, js_.data() + static_cast<size_t>(i) * nnz_, sizeof(double) * nnz_);
      }
      r.rc = rc;
      r.error = err;
      r.done = true;
    }
    ++launches_;
    requests_ += n;
    launching_ = false;
    if (n_pending_ > 0) oldest_ = std::chrono::steady_clock::now();   // conservative: restart the clock for late arrivals
    cv_.notify_all();
  }

  const int n_, n_vars_, n_rows_, nnz_;
  Evaluator eval_;
  const std::chrono::microseconds max_wait_;
  mutable std::mutex mu_;
  std::condition_variable cv_;
  std::vector<char> active_;
  int n_active_;
  std::vector<Request> req_;
  int n_pending_ = 0;
  bool launching_ = false;
  unsigned long long round_ = 0;
  std::chrono::steady_clock::time_point oldest_{};
  long long launches_ = 0, requests_ = 0;
  std::vector<double> xs_, gs_, js_;
};

}  // namespace towr_b200

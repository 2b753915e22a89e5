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
// Requests that only want g (line-search trials) and requests that want the Jacobian go out as separate evaluator calls
// of one round: nobody pays the 8 * nnz bytes of D2H for values it did not ask for.
//
// The evaluator is a callable so that the policy can be tested without a device; BindBatch() binds the product path
// (towr_b200_batch_eval, one H2D + launch + D2H for all pending problems) and PinnedStaging() gives the coalescer
// page-locked staging buffers through the C ABI.  There is no CPU fallback in the product.
// ShardedCoalescer (below) runs one EvalCoalescer per device for a batch that is sharded over several GPUs.
#pragma once
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <new>
#include <stdexcept>
#include <string>
#include <vector>

#include "towr_b200.h"

namespace towr_b200 {

// staging memory of a coalescer: alloc(bytes) / free(ptr); the default is the heap, EvalCoalescer::PinnedStaging()
// page-locked memory
struct CoalescerStaging {
  std::function<void*(size_t)> alloc = [](size_t b) { return std::malloc(b); };
  std::function<void(void*)> free = [](void* p) { std::free(p); };
};

class EvalCoalescer {
 public:
  // evaluator(problem_ids, n, what, x[n*n_vars], g[n*n_rows], jac[n*nnz]) -> 0 on success (the towr_b200_batch_eval contract)
  using Evaluator = std::function<int(const int*, int, int, const double*, double*, double*)>;

  using Staging = CoalescerStaging;
  static Staging PinnedStaging() {
    Staging s;
    s.alloc = [](size_t b) { void* p = nullptr; return towr_b200_host_alloc(b, &p) == TOWR_B200_OK ? p : nullptr; };
    s.free = [](void* p) { towr_b200_host_free(p); };
    return s;
  }

  EvalCoalescer(int n_problems, int n_vars, int n_rows, int nnz, Evaluator eval,
                std::chrono::microseconds max_wait = std::chrono::microseconds(200), Staging staging = Staging())
      : n_(n_problems), n_vars_(n_vars), n_rows_(n_rows), nnz_(nnz), eval_(std::move(eval)), max_wait_(max_wait),
        active_(n_problems, 1), n_active_(n_problems), req_(n_problems), xs_(staging), gs_(staging), js_(staging) {}

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

  // growable staging buffer in caller-chosen memory
  class Buffer {
   public:
    explicit Buffer(Staging st) : st_(std::move(st)) {}
    ~Buffer() { if (p_) st_.free(p_); }
    Buffer(const Buffer&) = delete;
    Buffer& operator=(const Buffer&) = delete;
    double* Reserve(size_t n) {
      if (n > cap_) {
        if (p_) st_.free(p_);
        p_ = nullptr; cap_ = 0;
        p_ = static_cast<double*>(st_.alloc(sizeof(double) * n));
        if (!p_) throw std::bad_alloc();
        cap_ = n;
      }
      return p_;
    }
   private:
    Staging st_;
    double* p_ = nullptr;
    size_t cap_ = 0;
  };

  // mu_ held on entry and exit; released around the evaluator calls.  Whatever happens in between - an allocation that
  // fails, an evaluator that throws - every gathered request is marked done (with the error) and launching_ is reset:
  // no solver thread is left waiting for a launch that will never finish.
  void Launch(std::unique_lock<std::mutex>& lk) {
    launching_ = true;
    std::vector<int> ids;
    int rc = 0;
    std::string err;
    try {
      // this round's requests, G-only ones first: the two groups go out as separate evaluator calls
      std::vector<int> g_only, with_jac;
      for (int p = 0; p < n_; ++p)
        if (req_[p].pending && !req_[p].done) ((req_[p].what & TOWR_B200_EVAL_JAC) ? with_jac : g_only).push_back(p);
      ids = g_only;
      ids.insert(ids.end(), with_jac.begin(), with_jac.end());
      const int n = static_cast<int>(ids.size()), ng = static_cast<int>(g_only.size()), nj = n - ng;
      int what_j = 0;
      for (int p : with_jac) what_j |= req_[p].what;
      const bool any_g = ng > 0 || (what_j & TOWR_B200_EVAL_G);
      double* xs = xs_.Reserve(static_cast<size_t>(n) * n_vars_);
      double* gs = any_g ? gs_.Reserve(static_cast<size_t>(n) * n_rows_) : nullptr;
      double* js = nj > 0 ? js_.Reserve(static_cast<size_t>(nj) * nnz_) : nullptr;
      for (int i = 0; i < n; ++i) std::memcpy(xs + static_cast<size_t>(i) * n_vars_, req_[ids[i]].x, sizeof(double) * n_vars_);
      n_pending_ = 0;           // later arrivals form the next launch
      ++round_;
      lk.unlock();
      try {
        if (ng > 0) rc = eval_(ids.data(), ng, TOWR_B200_EVAL_G, xs, gs, nullptr);
        if (rc == 0 && nj > 0)
          rc = eval_(ids.data() + ng, nj, what_j, xs + static_cast<size_t>(ng) * n_vars_,
                     (what_j & TOWR_B200_EVAL_G) ? gs + static_cast<size_t>(ng) * n_rows_ : nullptr, js);
        if (rc != 0) err = towr_b200_last_error();
      } catch (const std::exception& e) {
        rc = -1;
        err = e.what();
      }
      lk.lock();
      if (rc == 0)
        for (int i = 0; i < n; ++i) {
          Request& r = req_[ids[i]];
          if ((r.what & TOWR_B200_EVAL_G) && r.g) std::memcpy(r.g, gs + static_cast<size_t>(i) * n_rows_, sizeof(double) * n_rows_);
          if ((r.what & TOWR_B200_EVAL_JAC) && r.jac) std::memcpy(r.jac, js + static_cast<size_t>(i - ng) * nnz_, sizeof(double) * nnz_);
        }
      launches_ += (ng > 0) + (nj > 0);
    } catch (const std::exception& e) {       // gathering failed (bad_alloc ...): mu_ is held here, by construction
      if (!lk.owns_lock()) lk.lock();
      rc = -1;
      err = e.what();
      if (ids.empty())
        for (int p = 0; p < n_; ++p)
          if (req_[p].pending && !req_[p].done) ids.push_back(p);
      n_pending_ = 0;
    }
    for (int p : ids) {
      Request& r = req_[p];
      r.rc = rc;
      r.error = err;
      r.done = true;
    }
    requests_ += static_cast<long long>(ids.size());
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
  Buffer xs_, gs_, js_;
};

// One EvalCoalescer per device for a batch sharded over several GPUs (contiguous problem ranges, as bench.py shards its
// batch): solver thread `problem` talks to the coalescer of the device that holds it.  No data-path collective: the
// shards never exchange anything.
class ShardedCoalescer {
 public:
  // batches[d] holds problems [first[d], first[d] + towr_b200_batch_size(batches[d])) of the job
  ShardedCoalescer(const std::vector<towr_b200_batch*>& batches, int n_vars, int n_rows, int nnz,
                   std::chrono::microseconds max_wait = std::chrono::microseconds(200)) {
    int first = 0;
    for (towr_b200_batch* b : batches) {
      const int n = towr_b200_batch_size(b);
      first_.push_back(first);
      shards_.emplace_back(new EvalCoalescer(n, n_vars, n_rows, nnz, EvalCoalescer::BindBatch(b), max_wait, EvalCoalescer::PinnedStaging()));
      first += n;
    }
    first_.push_back(first);
  }
  int size() const { return first_.back(); }
  void Eval(int problem, int what, const double* x, double* g_out, double* jac_out) {
    const int d = Shard(problem);
    shards_[d]->Eval(problem - first_[d], what, x, g_out, jac_out);
  }
  void Deactivate(int problem) {
    const int d = Shard(problem);
    shards_[d]->Deactivate(problem - first_[d]);
  }
  long long launches() const { long long n = 0; for (auto& s : shards_) n += s->launches(); return n; }
  long long requests() const { long long n = 0; for (auto& s : shards_) n += s->requests(); return n; }

 private:
  int Shard(int problem) const {
    if (problem < 0 || problem >= first_.back()) throw std::runtime_error("ShardedCoalescer: problem out of range");
    int d = 0;
    while (problem >= first_[d + 1]) ++d;
    return d;
  }
  std::vector<int> first_;
  std::vector<std::unique_ptr<EvalCoalescer>> shards_;
};

}  // namespace towr_b200

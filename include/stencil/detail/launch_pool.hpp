#pragma once
/* LaunchPool: enqueue the per-GPU work of one step from one host thread PER GPU instead of one thread walking all
   GPUs (cudaSetDevice + waits + a launch with ~1 KB of parameters + an event record per GPU add up to tens of
   microseconds on an 8-GPU box, which a host-synchronised loop — the reference's deployment: one rank, all GPUs,
   /root/reference/src/stencil.cu:670-864 — pays in full every iteration).

   run(n, fn) calls fn(0) .. fn(n-1) exactly once each and returns when all have returned: fn(0) on the calling
   thread, fn(i) on worker i, which is created on first use and lives as long as the pool. Workers spin for about a
   millisecond after their last job before they go to sleep on a condition variable, so a tight iteration loop never
   pays a wake-up. fn must be safe to run concurrently for different i (CUDA runtime calls on different devices are). */
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace stencil {
namespace detail {

class LaunchPool {
  struct Worker {
    std::thread thread;
    std::atomic<unsigned long long> posted{0}; // jobs handed to this worker
    std::atomic<unsigned long long> finished{0};
  };
  std::vector<Worker *> workers_;
  const std::function<void(size_t)> *job_ = nullptr;
  std::mutex mu_;
  std::condition_variable cv_;
  std::atomic<bool> stop_{false};
  std::atomic<int> sleepers_{0};

  void loop(Worker *w, size_t index) {
    unsigned long long seen = 0;
    for (;;) {
      /* wait for the next job: spin first, then sleep */
      const auto t0 = std::chrono::steady_clock::now();
      unsigned spins = 0;
      while (w->posted.load(std::memory_order_acquire) == seen && !stop_.load(std::memory_order_relaxed)) {
        if ((++spins & 1023u) == 0 && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(1)) {
          std::unique_lock<std::mutex> lk(mu_);
          sleepers_.fetch_add(1);
          cv_.wait(lk, [&] { return w->posted.load(std::memory_order_acquire) != seen || stop_.load(); });
          sleepers_.fetch_sub(1);
        }
      }
      if (w->posted.load(std::memory_order_acquire) == seen) return; // stop
      ++seen;
      (*job_)(index);
      w->finished.store(seen, std::memory_order_release);
    }
  }

public:
  LaunchPool() {}
  LaunchPool(const LaunchPool &) = delete;
  LaunchPool &operator=(const LaunchPool &) = delete;
  ~LaunchPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_.store(true);
    }
    cv_.notify_all();
    for (Worker *w : workers_) {
      if (w->thread.joinable()) w->thread.join();
      delete w;
    }
  }

  void run(size_t n, const std::function<void(size_t)> &fn) {
    if (n == 0) return;
    if (n == 1) {
      fn(0);
      return;
    }
    while (workers_.size() < n - 1) {
      Worker *w = new Worker;
      const size_t index = workers_.size() + 1;
      workers_.push_back(w);
      w->thread = std::thread([this, w, index] { loop(w, index); });
    }
    job_ = &fn;
    for (size_t i = 0; i + 1 < n; ++i) workers_[i]->posted.fetch_add(1, std::memory_order_release);
    if (sleepers_.load() > 0) {
      std::lock_guard<std::mutex> lk(mu_);
      cv_.notify_all();
    }
    fn(0);
    for (size_t i = 0; i + 1 < n; ++i) {
      Worker *w = workers_[i];
      const unsigned long long want = w->posted.load(std::memory_order_relaxed);
      while (w->finished.load(std::memory_order_acquire) != want) std::this_thread::yield();
    }
  }
};

} // namespace detail
} // namespace stencil

// Persistent host worker threads for the per-tree planning of a batch (one pool per family).
// run(n, fn) calls fn(i, w) for i in [0, n) on the workers (w = 0..workers-1) and on the calling
// thread (w = workers) and returns when all calls are done; tasks are handed out through one atomic
// counter.
#pragma once
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

namespace lhb {

class WorkerPool {
 public:
  explicit WorkerPool(int workers) {
    for (int i = 0; i < workers; ++i) threads_.emplace_back([this, i] { loop(i); });
  }
  ~WorkerPool() {
    {
      std::lock_guard<std::mutex> lk(m_);
      stop_ = true;
    }
    wake_.notify_all();
    for (auto& t : threads_) t.join();
  }
  WorkerPool(const WorkerPool&) = delete;
  WorkerPool& operator=(const WorkerPool&) = delete;

  int workers() const { return (int)threads_.size(); }

  void run(int n, const std::function<void(int, int)>& fn) {
    if (n <= 0) return;
    if (threads_.empty() || n == 1) {
      for (int i = 0; i < n; ++i) fn(i, workers());
      return;
    }
    {
      std::lock_guard<std::mutex> lk(m_);
      fn_ = &fn;
      n_ = n;
      next_.store(0);
      active_ = (int)threads_.size();
      ++generation_;
    }
    wake_.notify_all();
    drain(workers());
    std::unique_lock<std::mutex> lk(m_);
    done_.wait(lk, [this] { return active_ == 0; });
    fn_ = nullptr;
  }

 private:
  void drain(int w) {
    for (;;) {
      const int i = next_.fetch_add(1);
      if (i >= n_) return;
      (*fn_)(i, w);
    }
  }
  void loop(int w) {
    unsigned seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lk(m_);
        wake_.wait(lk, [&] { return stop_ || generation_ != seen; });
        if (stop_) return;
        seen = generation_;
      }
      drain(w);
      std::lock_guard<std::mutex> lk(m_);
      if (--active_ == 0) done_.notify_one();
    }
  }

  std::vector<std::thread> threads_;
  std::mutex m_;
  std::condition_variable wake_, done_;
  const std::function<void(int, int)>* fn_ = nullptr;
  int n_ = 0, active_ = 0;
  unsigned generation_ = 0;
  std::atomic<int> next_{0};
  bool stop_ = false;
};

}  // namespace lhb

// Small fixed-size thread pool (push + wait-for-idle).  The reference vendors BS::thread_pool
// (common/bs_thread_pool.h, used at controller.cc:39 and controller_servicer.cc:113); this is an
// independent minimal implementation of the two calls the controller needs.
#pragma once
#include <condition_variable>
#include <functional>
#include <mutex>
#include <queue>
#include <thread>
#include <vector>

namespace mfl::core {

class ThreadPool {
 public:
  explicit ThreadPool(unsigned n) {
    for (unsigned i = 0; i < (n ? n : 1); ++i) workers_.emplace_back([this] { Run(); });
  }
  ~ThreadPool() { Stop(); }
  void Push(std::function<void()> fn) {
    {
      std::lock_guard<std::mutex> l(mu_);
      if (stopping_) return;
      q_.push(std::move(fn));
      ++pending_;
    }
    cv_.notify_one();
  }
  void WaitIdle() {
    std::unique_lock<std::mutex> l(mu_);
    idle_cv_.wait(l, [this] { return pending_ == 0; });
  }
  void Stop() {
    {
      std::lock_guard<std::mutex> l(mu_);
      if (stopping_) return;
      stopping_ = true;
    }
    cv_.notify_all();
    for (auto& t : workers_)
      if (t.joinable()) t.join();
  }

 private:
  void Run() {
    for (;;) {
      std::function<void()> fn;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [this] { return stopping_ || !q_.empty(); });
        if (q_.empty()) return;  // stopping and drained
        fn = std::move(q_.front());
        q_.pop();
      }
      try { fn(); } catch (...) {}
      {
        std::lock_guard<std::mutex> l(mu_);
        --pending_;
      }
      idle_cv_.notify_all();
    }
  }
  std::mutex mu_;
  std::condition_variable cv_, idle_cv_;
  std::queue<std::function<void()>> q_;
  std::vector<std::thread> workers_;
  size_t pending_ = 0;
  bool stopping_ = false;
};

}  // namespace mfl::core

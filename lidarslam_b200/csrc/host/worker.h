// One helper thread per handle that runs the host side of work the caller does not have to wait for:
// the map update of frame t (its read-backs need a host thread) while the caller already returns with the pose,
// and the upload + keypoint extraction of a frame handed in ahead of time (lsb_prefetch_frames).  Jobs run in
// the order they were posted; a ticket identifies a job, wait(ticket) returns once it is over and rethrows
// what it threw.  The helper spins for a short while after a job (frames arrive back to back), then sleeps.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <deque>
#include <exception>
#include <functional>
#include <map>
#include <mutex>
#include <thread>
#include <cuda_runtime.h>

namespace lsb {

class AsyncWorker {
 public:
  explicit AsyncWorker(int device) : device_(device) { th_ = std::thread([this] { loop(); }); }
  ~AsyncWorker()
  {
    { std::lock_guard<std::mutex> g(m_); stop_ = true; }
    cv_.notify_one();
    if (th_.joinable()) th_.join();
  }
  AsyncWorker(const AsyncWorker&) = delete;
  AsyncWorker& operator=(const AsyncWorker&) = delete;

  // spin = false: the helper sleeps at once when idle (many handles sharing the host cores)
  unsigned post(std::function<void()> fn, bool spin = true)
  {
    unsigned t;
    {
      std::lock_guard<std::mutex> g(m_);
      t = ++posted_;
      q_.emplace_back(std::move(fn));
      spin_ = spin;
    }
    pending_.fetch_add(1, std::memory_order_release);
    if (sleeping_.load()) cv_.notify_one();
    return t;
  }

  bool done(unsigned ticket) const { return (int)(completed_.load(std::memory_order_acquire) - ticket) >= 0; }

  void wait(unsigned ticket)
  {
    unsigned spins = 0;
    while (!done(ticket)) {
      relax();
      if ((++spins & 0xfffu) == 0) std::this_thread::yield();
    }
    std::exception_ptr e;
    {
      std::lock_guard<std::mutex> g(m_);
      auto it = err_.find(ticket);
      if (it != err_.end()) { e = it->second; err_.erase(it); }
    }
    if (e) std::rethrow_exception(e);
  }
  unsigned last_posted() const { return posted_; }   // caller's thread only
  // every job posted so far is over; errors of jobs nobody waited for are dropped
  void drain()
  {
    const unsigned t = posted_;
    while (!done(t)) relax();
    std::lock_guard<std::mutex> g(m_);
    err_.clear();
  }

 private:
  std::thread th_;
  std::mutex m_;
  std::condition_variable cv_;
  std::deque<std::function<void()>> q_;
  std::map<unsigned, std::exception_ptr> err_;
  unsigned posted_ = 0;
  std::atomic<unsigned> completed_{0};
  std::atomic<int> pending_{0};
  std::atomic<bool> sleeping_{false};
  bool stop_ = false, spin_ = true;
  int device_;

  static void relax()
  {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#else
    std::this_thread::yield();
#endif
  }

  void loop()
  {
    cudaSetDevice(device_);
    const auto linger = std::chrono::milliseconds(20);
    auto last = std::chrono::steady_clock::now();
    unsigned ticket = 0;
    while (true) {
      unsigned spins = 0;
      while (pending_.load(std::memory_order_acquire) == 0) {
        relax();
        bool nap = false;
        if ((++spins & 1023u) == 0) {
          std::lock_guard<std::mutex> g(m_);
          if (stop_) return;
          nap = !spin_ || std::chrono::steady_clock::now() - last > linger;
        }
        if (nap) {
          std::unique_lock<std::mutex> g(m_);
          sleeping_.store(true);
          cv_.wait(g, [&] { return !q_.empty() || stop_; });
          sleeping_.store(false);
          if (stop_ && q_.empty()) return;
        }
      }
      std::function<void()> fn;
      {
        std::lock_guard<std::mutex> g(m_);
        if (q_.empty()) continue;
        fn = std::move(q_.front());
        q_.pop_front();
      }
      pending_.fetch_sub(1, std::memory_order_acq_rel);
      ++ticket;
      try { fn(); }
      catch (...) { std::lock_guard<std::mutex> g(m_); err_[ticket] = std::current_exception(); }
      last = std::chrono::steady_clock::now();
      completed_.store(ticket, std::memory_order_release);
    }
  }
};

}  // namespace lsb

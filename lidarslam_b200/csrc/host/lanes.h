// Host lanes: the three keypoint types (edges, planes, blobs) are independent RollingGrid /
// KeypointsMatcher instances in the reference (Slam.h:553, Slam.cxx:1003-1037, :1530-1560), which
// walks them in an OpenMP `parallel for`.  Here each type owns a CUDA stream; the lanes give each
// stream its own enqueuing host thread, so that the per-type launch sequences (and their small
// read-backs) do not queue behind one another on a single host thread.
// Two helper threads per handle: the caller runs the first type itself.  Helpers spin for a short
// while after a task (frames arrive back to back), then sleep on a condition variable.
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <exception>
#include <functional>
#include <mutex>
#include <thread>
#include <cuda_runtime.h>

namespace lsb {

class TypeLanes {
 public:
  explicit TypeLanes(int device) : device_(device)
  {
    for (int i = 0; i < 2; ++i) w_[i].th = std::thread([this, i] { loop(w_[i]); });
  }
  ~TypeLanes()
  {
    for (auto& w : w_) {
      { std::lock_guard<std::mutex> g(w.m); w.stop.store(true); }
      w.cv.notify_one();
      if (w.th.joinable()) w.th.join();
    }
  }
  TypeLanes(const TypeLanes&) = delete;
  TypeLanes& operator=(const TypeLanes&) = delete;

  // fn(order[0]) on the caller, fn(order[1]) and fn(order[2]) on the helpers; returns when the
  // three calls are over and rethrows the first exception (caller's first).
  void run(const int order[3], const std::function<void(int)>& fn)
  {
    for (int i = 0; i < 2; ++i) post(w_[i], order[i + 1], &fn);
    std::exception_ptr mine;
    try { fn(order[0]); } catch (...) { mine = std::current_exception(); }
    for (auto& w : w_) wait(w);
    std::exception_ptr first = mine;
    for (auto& w : w_) {
      if (w.err && !first) first = w.err;
      w.err = nullptr;
    }
    if (first) std::rethrow_exception(first);
  }

 private:
  struct Worker {
    std::thread th;
    std::mutex m;
    std::condition_variable cv;
    std::atomic<unsigned> req{0}, ack{0};
    std::atomic<bool> sleeping{false}, stop{false};
    const std::function<void(int)>* fn = nullptr;
    int arg = 0;
    std::exception_ptr err;
  };
  Worker w_[2];
  int device_;

  static void relax()
  {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#else
    std::this_thread::yield();
#endif
  }

  void post(Worker& w, int arg, const std::function<void(int)>* fn)
  {
    w.fn = fn;
    w.arg = arg;
    w.req.fetch_add(1);                         // seq_cst: ordered against `sleeping`
    if (w.sleeping.load()) {
      { std::lock_guard<std::mutex> g(w.m); }   // the helper is inside cv.wait (or has seen req)
      w.cv.notify_one();
    }
  }

  static void wait(Worker& w)
  {
    const unsigned want = w.req.load();
    while (w.ack.load(std::memory_order_acquire) != want) relax();
  }

  void loop(Worker& w)
  {
    cudaSetDevice(device_);
    unsigned seen = 0;
    const auto linger = std::chrono::milliseconds(20);
    auto last = std::chrono::steady_clock::now();
    while (true) {
      // spin while frames are in flight, sleep when the handle goes idle
      unsigned spins = 0;
      while (w.req.load(std::memory_order_acquire) == seen && !w.stop.load(std::memory_order_relaxed)) {
        relax();
        if ((++spins & 1023u) == 0 && std::chrono::steady_clock::now() - last > linger) {
          std::unique_lock<std::mutex> g(w.m);
          w.sleeping.store(true);
          w.cv.wait(g, [&] { return w.req.load() != seen || w.stop.load(); });
          w.sleeping.store(false);
        }
      }
      if (w.stop.load()) return;
      seen = w.req.load(std::memory_order_acquire);
      try { (*w.fn)(w.arg); } catch (...) { w.err = std::current_exception(); }
      last = std::chrono::steady_clock::now();
      w.ack.store(seen, std::memory_order_release);
    }
  }
};

}  // namespace lsb

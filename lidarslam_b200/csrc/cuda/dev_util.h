// Device-side plumbing shared by the kernels of liblidarslam_b200: error
// handling, growable device buffers, launch helpers.  No torch types: the
// library owns its device memory and one CUDA stream per lsb_slam handle.
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <cstdint>
#include <cstddef>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <stdexcept>
#include "../../../include/lidarslam_b200.h"

namespace lsb {

struct CudaError : std::runtime_error {
  int code;
  CudaError(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

inline void cuda_check(cudaError_t e, const char* what, const char* file, int line)
{
  if (e != cudaSuccess) {
    char buf[512];
    std::snprintf(buf, sizeof(buf), "%s failed at %s:%d: %s", what, file, line, cudaGetErrorString(e));
    throw CudaError(LSB_E_CUDA, buf);
  }
}
#define LSB_CUDA(x) ::lsb::cuda_check((x), #x, __FILE__, __LINE__)
#define LSB_LAUNCH_CHECK() ::lsb::cuda_check(cudaGetLastError(), "kernel launch", __FILE__, __LINE__)

// Growable device array on the stream-ordered allocator: growth is a cudaMallocAsync /
// cudaFreeAsync pair ordered on the handle's stream (no device synchronisation); the default
// memory pool keeps freed blocks (release threshold raised in lsb_slam), so a growing map costs
// O(log) cheap reallocations and steady state allocates nothing.  Content is kept only when asked.
template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t cap = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  // A block of the stream-ordered allocator goes back to its pool only through cudaFreeAsync: after cudaFree the
  // pool keeps counting it as used (measured, tools/probes/pool_free_probe.cu: 161 MiB lost per destroyed OS-128
  // handle).  The owner may have destroyed its streams already, so the free is ordered on the legacy stream,
  // behind a device-wide synchronisation (what cudaFree did implicitly).
  ~DevBuf()
  {
    if (!p) return;
    cudaDeviceSynchronize();
    if (cudaFreeAsync(p, cudaStreamLegacy) != cudaSuccess) { cudaGetLastError(); cudaFree(p); }
  }
  void reserve(size_t n, cudaStream_t s = 0, bool keep = false, size_t keep_n = 0)
  {
    if (n <= cap) return;
    // geometric growth (x2, floor 64 Ki elements)
    size_t ncap = cap ? cap : 65536;
    while (ncap < n) ncap = ncap * 2;
    T* q = nullptr;
    LSB_CUDA(cudaMallocAsync(&q, ncap * sizeof(T), s));
    if (p) {
      if (keep && keep_n) LSB_CUDA(cudaMemcpyAsync(q, p, keep_n * sizeof(T), cudaMemcpyDeviceToDevice, s));
      LSB_CUDA(cudaFreeAsync(p, s));
    }
    p = q;
    cap = ncap;
  }
  void swap(DevBuf& o) { T* tp = p; p = o.p; o.p = tp; size_t tc = cap; cap = o.cap; o.cap = tc; }
};

// Pinned host staging buffer (results read back by the host driver).
template <typename T>
struct PinnedBuf {
  T* p = nullptr;
  size_t cap = 0;
  PinnedBuf() = default;
  PinnedBuf(const PinnedBuf&) = delete;
  PinnedBuf& operator=(const PinnedBuf&) = delete;
  ~PinnedBuf() { if (p) cudaFreeHost(p); }
  void reserve(size_t n)
  {
    if (n <= cap) return;
    if (p) cudaFreeHost(p);
    LSB_CUDA(cudaMallocHost(&p, n * sizeof(T)));
    std::memset(p, 0, n * sizeof(T));       // sequence words polled by host threads must not start out as garbage
    cap = n;
  }
};

inline unsigned div_up(size_t a, size_t b) { return (unsigned)((a + b - 1) / b); }

// CUDA-event bracket around selected kernel launches (lsb_set_profiling).
struct KernelTimer {
  bool enabled = false;
  std::vector<cudaEvent_t> ev;
  size_t used = 0;
  unsigned launches = 0;
  ~KernelTimer() { for (auto e : ev) cudaEventDestroy(e); }
  void begin(cudaStream_t s)
  {
    ++launches;
    if (!enabled) return;
    while (used + 2 > ev.size()) { cudaEvent_t e; LSB_CUDA(cudaEventCreate(&e)); ev.push_back(e); }
    LSB_CUDA(cudaEventRecord(ev[used], s));
  }
  void end(cudaStream_t s)
  {
    if (!enabled) return;
    LSB_CUDA(cudaEventRecord(ev[used + 1], s));
    used += 2;
  }
  // call after the stream is synchronised; returns ms and resets (each: the individual durations in us, in launch order)
  float collect(unsigned* n_launches, std::vector<float>* each = nullptr)
  {
    float total = 0.f;
    if (each) for (size_t i = 0; i + 1 < used; i += 2) { float ms = 0.f; cudaEventElapsedTime(&ms, ev[i], ev[i + 1]); each->push_back(ms * 1e3f); }
    for (size_t i = 0; i + 1 < used; i += 2) { float ms = 0.f; cudaEventElapsedTime(&ms, ev[i], ev[i + 1]); total += ms; }
    used = 0;
    if (n_launches) *n_launches = launches;
    launches = 0;
    return total;
  }
};

// Launch counter of one handle (lsb_frame_stats.kernel_launches).
struct LaunchCounter {   // bumped from the helper threads (host/worker.h) as well
  std::atomic<unsigned> n{0};
  void add(unsigned k = 1) { n.fetch_add(k, std::memory_order_relaxed); }
};

}  // namespace lsb

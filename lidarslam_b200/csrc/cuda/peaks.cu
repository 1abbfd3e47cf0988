// Measured arithmetic peaks of the device (diagnostic for bench.py's roofline denominators; SURVEY 8(d) asks for a
// dependent-FMA microbenchmark on the box instead of a nominal figure).  Every thread runs 8 independent chains of
// dependent operations (enough ILP to cover the pipe latency at full occupancy); four variants:
//   [0] FP64 fused multiply-add            (2 flops per instruction)
//   [1] FP64 multiply then add, not fused  (what this library issues: it is built with -fmad=false for bit parity with the CPU oracle)
//   [2] FP32 fused multiply-add
//   [3] FP32 multiply then add, not fused
// Result: TFLOP/s of each, best of several launches timed with CUDA events.
#include "dev_util.h"
#include <algorithm>

namespace lsb {

namespace {

constexpr int PEAK_ITERS = 4096;
constexpr int PEAK_CHAINS = 8;

template <typename T, bool FUSED>
__global__ void __launch_bounds__(256) k_peak(T* __restrict__ sink, T a, T b)
{
  T v[PEAK_CHAINS];
#pragma unroll
  for (int c = 0; c < PEAK_CHAINS; ++c) v[c] = (T)(threadIdx.x + c) * (T)1e-3;
  for (int i = 0; i < PEAK_ITERS; ++i) {
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; ++c) {
      if (FUSED) v[c] = (sizeof(T) == 8) ? (T)fma((double)v[c], (double)a, (double)b) : (T)fmaf((float)v[c], (float)a, (float)b);
      else v[c] = v[c] * a + b;      // -fmad=false: a multiply and an add
    }
  }
  T s = (T)0;
#pragma unroll
  for (int c = 0; c < PEAK_CHAINS; ++c) s += v[c];
  if (s == (T)123456789) sink[0] = s;   // keeps the chains alive
}

template <typename T, bool FUSED>
double measure(int sms, T* sink, cudaStream_t s)
{
  const int grid = sms * 8, threads = 256;
  cudaEvent_t e0, e1;
  LSB_CUDA(cudaEventCreate(&e0)); LSB_CUDA(cudaEventCreate(&e1));
  double best = 0.0;
  for (int rep = 0; rep < 6; ++rep) {
    LSB_CUDA(cudaEventRecord(e0, s));
    k_peak<T, FUSED><<<grid, threads, 0, s>>>(sink, (T)0.999999, (T)1e-7);
    LSB_LAUNCH_CHECK();
    LSB_CUDA(cudaEventRecord(e1, s));
    LSB_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    LSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double flops = 2.0 * PEAK_ITERS * PEAK_CHAINS * (double)grid * threads;
    if (rep > 0 && ms > 0.f) best = std::max(best, flops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return best;
}

}  // namespace

void measure_fp_peaks(int device, double out_tflops[4])
{
  LSB_CUDA(cudaSetDevice(device));
  int sms = 0;
  LSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
  cudaStream_t s;
  LSB_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  DevBuf<double> sink;
  sink.reserve(16, s);
  out_tflops[0] = measure<double, true>(sms, sink.p, s);
  out_tflops[1] = measure<double, false>(sms, sink.p, s);
  out_tflops[2] = measure<float, true>(sms, reinterpret_cast<float*>(sink.p), s);
  out_tflops[3] = measure<float, false>(sms, reinterpret_cast<float*>(sink.p), s);
  LSB_CUDA(cudaStreamSynchronize(s));
  cudaStreamDestroy(s);
}

}  // namespace lsb

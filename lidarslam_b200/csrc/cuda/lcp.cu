// LCP overlap estimator: reference slam_lib/src/ConfidenceEstimators.cxx:27-65 and
// Slam::EstimateOverlap (Slam.cxx:1370-1388).  The nearest-neighbour distances come from the kNN
// kernels (k = 1, capped at the distance where the probability is exactly 0.f); this file holds
// the sample gather and the canonical float tree sum of common/lcp_core.h.
#include "../host/slam.cuh"
#include "../common/lcp_core.h"

namespace lsb {

namespace {

// query n = xyz of cloud[(size_t)(n / ratio)]
__global__ void k_lcp_gather(const lsb_point* __restrict__ cloud, int nb, float ratio, float* __restrict__ q_xyz)
{
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= nb) return;
  const float4 p = reinterpret_cast<const float4*>(cloud)[2 * lcp_sample_index(n, ratio)];
  q_xyz[3 * (size_t)n] = p.x; q_xyz[3 * (size_t)n + 1] = p.y; q_xyz[3 * (size_t)n + 2] = p.z;
}

__device__ __forceinline__ float fold32f(float v)
{
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
  return v;   // complete in lane 0
}

// one CTA per 1024-sample chunk: best probability over the maps, fold32 per warp, fold32 of the 32 warp sums
__global__ void __launch_bounds__(LSB_LCP_CHUNK)
k_lcp_chunks(const float* __restrict__ d2_0, const float* __restrict__ d2_1, const float* __restrict__ d2_2, float thr0, float thr1, float thr2,
             int nb, float* __restrict__ chunk_sum)
{
  __shared__ float s_group[32];
  const int i = blockIdx.x * LSB_LCP_CHUNK + threadIdx.x;
  float best = 0.f;
  if (i < nb) {
    if (d2_0) { const float p = lcp_probability(d2_0[i], thr0); best = (p > best) ? p : best; }
    if (d2_1) { const float p = lcp_probability(d2_1[i], thr1); best = (p > best) ? p : best; }
    if (d2_2) { const float p = lcp_probability(d2_2[i], thr2); best = (p > best) ? p : best; }
  }
  const float g = fold32f(best);
  if ((threadIdx.x & 31) == 0) s_group[threadIdx.x >> 5] = g;
  __syncthreads();
  if (threadIdx.x < 32) {
    const float c = fold32f(s_group[threadIdx.x]);
    if (threadIdx.x == 0) chunk_sum[blockIdx.x] = c;
  }
}

// one warp: fold32 of the chunk sums in blocks of 32, block sums added in order, mean
__global__ void k_lcp_final(const float* __restrict__ chunk_sum, int nchunks, int nb, float* __restrict__ out)
{
  const int lane = threadIdx.x;
  float lcp = 0.f;
  for (int b = 0; b * 32 < nchunks; ++b) {
    const int c = b * 32 + lane;
    const float v = fold32f(c < nchunks ? chunk_sum[c] : 0.f);
    lcp = (b == 0) ? v : lcp + v;
  }
  if (lane == 0) out[0] = lcp / (float)nb;
}

}  // namespace

void dev_lcp_gather(const lsb_point* d_cloud, int nb, float ratio, float* d_q_xyz, cudaStream_t s, LaunchCounter& lc)
{
  k_lcp_gather<<<div_up((size_t)nb, 256), 256, 0, s>>>(d_cloud, nb, ratio, d_q_xyz); LSB_LAUNCH_CHECK(); lc.add();
}

void dev_lcp_reduce(const float* const d2[3], const float thr[3], int nb, DevBuf<float>& chunk_sum, float* d_out, cudaStream_t s, LaunchCounter& lc)
{
  const int nchunks = (nb + LSB_LCP_CHUNK - 1) / LSB_LCP_CHUNK;
  chunk_sum.reserve((size_t)nchunks, s);
  k_lcp_chunks<<<nchunks, LSB_LCP_CHUNK, 0, s>>>(d2[0], d2[1], d2[2], thr[0], thr[1], thr[2], nb, chunk_sum.p); LSB_LAUNCH_CHECK(); lc.add();
  k_lcp_final<<<1, 32, 0, s>>>(chunk_sum.p, nchunks, nb, d_out); LSB_LAUNCH_CHECK(); lc.add();
}

}  // namespace lsb

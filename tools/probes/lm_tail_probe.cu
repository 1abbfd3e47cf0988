// Probe (not product): cycle counts of the two serial pieces at the end of k_lm_eval —
// the in-order sum of the chunk partials and the trust-region control step.
#include "../../lidarslam_b200/csrc/cuda/match.cu"
#include <cstdio>
#include <vector>
using namespace lsb;

__global__ void probe(LmState* state, LmPoint* point, unsigned* counter, const double* part, int nchunks, long long* t)
{
  __shared__ double s_E[LSB_LM_NSUM];
  __shared__ double s_tile[LM_TILE * LSB_LM_NSUM];
  __shared__ __align__(16) unsigned long long s_state_raw[sizeof(LmState) / 8];
  const long long t0 = clock64();
  double e = 0.0;
  for (int c0 = 0; c0 < nchunks; c0 += LM_TILE) {
    const int nt = min(LM_TILE, nchunks - c0);
    for (int q = threadIdx.x; q < nt * LSB_LM_NSUM; q += blockDim.x) s_tile[q] = __ldcg(part + (size_t)c0 * LSB_LM_NSUM + q);
    __syncthreads();
    if (threadIdx.x < LSB_LM_NSUM) {
      int c = 0;
      if (c0 == 0) { e = s_tile[threadIdx.x]; c = 1; }
      for (; c < nt; ++c) e = e + s_tile[c * LSB_LM_NSUM + threadIdx.x];
    }
    __syncthreads();
  }
  if (threadIdx.x < LSB_LM_NSUM) s_E[threadIdx.x] = e;
  __syncthreads();
  const long long t1 = clock64();
  if (threadIdx.x < 32) lm_control(state, point, counter, s_E, 1, s_state_raw);
  __syncthreads();
  const long long t2 = clock64();
  if (threadIdx.x < 32) lm_control(state, point, counter, s_E, 0, s_state_raw);
  __syncthreads();
  const long long t3 = clock64();
  if (threadIdx.x == 0) { t[0] = t1 - t0; t[1] = t2 - t1; t[2] = t3 - t2; }
}

// lm_control's steps with a time stamp between them (same code, same order)
__global__ void probe_control(LmState* state, LmPoint* point, const double* E_in, long long* t)
{
  __shared__ double sE[LSB_LM_NSUM];
  __shared__ __align__(16) unsigned long long s_state_raw[sizeof(LmState) / 8];
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int NW = (int)(sizeof(LmState) / 8);
  const int lane = threadIdx.x & 31;
  if (lane < LSB_LM_NSUM) sE[lane] = E_in[lane];
  __syncwarp();
  const long long t0 = clock64();
  for (int w = lane; w < NW; w += 32) s_state_raw[w] = reinterpret_cast<const unsigned long long*>(state)[w];
  __syncwarp();
  const long long t1 = clock64();
  LmState& s = *reinterpret_cast<LmState*>(s_state_raw);
  int propose = 0;
  if (lane == 0) propose = lm_first_update(s, sE) ? 1 : 0;
  __syncwarp();
  propose = __shfl_sync(FULL, propose, 0);
  const long long t2 = clock64();
  if (propose) lm_propose_warp(s);
  __syncwarp();
  const long long t3 = clock64();
  for (int w = lane; w < NW; w += 32) reinterpret_cast<unsigned long long*>(state)[w] = s_state_raw[w];
  double xc[6];
  for (int q = 0; q < 6; ++q) xc[q] = s.xc[q];
  const double tv = (lane < 6) ? lm_point_trig(xc, lane) : 0.0;
  double trig[6];
  for (int q = 0; q < 6; ++q) trig[q] = __shfl_sync(FULL, tv, q);
  const long long t4 = clock64();
  if (lane == 0) {
    LmPoint pt;
    lm_make_point_from_trig(xc, trig, pt);
    *point = pt;
  }
  __syncwarp();
  const long long t5 = clock64();
  if (lane == 0) { t[0] = t1 - t0; t[1] = t2 - t1; t[2] = t3 - t2; t[3] = t4 - t3; t[4] = t5 - t4; t[5] = propose; }
}

// dependent-chain latency of DADD / DMUL / DFMA-free code
__global__ void chain(double* out, long long* t, int n)
{
  double a = out[0], b = out[1];
  const long long t0 = clock64();
  for (int i = 0; i < n; ++i) a = a + b;
  const long long t1 = clock64();
  for (int i = 0; i < n; ++i) a = a * b;
  const long long t2 = clock64();
  for (int i = 0; i < n; ++i) a = b / a;
  const long long t3 = clock64();
  for (int i = 0; i < n; ++i) a = sqrt(a + b);
  const long long t4 = clock64();
  out[2] = a;
  t[0] = t1 - t0; t[1] = t2 - t1; t[2] = t3 - t2; t[3] = t4 - t3;
}

int main()
{
  const int nchunks = 116;
  std::vector<double> part((size_t)nchunks * LSB_LM_NSUM, 0.0);
  // a plausible SPD normal matrix: diag heavy
  for (int c = 0; c < nchunks; ++c) {
    double* p = &part[(size_t)c * LSB_LM_NSUM];
    int o = 0;
    for (int i = 0; i < 6; ++i) for (int j = i; j < 6; ++j) p[o++] = (i == j) ? 50.0 + i : 0.3 * (i + j);
    for (int i = 0; i < 6; ++i) p[21 + i] = 0.01 * (i + 1);
    p[27] = 10.0;
  }
  double* d_part; LmState* d_state; LmPoint* d_point; unsigned* d_counter; long long* d_t; double* d_out;
  cudaMalloc(&d_part, part.size() * 8); cudaMalloc(&d_state, sizeof(LmState)); cudaMalloc(&d_point, sizeof(LmPoint));
  cudaMalloc(&d_counter, 8); int* d_hist; cudaMalloc(&d_hist, 256); cudaMemset(d_hist, 1, 256); cudaMalloc(&d_t, 64); cudaMalloc(&d_out, 64);
  cudaMemcpy(d_part, part.data(), part.size() * 8, cudaMemcpyHostToDevice);
  for (int rep = 0; rep < 3; ++rep) {
    k_lm_begin<<<1, 1>>>(d_state, d_point, d_counter, 0.1, 0.2, 0.3, 0.01, 0.02, 0.03, 15, 0, d_hist, 1u);
    probe<<<1, 256>>>(d_state, d_point, d_counter, d_part, nchunks, d_t);
    long long t[4];
    cudaMemcpy(t, d_t, 32, cudaMemcpyDeviceToHost);
    printf("rep %d: chunk-sum %lld cycles, control(first) %lld cycles, control(next) %lld cycles  (%s)\n", rep, t[0], t[1], t[2], cudaGetErrorString(cudaGetLastError()));
  }
  for (int rep = 0; rep < 3; ++rep) {
    k_lm_begin<<<1, 1>>>(d_state, d_point, d_counter, 0.1, 0.2, 0.3, 0.01, 0.02, 0.03, 15, 0, d_hist, 1u);
    probe_control<<<1, 32>>>(d_state, d_point, d_part, d_t);
    long long t[6];
    cudaMemcpy(t, d_t, 48, cudaMemcpyDeviceToHost);
    printf("control: load state %lld, first_update %lld, propose_warp %lld (proposed=%lld), write-back+trig %lld, make_point+store %lld cycles\n",
           t[0], t[1], t[2], t[5], t[3], t[4]);
  }
  double init[2] = {1.0000001, 1.0000002};
  cudaMemcpy(d_out, init, 16, cudaMemcpyHostToDevice);
  chain<<<1, 1>>>(d_out, d_t, 1000);
  long long t[4];
  cudaMemcpy(t, d_t, 32, cudaMemcpyDeviceToHost);
  printf("dependent latency per op (cycles): dadd %.1f dmul %.1f ddiv %.1f dsqrt+add %.1f\n", t[0] / 1000., t[1] / 1000., t[2] / 1000., t[3] / 1000.);
  return 0;
}

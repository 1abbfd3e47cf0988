// Probe (not product): cost of the persistent LM kernel (k_lm_solve, used unmodified) per
// cost/Jacobian evaluation, on synthetic records of the OS-128 problem size: the kernel is timed
// with max_iter = 0, 1, 2, ... (the synthetic problem converges after a few steps, the slope over
// the first ones is the cost of one evaluation + barrier + control step).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -I../../include -o lm_eval_probe lm_eval_probe.cu ../../lidarslam_b200/csrc/cuda/knn.cu
#define LSB_LM_PROBE 1
#include "../../lidarslam_b200/csrc/cuda/match.cu"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include <algorithm>
#include <functional>
using namespace lsb;

__global__ void k_empty(int) {}

static float time_launches(int reps, const std::function<void()>& f)
{
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int i = 0; i < reps; ++i) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms * 1e3f / reps;
}

int main(int argc, char** argv)
{
  const int n = argc > 1 ? atoi(argv[1]) : 118000;
  const size_t stride = (size_t)(n + 31) / 32 * 32;
  const int nchunks = (n + LSB_LM_CHUNK - 1) / LSB_LM_CHUNK;
  std::vector<double> rec(13 * stride, 0.0);
  srand(7);
  auto rnd = [] { return rand() / (double)RAND_MAX - 0.5; };
  for (int i = 0; i < n; ++i) {
    double nx = rnd(), ny = rnd(), nz = rnd();
    const double nn = std::sqrt(nx * nx + ny * ny + nz * nz) + 1e-9;
    nx /= nn; ny /= nn; nz /= nn;
    const double A[9] = {nx * nx, nx * ny, nx * nz, nx * ny, ny * ny, ny * nz, nx * nz, ny * nz, nz * nz};
    const double P[3] = {40 * rnd(), 40 * rnd(), 6 * rnd()};
    LmRecord r;
    const double X[3] = {P[0] + 0.05 * rnd(), P[1] + 0.05 * rnd(), P[2] + 0.05 * rnd()};
    lm_record_set(r, A, P, X, (i % 17 == 0) ? -1.0 : 0.8);
    for (int q = 0; q < 6; ++q) rec[q * stride + i] = r.M[q];
    for (int q = 0; q < 3; ++q) { rec[(6 + q) * stride + i] = r.P[q]; rec[(9 + q) * stride + i] = r.X[q]; }
    rec[12 * stride + i] = r.w;
  }
  double* d_rec; double* d_part; LmState* d_state; unsigned* d_bar; int* d_hist;
  const int pstride = (nchunks + 31) / 32 * 32;
  cudaMalloc(&d_rec, rec.size() * 8); cudaMalloc(&d_part, (size_t)2 * pstride * LSB_LM_NSUM * 8);
  cudaMalloc(&d_state, sizeof(LmState)); cudaMalloc(&d_bar, 8); cudaMemset(d_bar, 0, 8);
  cudaMalloc(&d_hist, 3 * LSB_MATCH_N_STATUS * 4);
  std::vector<int> hist(3 * LSB_MATCH_N_STATUS, 1000);
  cudaMemcpy(d_hist, hist.data(), hist.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_rec, rec.data(), rec.size() * 8, cudaMemcpyHostToDevice);
  constexpr size_t kSmem = (size_t)REC_PLANES * LSB_LM_CHUNK * sizeof(double);
  cudaFuncSetAttribute(k_lm_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
  int per_sm = 0, sms = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_lm_solve, LM_THREADS, kSmem);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int share = argc > 2 ? atoi(argv[2]) : 1;      // > 1: the grid a handle gets when that many share the GPU
  const int grid = std::min(nchunks, std::max(per_sm * sms / share, 1));
  LmStart start = {{0.01, 0.02, 0.03, 0.001, 0.002, 0.003}};
  LmSensorTerms terms;
  lm_sensor_terms_clear(terms);
  auto solve = [&](int max_iter) {
    cudaLaunchAttribute attr;
    attr.id = cudaLaunchAttributeCooperative;
    attr.val.cooperative = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(LM_THREADS); cfg.dynamicSmemBytes = kSmem; cfg.attrs = &attr; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, k_lm_solve, (const double*)d_rec, stride, n, nchunks, pstride, 1.0, start, max_iter, 0, (const int*)d_hist, 100u,
                       d_state, d_part, d_bar, terms, (MatchWorkspace::LmHostBlock*)nullptr, 0u, (IcpState*)nullptr, IcpCtl{});
  };
  printf("n = %d records, %d chunks of %d, grid %d x %d threads (%d CTA/SM)\n", n, nchunks, LSB_LM_CHUNK, grid, LM_THREADS, per_sm);
  printf("empty launch   %7.2f us\n", time_launches(200, [&] { k_empty<<<grid, LM_THREADS>>>(0); }));
  float prev = 0.f;
  for (int mi : {0, 1, 2, 3, 15}) {
    const float t = time_launches(200, [&] { solve(mi); });
    LmState st; cudaMemcpy(&st, d_state, sizeof(st), cudaMemcpyDeviceToHost);
    printf("max_iter %2d    %7.2f us  (+%6.2f)  LM iter %d, term %d, cost %.6g  %s\n", mi, t, t - prev, st.iter, st.term, st.cost,
           cudaGetErrorString(cudaGetLastError()));
    prev = t;
  }
  // phases of each evaluation on CTA 0 (globaltimer, ns)
  solve(3); cudaDeviceSynchronize();
  unsigned long long st[64 * 8];
  cudaMemcpyFromSymbol(st, g_lm_probe, sizeof(st));
  for (int e = 0; e < 4; ++e)
    printf("eval %d: evaluate+chunk fold %5.2f us | barrier %5.2f | final fold %5.2f | control %5.2f (update %4.2f, propose %4.2f, next point %4.2f)\n", e,
           (st[e * 8 + 1] - st[e * 8]) * 1e-3, (st[e * 8 + 2] - st[e * 8 + 1]) * 1e-3, (st[e * 8 + 3] - st[e * 8 + 2]) * 1e-3,
           (st[e * 8 + 4] - st[e * 8 + 3]) * 1e-3, (st[e * 8 + 5] - st[e * 8 + 3]) * 1e-3, (st[e * 8 + 6] - st[e * 8 + 5]) * 1e-3,
           (st[e * 8 + 4] - st[e * 8 + 6]) * 1e-3);
  return 0;
}

// Probe (not product): where one k_lm_eval launch spends its time, on synthetic records of the
// OS-128 problem size.  Uses the product kernel unmodified:
//   full      = evaluation + last-CTA tail (chunk-order sum + trust-region control)
//   main only = the ticket counter is preset so that no CTA is "last" (no tail)
//   empty     = a kernel of the same grid that does nothing (launch floor)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -o lm_eval_probe lm_eval_probe.cu
#include "../../lidarslam_b200/csrc/cuda/match.cu"
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cmath>
#include <functional>
using namespace lsb;

__global__ void k_empty(int) {}
__global__ void k_set(unsigned* c, unsigned v) { *c = v; }

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
  double* d_rec; double* d_part; LmState* d_state; LmPoint* d_point; unsigned* d_counter; int* d_hist;
  cudaMalloc(&d_rec, rec.size() * 8); const int pstride = (nchunks + 31) / 32 * 32; cudaMalloc(&d_part, (size_t)pstride * LSB_LM_NSUM * 8);
  cudaMalloc(&d_state, sizeof(LmState)); cudaMalloc(&d_point, sizeof(LmPoint)); cudaMalloc(&d_counter, 8);
  cudaMalloc(&d_hist, 3 * LSB_MATCH_N_STATUS * 4);
  std::vector<int> hist(3 * LSB_MATCH_N_STATUS, 1000);
  cudaMemcpy(d_hist, hist.data(), hist.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(d_rec, rec.data(), rec.size() * 8, cudaMemcpyHostToDevice);
  auto begin = [&] { k_lm_begin<<<1, 1>>>(d_state, d_point, d_counter, 0.01, 0.02, 0.03, 0.001, 0.002, 0.003, 15, 0, d_hist, 100u); };
  auto eval = [&](int first) { k_lm_eval<<<nchunks, LM_THREADS>>>(d_rec, stride, n, nchunks, pstride, 1.0, d_state, d_point, d_part, d_counter, first); };
  printf("n = %d records, %d chunks of %d, %d threads per CTA\n", n, nchunks, LSB_LM_CHUNK, LM_THREADS);
  const float t_empty = time_launches(200, [&] { k_empty<<<nchunks, LM_THREADS>>>(0); });
  const float t_begin = time_launches(200, [&] { begin(); });
  const float t_full = time_launches(200, [&] { begin(); eval(1); });
  const float t_main = time_launches(200, [&] { begin(); k_set<<<1, 1>>>(d_counter, 1u << 30); eval(1); });
  const float t_set = time_launches(200, [&] { begin(); k_set<<<1, 1>>>(d_counter, 1u << 30); });
  printf("empty launch             %7.2f us\n", t_empty);
  printf("k_lm_begin               %7.2f us\n", t_begin);
  printf("begin + eval (full)      %7.2f us  -> eval  %7.2f us\n", t_full, t_full - t_begin);
  printf("begin + set + eval(main) %7.2f us  -> main  %7.2f us  (tail %7.2f us)\n", t_main, t_main - t_set, (t_full - t_begin) - (t_main - t_set));
  // product-like sequence: 1 + 15 launches back to back (they may finish early when the LM converges)
  const float t_seq = time_launches(50, [&] { begin(); for (int e = 0; e <= 15; ++e) eval(e == 0); });
  LmState st; cudaMemcpy(&st, d_state, sizeof(st), cudaMemcpyDeviceToHost);
  printf("begin + 16 evals         %7.2f us  (LM iter %d, term %d, cost %.6g)  %s\n", t_seq, st.iter, st.term, st.cost, cudaGetErrorString(cudaGetLastError()));
  return 0;
}

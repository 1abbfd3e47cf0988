// Probe (not product): cost of the single-thread steps the device-resident ICP loop runs at the tail / head of k_lm_solve
// (cuda/icp.cuh icp_control, the XYZRPY prior), and the latency of dependent FP64 operations on this GPU.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -o icp_control_probe icp_control_probe.cu
#include "../../lidarslam_b200/csrc/cuda/icp.cuh"
#include <cstdio>
#include <cstring>
using namespace lsb;

__global__ void k_probe(IcpState* st, LmState* lm, long long* out, double* sink)
{
  __shared__ IcpState s;
  __shared__ LmState l;
  s = *st; l = *lm;
  IcpCtl c; c.stage = ICP_STAGE_LOC; c.icp_iter = 0; c.icp_max_iter = 3; c.min_matches = 20;
  long long t0 = clock64();
  for (int i = 0; i < 16; ++i) { s.done = 0; l.x[0] += 1e-9; icp_control(s, c, l, 1000u); }
  long long t1 = clock64();
  double x0[6], acc = 0;
  for (int i = 0; i < 16; ++i) { s.Tworld.t.x += 1e-9; s.Tworld.R.m[1] += 1e-12; iso_to_xyzrpy(s.Tworld, x0); acc += x0[3] + x0[5]; }
  long long t2 = clock64();
  double a = l.x[1] + 1.0;
  for (int i = 0; i < 1024; ++i) a = a + 1.000001;
  long long t3 = clock64();
  for (int i = 0; i < 1024; ++i) a = a * 1.000001;
  long long t4 = clock64();
  for (int i = 0; i < 256; ++i) a = 1.000001 / a;
  long long t5 = clock64();
  for (int i = 0; i < 256; ++i) a = ::sqrt(a + 1.0);
  long long t6 = clock64();
  for (int i = 0; i < 64; ++i) a = dm::sin(a);
  long long t7 = clock64();
  for (int i = 0; i < 64; ++i) a = dm::atan2(a, 0.7);
  long long t8 = clock64();
  out[0] = (t1 - t0) / 16; out[1] = (t2 - t1) / 16; out[2] = (t3 - t2) / 1024; out[3] = (t4 - t3) / 1024;
  out[4] = (t5 - t4) / 256; out[5] = (t6 - t5) / 256; out[6] = (t7 - t6) / 64; out[7] = (t8 - t7) / 64;
  *sink = a + acc + s.Tworld.t.x;
  *st = s;
}

int main()
{
  IcpState h; std::memset(&h, 0, sizeof(h));
  h.Tworld = h.PreviousTworld = h.Trelative = Iso3d::identity();
  h.Tworld.t = Vec3d(0.8, 0.02, 0.001);
  h.Within.reset(); h.Within.set_times(-0.1, 0.0); h.Pending = h.Within;
  h.prev_pose_time = 1.0; h.curr_pose_time = 1.1; h.max_extrapolation_ratio = 3.0; h.undistortion = 2;
  LmState l; std::memset(&l, 0, sizeof(l));
  l.x[0] = 0.81; l.x[1] = 0.021; l.x[3] = 0.001; l.x[4] = -0.002; l.x[5] = 0.01; l.num_successful = 3; l.iter = 15;
  IcpState* d; LmState* dl; long long* o; double* sink;
  cudaMalloc(&d, sizeof(h)); cudaMalloc(&dl, sizeof(l)); cudaMalloc(&o, 64); cudaMalloc(&sink, 8);
  cudaMemcpy(d, &h, sizeof(h), cudaMemcpyHostToDevice); cudaMemcpy(dl, &l, sizeof(l), cudaMemcpyHostToDevice);
  k_probe<<<1, 1>>>(d, dl, o, sink);
  long long r[8];
  if (cudaMemcpy(r, o, 64, cudaMemcpyDeviceToHost) != cudaSuccess) { std::printf("failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
  std::printf("cycles (SM clock): icp_control %lld  iso_to_xyzrpy %lld  dadd %lld  dmul %lld  ddiv %lld  dsqrt(+add) %lld  dm::sin %lld  dm::atan2 %lld\n",
              r[0], r[1], r[2], r[3], r[4], r[5], r[6], r[7]);
  std::printf("sizeof(IcpState) %zu sizeof(LmState) %zu\n", sizeof(IcpState), sizeof(LmState));
  return 0;
}

// Driver-cloud adaptors on the device (see convert.cuh).  The reference's loops are sequential
// because of two carried states; both have a closed parallel form:
//  * SpinningFrameAdvancementEstimator (Utilities.h:58-108) keeps the previous advancement of each
//    ring and adds 1 when the wrapped azimuth a falls below it.  Since a < 1, once a ring has been
//    corrected every later point of the ring is corrected too: the correction of a point is
//    "some point of my ring, up to me, has a smaller a than its predecessor in the ring".  Points
//    are bucketed by ring with a stable radix sort (input order kept inside a ring), descents are
//    found by comparing neighbours, and the first descent of each ring is an atomicMin.
//  * the dual-return filter of the Robosense node compares a point with the last KEPT point
//    (RobosenseToLidarNode.cxx:86-88).  A dropped point equals the kept one before it, so comparing
//    with the previous FINITE point gives the same answer; that index is an exclusive max-scan.
#include "convert.cuh"
#include <cub/cub.cuh>

namespace lsb {
namespace {

constexpr int TPB = 256;

__global__ void k_velodyne_points(const lsb_velodyne_point* __restrict__ in, int n, const int32_t* __restrict__ mapping, int n_mapping,
                                  int device_id, int time_valid, lsb_point* __restrict__ out, double* __restrict__ adv,
                                  uint16_t* __restrict__ key, uint32_t* __restrict__ idx)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const lsb_velodyne_point v = in[i];
  const unsigned laser = conv_map_laser_id(v.ring, mapping, n_mapping);
  out[i] = conv_make_point(v.x, v.y, v.z, v.intensity, laser, device_id, time_valid ? (double)v.time : 0.);
  if (!time_valid) {
    const lsb_velodyne_point f = in[0];                       // first point of the frame: advancement 0
    adv[i] = conv_wrap_advancement(conv_point_advancement(v.x, v.y), conv_point_advancement(f.x, f.y));
    key[i] = (uint16_t)laser;
    idx[i] = (uint32_t)i;
  }
}

__global__ void k_fill_u32(uint32_t* p, int n, uint32_t v)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// j runs over the ring-bucketed order; first[r] = first position of ring r whose a is below its predecessor's
__global__ void k_velodyne_descents(const uint16_t* __restrict__ key, const uint32_t* __restrict__ idx, const double* __restrict__ adv, int n,
                                    uint32_t* __restrict__ first)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j <= 0 || j >= n) return;
  const uint16_t r = key[j];
  if (key[j - 1] == r && adv[idx[j]] < adv[idx[j - 1]]) atomicMin(&first[r], (uint32_t)j);
}

__global__ void k_velodyne_times(const uint16_t* __restrict__ key, const uint32_t* __restrict__ idx, const double* __restrict__ adv, int n,
                                 const uint32_t* __restrict__ first, int first_packet, double rpm, lsb_point* __restrict__ out)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t i = idx[j];
  double fa = adv[i];
  if ((uint32_t)j >= first[key[j]]) fa += 1.;
  out[i].time = conv_time_from_advancement(fa, first_packet, rpm);
}

__device__ __forceinline__ bool finite3(const lsb_xyzi_point& p) { return isfinite(p.x) && isfinite(p.y) && isfinite(p.z); }

__global__ void k_rs_finite(const lsb_xyzi_point* __restrict__ in, int n, int32_t* __restrict__ fin)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) fin[i] = finite3(in[i]) ? i : -1;
}

__global__ void k_rs_flags(const lsb_xyzi_point* __restrict__ in, int n, const int32_t* __restrict__ last_finite, uint32_t* __restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const lsb_xyzi_point p = in[i];
  bool keep = finite3(p);
  const int j = last_finite[i];
  if (keep && j >= 0) {
    const lsb_xyzi_point q = in[j];
    if (p.x == q.x && p.y == q.y && p.z == q.z) keep = false;   // std::equal on the three floats
  }
  flags[i] = keep ? 1u : 0u;
}

__global__ void k_rs_scatter(const lsb_xyzi_point* __restrict__ in, int n, const uint32_t* __restrict__ flags, const uint32_t* __restrict__ pos,
                             unsigned width, unsigned height, unsigned points_per_ring, const int32_t* __restrict__ mapping, int n_mapping,
                             int device_id, double rpm, lsb_point* __restrict__ out, uint32_t* __restrict__ count)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (i == n - 1) *count = pos[i] + flags[i];
  if (!flags[i]) return;
  const lsb_xyzi_point p = in[i];
  unsigned laser = (unsigned)i / width;                          // RobosenseToLidarNode.cxx:118
  laser = mapping ? conv_map_laser_id(laser, mapping, n_mapping) : (height == 16 ? conv_rs16_laser_id(laser) : laser);
  out[pos[i]] = conv_make_point(p.x, p.y, p.z, p.intensity, laser, device_id, conv_rs_time((unsigned)i, points_per_ring, rpm));
}

struct MaxOp { __device__ __forceinline__ int32_t operator()(int32_t a, int32_t b) const { return a > b ? a : b; } };

}  // namespace

void ConvertWorkspace::upload_mapping(const lsb_conversion_params& p, cudaStream_t s)
{
  if (p.laser_id_mapping && p.n_laser_id_mapping > 0) {
    mapping.reserve((size_t)p.n_laser_id_mapping, s);
    LSB_CUDA(cudaMemcpyAsync(mapping.p, p.laser_id_mapping, (size_t)p.n_laser_id_mapping * sizeof(int32_t), cudaMemcpyHostToDevice, s));
  }
}

size_t ConvertWorkspace::velodyne(const lsb_velodyne_point* host, size_t n, const lsb_conversion_params& p, cudaStream_t s, LaunchCounter& lc)
{
  n_out = 0;
  if (n == 0) return 0;                                          // VelodyneToLidarNode.cxx:53-57 frame ignored
  const int N = (int)n;
  const bool has_map = p.laser_id_mapping && p.n_laser_id_mapping > 0;
  // VelodyneToLidarNode.cxx:72: the time field is usable if it spans more than 1e-8 s (float difference)
  const int time_valid = (host[n - 1].time - host[0].time > 1e-8) ? 1 : 0;
  raw.reserve(n * sizeof(lsb_velodyne_point), s);
  out.reserve(n, s);
  upload_mapping(p, s);
  LSB_CUDA(cudaMemcpyAsync(raw.p, host, n * sizeof(lsb_velodyne_point), cudaMemcpyHostToDevice, s));
  const lsb_velodyne_point* in = reinterpret_cast<const lsb_velodyne_point*>(raw.p);
  if (!time_valid) { adv.reserve(n, s); key_a.reserve(n, s); key_b.reserve(n, s); idx_a.reserve(n, s); idx_b.reserve(n, s); first.reserve(65536, s); }
  k_velodyne_points<<<div_up(n, TPB), TPB, 0, s>>>(in, N, has_map ? mapping.p : nullptr, has_map ? p.n_laser_id_mapping : 0, p.device_id, time_valid,
                                                   out.p, adv.p, key_a.p, idx_a.p); LSB_LAUNCH_CHECK(); lc.add();
  if (!time_valid) {
    size_t tb = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tb, key_a.p, key_b.p, idx_a.p, idx_b.p, N, 0, 16, s);
    cub_tmp.reserve(tb, s);
    LSB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tb, key_a.p, key_b.p, idx_a.p, idx_b.p, N, 0, 16, s)); lc.add(3);
    k_fill_u32<<<div_up(65536, TPB), TPB, 0, s>>>(first.p, 65536, 0xffffffffu); LSB_LAUNCH_CHECK(); lc.add();
    k_velodyne_descents<<<div_up(n, TPB), TPB, 0, s>>>(key_b.p, idx_b.p, adv.p, N, first.p); LSB_LAUNCH_CHECK(); lc.add();
    k_velodyne_times<<<div_up(n, TPB), TPB, 0, s>>>(key_b.p, idx_b.p, adv.p, N, first.p, p.timestamp_first_packet, p.rpm, out.p);
    LSB_LAUNCH_CHECK(); lc.add();
  }
  n_out = n;
  return n;
}

size_t ConvertWorkspace::robosense(const lsb_xyzi_point* host, uint32_t width, uint32_t height, const lsb_conversion_params& p, cudaStream_t s,
                                   LaunchCounter& lc)
{
  n_out = 0;
  const size_t n = (size_t)width * height;
  if (n == 0) return 0;                                          // RobosenseToLidarNode.cxx:61-65
  const int N = (int)n;
  const bool has_map = p.laser_id_mapping && p.n_laser_id_mapping > 0;
  raw.reserve(n * sizeof(lsb_xyzi_point), s);
  out.reserve(n, s);
  last_finite_in.reserve(n, s); last_finite.reserve(n, s); flags.reserve(n, s); scan.reserve(n, s); first.reserve(65536, s);
  h_count.reserve(1);
  upload_mapping(p, s);
  LSB_CUDA(cudaMemcpyAsync(raw.p, host, n * sizeof(lsb_xyzi_point), cudaMemcpyHostToDevice, s));
  const lsb_xyzi_point* in = reinterpret_cast<const lsb_xyzi_point*>(raw.p);
  k_rs_finite<<<div_up(n, TPB), TPB, 0, s>>>(in, N, last_finite_in.p); LSB_LAUNCH_CHECK(); lc.add();
  size_t tb = 0, tb2 = 0;
  cub::DeviceScan::ExclusiveScan(nullptr, tb, last_finite_in.p, last_finite.p, MaxOp(), (int32_t)-1, N, s);
  cub::DeviceScan::ExclusiveSum(nullptr, tb2, flags.p, scan.p, N, s);
  cub_tmp.reserve(std::max(tb, tb2), s);
  LSB_CUDA(cub::DeviceScan::ExclusiveScan(cub_tmp.p, tb, last_finite_in.p, last_finite.p, MaxOp(), (int32_t)-1, N, s)); lc.add(2);
  k_rs_flags<<<div_up(n, TPB), TPB, 0, s>>>(in, N, last_finite.p, flags.p); LSB_LAUNCH_CHECK(); lc.add();
  LSB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb2, flags.p, scan.p, N, s)); lc.add(2);
  k_rs_scatter<<<div_up(n, TPB), TPB, 0, s>>>(in, N, flags.p, scan.p, width, height, (unsigned)(n / height), has_map ? mapping.p : nullptr,
                                              has_map ? p.n_laser_id_mapping : 0, p.device_id, p.rpm, out.p, first.p); LSB_LAUNCH_CHECK(); lc.add();
  LSB_CUDA(cudaMemcpyAsync(h_count.p, first.p, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  LSB_CUDA(cudaStreamSynchronize(s));
  n_out = h_count.p[0];
  return n_out;
}

}  // namespace lsb

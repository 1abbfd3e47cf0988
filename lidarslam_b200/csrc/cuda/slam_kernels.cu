// K6: rigid transform / motion undistortion / small per-cloud reductions.
// Reference: Utils::TransformPoint (Utilities.h:274-278, FP64 math, FP32 store),
// LinearTransformInterpolator (MotionModel.h:36-136), Slam::RefineUndistortion
// (Slam.cxx:1337-1351), Slam::AggregateFrames (Slam.cxx:1512-1578).
#include "../host/slam.cuh"
#include <climits>
#include <cstring>

namespace lsb {

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ void load_xyz(const lsb_point* p, float& x, float& y, float& z)
{
  const float4 v = reinterpret_cast<const float4*>(p)[0];
  x = v.x; y = v.y; z = v.z;
}
__device__ __forceinline__ void store_xyz(lsb_point* p, float x, float y, float z)
{
  float4 v = reinterpret_cast<float4*>(p)[0];
  v.x = x; v.y = y; v.z = z;
  reinterpret_cast<float4*>(p)[0] = v;
}

__device__ __forceinline__ int f2ord_k(float f) { const int i = __float_as_int(f); return (i >= 0) ? i : (i ^ 0x7fffffff); }

__global__ void k_transform(lsb_point* __restrict__ pts, int n, Iso3d T)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float x, y, z;
  load_xyz(pts + i, x, y, z);
  transform_point_f(T, x, y, z);
  store_xyz(pts + i, x, y, z);
}

// k_transform with the bounding box of the transformed cloud (pcl::getMinMax3D of RollingGrid::Add, RollingGrid.cxx:171-172)
// reduced on the way: the box is that of the floats just stored, i.e. what a separate pass over the cloud would find
__global__ void k_transform_bbox(lsb_point* __restrict__ pts, int n, Iso3d T, int* __restrict__ bb)
{
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    float x, y, z;
    load_xyz(pts + i, x, y, z);
    transform_point_f(T, x, y, z);
    store_xyz(pts + i, x, y, z);
    const int v[3] = {f2ord_k(x), f2ord_k(y), f2ord_k(z)};
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = min(mn[a], v[a]); mx[a] = max(mx[a], v[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    for (int off = 16; off; off >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], off));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], off));
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(bb + a, mn[a]); atomicMax(bb + 3 + a, mx[a]); }
  }
}

// Box of a cloud without an initialisation launch: every block leaves its partial box (ordered-int encoding), the last block
// to arrive (ticket counter, reset by that block) folds the `nb` partials from `first` on and publishes the box in device
// memory and in pinned host memory (UVA: the host reads it after the stream's next synchronisation, no copy enqueued).
__device__ __forceinline__ void box_reduce_publish(int mn[3], int mx[3], int first, int nb, int* __restrict__ partial,
                                                   unsigned* __restrict__ counter, int* __restrict__ d_bb, int* __restrict__ h_bb, int seq)
{
  __shared__ int sh[TPB / 32][6];
  __shared__ bool is_last;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < 3; ++q) {
    for (int off = 16; off; off >>= 1) {
      mn[q] = min(mn[q], __shfl_xor_sync(0xffffffffu, mn[q], off));
      mx[q] = max(mx[q], __shfl_xor_sync(0xffffffffu, mx[q], off));
    }
    if (lane == 0) { sh[wid][q] = mn[q]; sh[wid][3 + q] = mx[q]; }
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    int r = sh[0][threadIdx.x];
    for (int k = 1; k < TPB / 32; ++k) r = (threadIdx.x < 3) ? min(r, sh[k][threadIdx.x]) : max(r, sh[k][threadIdx.x]);
    partial[(size_t)blockIdx.x * 6 + threadIdx.x] = r;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == (unsigned)(nb - 1);
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  int r[6] = {INT_MAX, INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MIN};
  for (int b = threadIdx.x; b < nb; b += TPB) {
    const int* pp = partial + (size_t)(first + b) * 6;
#pragma unroll
    for (int q = 0; q < 3; ++q) { r[q] = min(r[q], __ldcg(pp + q)); r[3 + q] = max(r[3 + q], __ldcg(pp + 3 + q)); }
  }
#pragma unroll
  for (int q = 0; q < 6; ++q) {
    for (int off = 16; off; off >>= 1) {
      const int o = __shfl_xor_sync(0xffffffffu, r[q], off);
      r[q] = (q < 3) ? min(r[q], o) : max(r[q], o);
    }
  }
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int q = 0; q < 6; ++q) sh[wid][q] = r[q];
  }
  __syncthreads();
  if (threadIdx.x < 6) {
    int f = sh[0][threadIdx.x];
    for (int k = 1; k < TPB / 32; ++k) f = (threadIdx.x < 3) ? min(f, sh[k][threadIdx.x]) : max(f, sh[k][threadIdx.x]);
    d_bb[threadIdx.x] = f;
    h_bb[threadIdx.x] = f;
    __threadfence_system();
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    *counter = 0u;
    reinterpret_cast<volatile int*>(h_bb)[6] = seq;      // sequence word: a host thread may poll for the box instead of synchronising
  }
}

// Tail of the device-resident localization loop, ONE launch for the three keypoint clouds: the last undistortion refinement
// (k_undistort_state mode 1), the world copy of the keypoints under the pose the loop ended with (GetKeypoints(k, true):
// Slam.cxx:690-702, :1491-1509) and the bounding box RollingGrid::Add rolls by (RollingGrid.cxx:171-172).  The pose is read
// from the loop's state in device memory, so the launch is enqueued behind the loop before the host knows its result and the
// map update finds its input (cloud + box, the box in pinned host memory) ready.  The box is reduced without an
// initialisation launch: every block leaves its partial box, the last block of a type (ticket counter, reset by that block)
// folds them (box_reduce_publish).
__global__ void __launch_bounds__(TPB) k_finish_keypoints(FinishArgs a, const IcpState* __restrict__ st, int refine,
                                                          int* __restrict__ partial, unsigned* __restrict__ counter)
{
  const int t = (int)blockIdx.x >= a.b0[2] ? 2 : ((int)blockIdx.x >= a.b0[1] ? 1 : 0);
  const int lb = (int)blockIdx.x - a.b0[t], nb = a.b0[t + 1] - a.b0[t];
  const int i = lb * TPB + (int)threadIdx.x;
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  if (i < a.n[t]) {
    lsb_point* c = a.cur[t] + i;
    float4 v = reinterpret_cast<const float4*>(c)[0];
    const uint4 w = reinterpret_cast<const uint4*>(c)[1];
    float x = v.x, y = v.y, z = v.z;
    if (refine && st->final_refine) {
      const Iso3d T = st->Pending.at(c->time);
      transform_point_f(T, x, y, z);
      v.x = x; v.y = y; v.z = z;
      reinterpret_cast<float4*>(c)[0] = v;
    }
    transform_point_f(st->Tworld, x, y, z);
    v.x = x; v.y = y; v.z = z;
    lsb_point* o = a.world[t] + i;
    reinterpret_cast<float4*>(o)[0] = v;
    reinterpret_cast<uint4*>(o)[1] = w;
    mn[0] = mx[0] = f2ord_k(x); mn[1] = mx[1] = f2ord_k(y); mn[2] = mx[2] = f2ord_k(z);
  }
  box_reduce_publish(mn, mx, a.b0[t], nb, partial, counter + t, a.d_bb[t], a.h_bb[t], a.seq[t]);
}

// Head of the sub-map selection of one keypoint type, one launch: the undistortion still pending on the cloud (k_undistort)
// and the box of the cloud under the predicted pose, pcl::transformPointCloud (PCL <= 1.10 explicit form) then
// pcl::getMinMax3D, published like the box above.
__global__ void __launch_bounds__(TPB) k_undistort_transformed_bbox(lsb_point* __restrict__ pts, int n, Interpolator it, int undistort, Iso3d T,
                                                                    int* __restrict__ partial, unsigned* __restrict__ counter,
                                                                    int* __restrict__ d_bb, int* __restrict__ h_bb, int seq)
{
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  const int i = blockIdx.x * TPB + threadIdx.x;
  if (i < n) {
    float fx, fy, fz;
    load_xyz(pts + i, fx, fy, fz);
    if (undistort) {
      const Iso3d U = it.at(pts[i].time);
      transform_point_f(U, fx, fy, fz);
      store_xyz(pts + i, fx, fy, fz);
    }
    const double x = fx, y = fy, z = fz;
    const float wx = (float)(T.R.m[0] * x + T.R.m[1] * y + T.R.m[2] * z + T.t.x);
    const float wy = (float)(T.R.m[3] * x + T.R.m[4] * y + T.R.m[5] * z + T.t.y);
    const float wz = (float)(T.R.m[6] * x + T.R.m[7] * y + T.R.m[8] * z + T.t.z);
    mn[0] = mx[0] = f2ord_k(wx); mn[1] = mx[1] = f2ord_k(wy); mn[2] = mx[2] = f2ord_k(wz);
  }
  box_reduce_publish(mn, mx, 0, (int)gridDim.x, partial, counter, d_bb, h_bb, seq);
}

__global__ void k_aggregate(lsb_point* __restrict__ pts, int n, double timeOffset, Iso3d T, int applyT)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  pts[i].time += timeOffset;
  if (applyT) {
    float x, y, z;
    load_xyz(pts + i, x, y, z);
    transform_point_f(T, x, y, z);
    store_xyz(pts + i, x, y, z);
  }
}

__global__ void k_undistort(lsb_point* __restrict__ pts, int n, Interpolator it, double timeOffset)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double t = pts[i].time;
  if (timeOffset != 0.0) { t += timeOffset; pts[i].time = t; }
  const Iso3d T = it.at(t);
  float x, y, z;
  load_xyz(pts + i, x, y, z);
  transform_point_f(T, x, y, z);
  store_xyz(pts + i, x, y, z);
}

// Undistortion refinement of the device-resident ICP loop (cuda/icp.cuh): the interpolator is the one icp_control left in
// device memory.  mode 0 = head of an ICP iteration (applies unless the loop is over), mode 1 = after the loop (applies
// when the last iteration asked for a final, non-deferred refinement).
__global__ void k_undistort_state(lsb_point* __restrict__ pts, int n, const IcpState* __restrict__ st, int mode)
{
  if (mode == 0 ? st->done != 0 : st->final_refine == 0) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Iso3d T = st->Pending.at(pts[i].time);
  float x, y, z;
  load_xyz(pts + i, x, y, z);
  transform_point_f(T, x, y, z);
  store_xyz(pts + i, x, y, z);
}

__global__ void k_icp_init(IcpState* st, IcpState init) { *st = init; }

__global__ void k_bbox_init(int* bb) { if (threadIdx.x < 3) bb[threadIdx.x] = INT_MAX; else if (threadIdx.x < 6) bb[threadIdx.x] = INT_MIN; }

}  // namespace

void dev_transform(lsb_point* d_pts, size_t n, const Iso3d& T, cudaStream_t s, LaunchCounter& lc)
{
  if (!n) return;
  k_transform<<<div_up(n, TPB), TPB, 0, s>>>(d_pts, (int)n, T); LSB_LAUNCH_CHECK(); lc.add();
}

void dev_transform_bbox_async(lsb_point* d_pts, size_t n, const Iso3d& T, DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox, cudaStream_t s,
                              LaunchCounter& lc)
{
  d_bbox.reserve(8, s); h_bbox.reserve(8);
  k_bbox_init<<<1, 32, 0, s>>>((int*)d_bbox.p); LSB_LAUNCH_CHECK(); lc.add();
  if (n) { k_transform_bbox<<<std::min(div_up(n, TPB), 1184u), TPB, 0, s>>>(d_pts, (int)n, T, (int*)d_bbox.p); LSB_LAUNCH_CHECK(); lc.add(); }
  LSB_CUDA(cudaMemcpyAsync(h_bbox.p, d_bbox.p, 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
}

void dev_finish_keypoints(FinishArgs a, const IcpState* st, bool refine, DevBuf<int>& partial, DevBuf<unsigned>& counter, cudaStream_t s,
                          LaunchCounter& lc)
{
  a.b0[0] = 0;
  for (int t = 0; t < 3; ++t) a.b0[t + 1] = a.b0[t] + (int)div_up((size_t)a.n[t], TPB);
  if (a.b0[3] == 0) return;
  partial.reserve((size_t)a.b0[3] * 6, s);
  if (!counter.p) { counter.reserve(4, s); LSB_CUDA(cudaMemsetAsync(counter.p, 0, 4 * sizeof(unsigned), s)); }
  k_finish_keypoints<<<a.b0[3], TPB, 0, s>>>(a, st, refine ? 1 : 0, partial.p, counter.p); LSB_LAUNCH_CHECK(); lc.add();
}

void dev_undistort_transformed_bbox(lsb_point* d_pts, size_t n, const Interpolator* pending, const Iso3d& T, DevBuf<int>& partial,
                                    DevBuf<unsigned>& counter, DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox, int seq, cudaStream_t s,
                                    LaunchCounter& lc)
{
  if (!n) return;
  const unsigned nb = div_up(n, TPB);
  d_bbox.reserve(8, s); h_bbox.reserve(8);
  partial.reserve((size_t)nb * 6, s);
  if (!counter.p) { counter.reserve(4, s); LSB_CUDA(cudaMemsetAsync(counter.p, 0, 4 * sizeof(unsigned), s)); }
  Interpolator none;
  none.reset();
  k_undistort_transformed_bbox<<<nb, TPB, 0, s>>>(d_pts, (int)n, pending ? *pending : none, pending ? 1 : 0, T, partial.p, counter.p,
                                                  reinterpret_cast<int*>(d_bbox.p), reinterpret_cast<int*>(h_bbox.p), seq);
  LSB_LAUNCH_CHECK(); lc.add();
}

void dev_aggregate(lsb_point* d_pts, size_t n, double timeOffset, const Iso3d& T, bool applyT, cudaStream_t s, LaunchCounter& lc)
{
  if (!n) return;
  k_aggregate<<<div_up(n, TPB), TPB, 0, s>>>(d_pts, (int)n, timeOffset, T, applyT ? 1 : 0); LSB_LAUNCH_CHECK(); lc.add();
}

void dev_undistort(lsb_point* d_pts, size_t n, const Interpolator& it, double timeOffset, cudaStream_t s, LaunchCounter& lc)
{
  if (!n) return;
  k_undistort<<<div_up(n, TPB), TPB, 0, s>>>(d_pts, (int)n, it, timeOffset); LSB_LAUNCH_CHECK(); lc.add();
}

void dev_undistort_state(lsb_point* d_pts, size_t n, const IcpState* st, int mode, cudaStream_t s, LaunchCounter& lc)
{
  if (!n) return;
  k_undistort_state<<<div_up(n, TPB), TPB, 0, s>>>(d_pts, (int)n, st, mode); LSB_LAUNCH_CHECK(); lc.add();
}

void dev_icp_init(IcpState* d_state, const IcpState& init, cudaStream_t s, LaunchCounter& lc)
{
  k_icp_init<<<1, 1, 0, s>>>(d_state, init); LSB_LAUNCH_CHECK(); lc.add();
}

}  // namespace lsb

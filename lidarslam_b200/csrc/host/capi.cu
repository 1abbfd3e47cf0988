// The C ABI of liblidarslam_b200 (include/lidarslam_b200.h): thin extern "C"
// layer over lsb_slam.  Every entry point catches exceptions and returns an
// LSB_E_* code with a thread-local message; there is no CPU fallback.
#include "slam.cuh"
#include <cstring>
#include <string>

using namespace lsb;

static thread_local std::string g_last_error;

extern "C" void lsb_internal_set_error(const char* m) { g_last_error = m ? m : ""; }

template <typename F>
static int guarded(F&& f)
{
  try { return f(); }
  catch (const CudaError& e) { g_last_error = e.what(); return e.code; }
  catch (const std::exception& e) { g_last_error = e.what(); return LSB_E_INVALID; }
}

static Iso3d FromRowMajor(const double* T)
{
  Iso3d r;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) r.R(i, j) = T[i * 4 + j];
  r.t = Vec3d(T[3], T[7], T[11]);
  return r;
}
static void ToRowMajor(const Iso3d& a, double* T)
{
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) T[i * 4 + j] = a.R(i, j);
    T[i * 4 + 3] = a.t[i];
  }
  T[12] = T[13] = T[14] = 0.; T[15] = 1.;
}
// device cloud -> host buffer; returns the count
static long CopyOut(lsb_slam* h, const lsb_point* d, size_t n, lsb_point* out, long cap)
{
  if (out && n && cap > 0) {
    const size_t c = std::min<size_t>(n, (size_t)cap);
    LSB_CUDA(cudaMemcpyAsync(out, d, c * sizeof(lsb_point), cudaMemcpyDeviceToHost, h->stream));
    LSB_CUDA(cudaStreamSynchronize(h->stream));
  }
  return (long)n;
}

extern "C" {

int lsb_device_count(void)
{
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
const char* lsb_last_error(void) { return g_last_error.c_str(); }

int lsb_create(lsb_slam** out, int cuda_device)
{
  if (!out) return LSB_E_INVALID;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    g_last_error = "no CUDA device available: liblidarslam_b200 has no CPU fallback";
    return LSB_E_NO_DEVICE;
  }
  if (cuda_device < 0 || cuda_device >= n) { g_last_error = "invalid CUDA device index"; return LSB_E_INVALID; }
  return guarded([&] { *out = new lsb_slam(cuda_device); return LSB_OK; });
}
void lsb_destroy(lsb_slam* h) { delete h; }
int lsb_reset(lsb_slam* h, int reset_log) { return guarded([&] { h->Reset(reset_log != 0); return LSB_OK; }); }

int lsb_set_params(lsb_slam* h, const lsb_params* p) { h->P = *p; return LSB_OK; }
int lsb_get_params(const lsb_slam* h, lsb_params* p) { *p = h->P; return LSB_OK; }
int lsb_set_extractor_params(lsb_slam* h, int device_id, const lsb_ke_params* p) { return guarded([&] { h->Extractor(device_id).P = *p; return LSB_OK; }); }
int lsb_get_extractor_params(const lsb_slam* h, int device_id, lsb_ke_params* p)
{
  auto it = h->KeyPointsExtractors.find(device_id);
  if (it == h->KeyPointsExtractors.end() || !it->second) { g_last_error = "no extractor for this device_id"; return LSB_E_INVALID; }
  *p = it->second->P;
  return LSB_OK;
}
int lsb_set_base_to_lidar(lsb_slam* h, int device_id, const double T[16]) { h->BaseToLidarOffsets[device_id] = FromRowMajor(T); return LSB_OK; }
int lsb_set_voxel_grid_leaf_size(lsb_slam* h, int k, double size)
{
  if (k < 0 || k > 2) return LSB_E_INVALID;
  h->LocalMaps[k].LeafSize = size;   // RollingGrid::SetLeafSize is a plain SetMacro (RollingGrid.h:106)
  return LSB_OK;
}
int lsb_set_voxel_grid_size(lsb_slam* h, int size) { return guarded([&] { for (auto& m : h->LocalMaps) m.SetGridSize(size, h->stream, h->lc); return LSB_OK; }); }
int lsb_set_voxel_grid_resolution(lsb_slam* h, double r) { return guarded([&] { for (auto& m : h->LocalMaps) m.SetVoxelResolution(r, h->stream, h->lc); return LSB_OK; }); }
int lsb_set_voxel_grid_min_frames_per_voxel(lsb_slam* h, unsigned n) { for (auto& m : h->LocalMaps) m.MinFramesPerVoxel = n; return LSB_OK; }
int lsb_set_voxel_grid_decaying_threshold(lsb_slam* h, double s) { for (auto& m : h->LocalMaps) m.DecayingThreshold = s; return LSB_OK; }
int lsb_set_voxel_grid_sampling_mode(lsb_slam* h, int k, int mode)
{
  if (k < 0 || k > 2) return LSB_E_INVALID;
  h->LocalMaps[k].Sampling = mode;
  return LSB_OK;
}

int lsb_add_frames(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  return guarded([&] { return h->AddFrames(n_frames, frames, n, stamp_us, seq, false); });
}
int lsb_add_frames_dev(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  return guarded([&] { return h->AddFrames(n_frames, frames, n, stamp_us, seq, true); });
}

void* lsb_host_alloc(size_t bytes) { void* p = nullptr; if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; } return p; }
void lsb_host_free(void* p) { if (p) cudaFreeHost(p); }
void* lsb_dev_alloc(lsb_slam* h, size_t bytes)
{
  void* p = nullptr;
  cudaSetDevice(h->device);
  if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
  return p;
}
void lsb_dev_free(lsb_slam* h, void* p) { cudaSetDevice(h->device); if (p) cudaFree(p); }
int lsb_dev_upload(lsb_slam* h, void* dst, const void* src, size_t bytes)
{
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    LSB_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, h->stream));
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}
int lsb_dev_flush_l2(lsb_slam* h)
{
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    const size_t bytes = 256u << 20;
    h->tmp_u8.reserve(bytes, h->stream);
    LSB_CUDA(cudaMemsetAsync(h->tmp_u8.p, (int)(h->NbrFrameProcessed & 0xff), bytes, h->stream));
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return LSB_OK;
  });
}

int lsb_get_world_transform(const lsb_slam* h, double T[16], double* time)
{
  if (h->LogTrajectory.empty()) { ToRowMajor(Iso3d::identity(), T); if (time) *time = 0.; return LSB_OK; }
  ToRowMajor(h->LogTrajectory.back().T, T);
  if (time) *time = h->LogTrajectory.back().time;
  return LSB_OK;
}
int lsb_get_latency_compensated_world_transform(const lsb_slam* h, double T[16], double* time)
{
  const size_t n = h->LogTrajectory.size();
  if (n < 2) return lsb_get_world_transform(h, T, time);
  const auto& prev = h->LogTrajectory[n - 2];
  const auto& cur = h->LogTrajectory[n - 1];
  if (time) *time = cur.time;
  if (std::abs(cur.time - prev.time) < 1e-6 || std::abs(h->Latency / (cur.time - prev.time)) > h->P.max_extrapolation_ratio) {
    ToRowMajor(cur.T, T);
    return LSB_W_EXTRAPOLATION;
  }
  ToRowMajor(linear_interpolation(prev.T, cur.T, cur.time + h->Latency, prev.time, cur.time), T);
  return LSB_OK;
}
int lsb_get_relative_transform(const lsb_slam* h, double T[16]) { ToRowMajor(h->Trelative, T); return LSB_OK; }
int lsb_get_covariance(const lsb_slam* h, double c[36]) { std::memcpy(c, h->LocalizationUncertainty.covariance, 36 * sizeof(double)); return LSB_OK; }
int lsb_get_registration_error(const lsb_slam* h, lsb_registration_error* e) { *e = h->LocalizationUncertainty; return LSB_OK; }
long lsb_get_trajectory(const lsb_slam* h, double* Ts, double* times, long cap)
{
  const long n = (long)h->LogTrajectory.size();
  for (long i = 0; i < std::min(n, cap); ++i) {
    if (Ts) ToRowMajor(h->LogTrajectory[i].T, Ts + 16 * i);
    if (times) times[i] = h->LogTrajectory[i].time;
  }
  return n;
}

long lsb_get_keypoints(lsb_slam* h, int type, int world, lsb_point* out, long cap)
{
  if (type < 0 || type > 2) return LSB_E_INVALID;
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    if (!world) return (int)CopyOut(h, h->CurrentKeypoints[type].buf.p, h->CurrentKeypoints[type].n, out, cap);
    const DCloud& w = h->WorldKeypoints(type);
    return (int)CopyOut(h, w.buf.p, w.n, out, cap);
  });
}
long lsb_get_map(lsb_slam* h, int type, int clean, lsb_point* out, long cap)
{
  if (type < 0 || type > 2) return LSB_E_INVALID;
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    RollingGridGpu& m = h->LocalMaps[type];
    if (!clean) return (int)CopyOut(h, m.pts.p, m.n_entries, out, cap);
    h->tmp_cloud.reserve(std::max<size_t>(m.n_entries, 1), h->stream);
    const size_t n = m.Get(true, h->tmp_cloud.p, h->stream, h->lc);
    return (int)CopyOut(h, h->tmp_cloud.p, n, out, cap);
  });
}
long lsb_get_submap(lsb_slam* h, int type, lsb_point* out, long cap)
{
  if (type < 0 || type > 2) return LSB_E_INVALID;
  return guarded([&] { LSB_CUDA(cudaSetDevice(h->device)); return (int)CopyOut(h, h->LocalMaps[type].submap.p, h->LocalMaps[type].n_submap, out, cap); });
}
long lsb_get_registered_frame(lsb_slam* h, lsb_point* out, long cap)
{
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    const size_t n = h->RegisteredFrame(h->tmp_cloud);
    return (int)CopyOut(h, h->tmp_cloud.p, n, out, cap);
  });
}
long lsb_get_match_debug(lsb_slam* h, int stage, int type, uint8_t* status, double* weight, long cap)
{
  if (type < 0 || type > 2 || stage < 0 || stage > 1) return LSB_E_INVALID;
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    lsb_slam::StageDebug& d = stage == 0 ? h->EgoDebug : h->LocDebug;
    const uint8_t* st = stage == 0 ? d.status.p : h->mw.status.p;
    const double* wt = stage == 0 ? d.weight.p : h->mw.weight.p;
    const size_t n = d.off[type + 1] - d.off[type];
    const size_t c = std::min<size_t>(n, cap > 0 ? (size_t)cap : 0);
    if (c && status) LSB_CUDA(cudaMemcpyAsync(status, st + d.off[type], c, cudaMemcpyDeviceToHost, h->stream));
    if (c && weight) LSB_CUDA(cudaMemcpyAsync(weight, wt + d.off[type], c * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
    LSB_CUDA(cudaStreamSynchronize(h->stream));
    return (int)n;
  });
}
long lsb_get_ke_debug(lsb_slam* h, int device_id, int field, float* out, long cap)
{
  auto it = h->KeyPointsExtractors.find(device_id);
  if (it == h->KeyPointsExtractors.end() || !it->second) { g_last_error = "no extractor for this device_id"; return LSB_E_INVALID; }
  return guarded([&] {
    LSB_CUDA(cudaSetDevice(h->device));
    ExtractorGpu& ke = *it->second;
    if (out && ke.n && cap > 0) {
      h->tmp_f.reserve(ke.n, h->stream);
      ke.debug_field(field, h->tmp_f.p, h->stream, h->lc);
      LSB_CUDA(cudaMemcpyAsync(out, h->tmp_f.p, std::min<size_t>(ke.n, (size_t)cap) * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
      LSB_CUDA(cudaStreamSynchronize(h->stream));
    }
    return (int)ke.n;
  });
}
int lsb_get_frame_stats(const lsb_slam* h, lsb_frame_stats* s) { *s = h->Stats; return LSB_OK; }
int lsb_set_profiling(lsb_slam* h, int on)
{
  h->mw.t_knn.enabled = h->mw.t_fit.enabled = h->mw.t_lm.enabled = (on != 0);
  h->serial_types = (on != 0);   // kernel durations are taken without overlap between the keypoint types
  return LSB_OK;
}
float lsb_get_overlap(const lsb_slam* h) { return h->OverlapEstimation; }
int lsb_get_comply_motion_limits(const lsb_slam* h) { return h->MotionLimits.Comply ? 1 : 0; }
unsigned int lsb_get_total_matched_keypoints(const lsb_slam* h) { return h->TotalMatchedKeypoints; }

int lsb_set_world_transform_from_guess(lsb_slam* h, const double T[16]) { h->SetWorldTransformFromGuess(FromRowMajor(T)); return LSB_OK; }
int lsb_clear_maps(lsb_slam* h) { return guarded([&] { h->ClearMaps(); return LSB_OK; }); }

}  // extern "C"

// The C ABI of liblidarslam_b200 (include/lidarslam_b200.h): thin extern "C"
// layer over lsb_slam.  Every entry point catches exceptions and returns an
// LSB_E_* code with a thread-local message; there is no CPU fallback.
#include "slam.cuh"
#include "pcd_io.h"
#include <cstring>
#include <cstdlib>
#include <ctime>
#include <string>

using namespace lsb;

static thread_local std::string g_last_error;

extern "C" void lsb_internal_set_error(const char* m) { g_last_error = m ? m : ""; }

template <typename F>
static int guarded(F&& f)
{
  try { return f(); }
  catch (const CudaError& e) { g_last_error = e.what(); return e.code; }
  catch (const std::bad_alloc&) { g_last_error = "out of host memory"; return LSB_E_CAPACITY; }
  catch (const std::exception& e) { g_last_error = e.what(); return LSB_E_INVALID; }
  catch (...) { g_last_error = "unknown exception"; return LSB_E_INVALID; }
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

// A handle drives four streams (one per keypoint type + its own).  The CUDA default of 8 hardware work queues makes the
// streams of several handles share queues, which serialises independent sequences (8 OS-128 sequences on one B200: 731
// frames/s with 8 queues, 1201 with 32, tools/probes/batch_probe.py).  The variable is read when the CUDA context is created,
// so it is set (never overriding the user's value) before the library's first CUDA call; a process that initialised CUDA
// earlier has to export it itself.
static void RequestWorkQueues()
{
  static const bool once = [] { setenv("CUDA_DEVICE_MAX_CONNECTIONS", "32", /*overwrite=*/0); return true; }();
  (void)once;
}

extern "C" {

int lsb_device_count(void)
{
  RequestWorkQueues();
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
const char* lsb_last_error(void) { return g_last_error.c_str(); }

int lsb_create(lsb_slam** out, int cuda_device)
{
  if (!out) return LSB_E_INVALID;
  *out = nullptr;
  RequestWorkQueues();
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    g_last_error = "no CUDA device available: liblidarslam_b200 has no CPU fallback";
    return LSB_E_NO_DEVICE;
  }
  if (cuda_device < 0 || cuda_device >= n) { g_last_error = "invalid CUDA device index"; return LSB_E_INVALID; }
  return guarded([&] { *out = new lsb_slam(cuda_device); return LSB_OK; });
}
void lsb_destroy(lsb_slam* h)
{
  if (h) guarded([&] { h->Quiesce(true); return LSB_OK; });   // an error of a job nobody waited for dies with the handle
  delete h;
}
int lsb_reset(lsb_slam* h, int reset_log) { return guarded([&] { h->Quiesce(true); h->Reset(reset_log != 0); return LSB_OK; }); }

// Entry points that read or change what the helper thread works on (maps, keypoints, extractors, statistics) wait for it
// first; setters also drop inputs staged with the old parameters.
int lsb_set_params(lsb_slam* h, const lsb_params* p) { return guarded([&] { h->Quiesce(true); h->P = *p; return LSB_OK; }); }
int lsb_get_params(const lsb_slam* h, lsb_params* p) { *p = h->P; return LSB_OK; }
int lsb_set_extractor_params(lsb_slam* h, int device_id, const lsb_ke_params* p)
{
  return guarded([&] {
    h->Quiesce(true);
    ExtractorGpu& ke = h->Extractor(device_id);
    // An azimuthal resolution outside the usable range [1e-6, pi/4] means "estimate it from the first frame"
    // (SpinningSensorKeypointExtractor.cxx:165-170).  A caller that pushes its whole parameter block again (the
    // C++ mirror does so before every frame) must not throw the estimate away: the reference keeps it for the
    // life of the extractor (:169), so an out-of-range incoming value leaves a usable stored one in place.
    const float est = ke.P.azimuthal_resolution;
    ke.P = *p;
    const bool incoming_unset = !(p->azimuthal_resolution >= 1e-6f && p->azimuthal_resolution <= 0.78539816339744830962f);
    const bool stored_usable = est >= 1e-6f && est <= 0.78539816339744830962f;
    if (incoming_unset && stored_usable) ke.P.azimuthal_resolution = est;
    return LSB_OK;
  });
}
int lsb_get_extractor_params(const lsb_slam* h, int device_id, lsb_ke_params* p)
{
  // no join of deferred / staged work (the C++ mirror reads the azimuth estimate back after every frame): the extractor state
  // changes hands under ke_mu, so the copy is consistent; with an input staged ahead it may already hold that input's estimate
  std::lock_guard<std::mutex> g(const_cast<lsb_slam*>(h)->ke_mu);
  auto it = h->KeyPointsExtractors.find(device_id);
  if (it == h->KeyPointsExtractors.end() || !it->second) { g_last_error = "no extractor for this device_id"; return LSB_E_INVALID; }
  *p = it->second->P;
  return LSB_OK;
}
int lsb_set_base_to_lidar(lsb_slam* h, int device_id, const double T[16])
{
  return guarded([&] { h->Quiesce(true); h->BaseToLidarOffsets[device_id] = FromRowMajor(T); return LSB_OK; });
}
int lsb_set_voxel_grid_leaf_size(lsb_slam* h, int k, double size)
{
  if (k < 0 || k > 2) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    h->LocalMaps[k].LeafSize = size;   // RollingGrid::SetLeafSize is a plain SetMacro (RollingGrid.h:106)
    return LSB_OK;
  });
}
int lsb_set_voxel_grid_size(lsb_slam* h, int size) { return guarded([&] { h->Quiesce(false); for (auto& m : h->LocalMaps) m.SetGridSize(size, h->stream, h->lc); return LSB_OK; }); }
int lsb_set_voxel_grid_resolution(lsb_slam* h, double r) { return guarded([&] { h->Quiesce(false); for (auto& m : h->LocalMaps) m.SetVoxelResolution(r, h->stream, h->lc); return LSB_OK; }); }
int lsb_set_voxel_grid_min_frames_per_voxel(lsb_slam* h, unsigned n)
{
  return guarded([&] { h->Quiesce(false); for (auto& m : h->LocalMaps) m.MinFramesPerVoxel = n; return LSB_OK; });
}
int lsb_set_voxel_grid_decaying_threshold(lsb_slam* h, double s)
{
  return guarded([&] { h->Quiesce(false); for (auto& m : h->LocalMaps) m.DecayingThreshold = s; return LSB_OK; });
}
int lsb_set_voxel_grid_sampling_mode(lsb_slam* h, int k, int mode)
{
  if (k < 0 || k > 2) return LSB_E_INVALID;
  if (mode < LSB_SAMPLING_FIRST || mode > LSB_SAMPLING_CENTROID) { g_last_error = "unknown SamplingMode"; return LSB_E_INVALID; }   // rejected here, not in the middle of a frame
  return guarded([&] { h->Quiesce(false); h->LocalMaps[k].Sampling = mode; return LSB_OK; });
}

int lsb_add_frames(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  return guarded([&] { return h->AddFrames(n_frames, frames, n, stamp_us, seq, false); });
}
int lsb_add_frames_dev(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  return guarded([&] { return h->AddFrames(n_frames, frames, n, stamp_us, seq, true); });
}
int lsb_prefetch_frames(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  return guarded([&] { return h->PrefetchFrames(n_frames, frames, n, stamp_us, seq, false); });
}
int lsb_prefetch_frames_dev(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  return guarded([&] { return h->PrefetchFrames(n_frames, frames, n, stamp_us, seq, true); });
}
int lsb_set_pipelining(lsb_slam* h, int defer_map_update)
{
  return guarded([&] { h->Quiesce(false); h->defer_map_update = defer_map_update != 0; return LSB_OK; });
}
int lsb_synchronize(lsb_slam* h) { return guarded([&] { h->SynchronizeAll(); return LSB_OK; }); }
int lsb_timer_begin(lsb_slam* h)
{
  return guarded([&] {
    h->SynchronizeAll();
    LSB_CUDA(cudaEventRecord(h->ev_timer[0], h->stream));
    return LSB_OK;
  });
}
int lsb_timer_end(lsb_slam* h, float* ms)
{
  return guarded([&] {
    h->Quiesce(false);                                   // helper jobs over; `stream` already waits for a deferred map update
    for (auto& sl : h->staged) if (sl.valid) LSB_CUDA(cudaStreamWaitEvent(h->stream, sl.ready, 0));
    LSB_CUDA(cudaEventRecord(h->ev_timer[1], h->stream));
    LSB_CUDA(cudaEventSynchronize(h->ev_timer[1]));
    float t = 0.f;
    LSB_CUDA(cudaEventElapsedTime(&t, h->ev_timer[0], h->ev_timer[1]));
    if (ms) *ms = t;
    return LSB_OK;
  });
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
    h->Quiesce(false);
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
int lsb_get_covariance(const lsb_slam* h, double c[36]) { std::memcpy(c, h->Uncertainty().covariance, 36 * sizeof(double)); return LSB_OK; }
int lsb_get_registration_error(const lsb_slam* h, lsb_registration_error* e) { *e = h->Uncertainty(); return LSB_OK; }
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
    h->Quiesce(false);
    if (!world) return (int)CopyOut(h, h->CurrentKeypoints[type].buf.p, h->CurrentKeypoints[type].n, out, cap);
    const DCloud& w = h->WorldKeypoints(type);
    return (int)CopyOut(h, w.buf.p, w.n, out, cap);
  });
}
long lsb_get_map(lsb_slam* h, int type, int clean, lsb_point* out, long cap)
{
  if (type < 0 || type > 2) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
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
  return guarded([&] { h->Quiesce(false); return (int)CopyOut(h, h->LocalMaps[type].submap.p, h->LocalMaps[type].n_submap, out, cap); });
}
long lsb_get_registered_frame(lsb_slam* h, lsb_point* out, long cap)
{
  return guarded([&] {
    h->Quiesce(false);
    const size_t n = h->RegisteredFrame(h->tmp_cloud);
    return (int)CopyOut(h, h->tmp_cloud.p, n, out, cap);
  });
}
long lsb_get_match_debug(lsb_slam* h, int stage, int type, uint8_t* status, double* weight, long cap)
{
  if (type < 0 || type > 2 || stage < 0 || stage > 1) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
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
  { const int rc = guarded([&] { h->Quiesce(false); return LSB_OK; }); if (rc < 0) return rc; }
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
int lsb_get_frame_stats(const lsb_slam* h, lsb_frame_stats* s)
{
  return guarded([&] { const_cast<lsb_slam*>(h)->Quiesce(false); *s = h->Stats; s->launches_total = h->lc.n; return LSB_OK; });   // a deferred map update completes ms_map_update / n_map
}
int lsb_set_profiling(lsb_slam* h, int on)
{
  { const int rc = guarded([&] { h->Quiesce(false); return LSB_OK; }); if (rc < 0) return rc; }
  h->mw.t_knn.enabled = h->mw.t_fit.enabled = h->mw.t_lm.enabled = (on != 0);
  h->serial_types = (on != 0);   // kernel durations are taken without overlap between the keypoint types
  return LSB_OK;
}
unsigned int lsb_get_azimuth_estimations(const lsb_slam* h, int device_id)
{
  std::lock_guard<std::mutex> g(const_cast<lsb_slam*>(h)->ke_mu);
  auto it = h->KeyPointsExtractors.find(device_id);
  return (it == h->KeyPointsExtractors.end() || !it->second) ? 0u : it->second->az_estimations;
}
float lsb_get_overlap(const lsb_slam* h) { return h->OverlapEstimation; }
int lsb_get_comply_motion_limits(const lsb_slam* h) { return h->MotionLimits.Comply ? 1 : 0; }
unsigned int lsb_get_total_matched_keypoints(const lsb_slam* h) { return h->TotalMatchedKeypoints; }

int lsb_set_world_transform_from_guess(lsb_slam* h, const double T[16])
{
  return guarded([&] { h->Quiesce(false); h->SetWorldTransformFromGuess(FromRowMajor(T)); return LSB_OK; });
}
int lsb_clear_maps(lsb_slam* h) { return guarded([&] { h->Quiesce(false); h->ClearMaps(); return LSB_OK; }); }

/* ---- external sensor constraints ------------------------------------ */
int lsb_add_wheel_odom_measurement(lsb_slam* h, const lsb_wheel_odom_measurement* m)
{
  if (!m) return LSB_E_INVALID;
  WheelOdomMeasurement w;
  w.Time = m->time; w.Distance = m->distance;
  h->WheelOdomManager.AddMeasurement(w);
  return LSB_OK;
}
int lsb_add_gravity_measurement(lsb_slam* h, const lsb_gravity_measurement* m)
{
  if (!m) return LSB_E_INVALID;
  GravityMeasurement g;
  g.Time = m->time;
  for (int i = 0; i < 3; ++i) g.Acceleration[i] = m->acceleration[i];
  h->ImuManager.AddMeasurement(g);
  return LSB_OK;
}
int lsb_clear_sensor_measurements(lsb_slam* h) { h->WheelOdomManager.Reset(); h->ImuManager.Reset(); return LSB_OK; }
int lsb_set_wheel_odom_weight(lsb_slam* h, double w) { h->WheelOdomManager.Weight = w; return LSB_OK; }
double lsb_get_wheel_odom_weight(const lsb_slam* h) { return h->WheelOdomManager.Weight; }
int lsb_set_gravity_weight(lsb_slam* h, double w) { h->ImuManager.Weight = w; return LSB_OK; }
double lsb_get_gravity_weight(const lsb_slam* h) { return h->ImuManager.Weight; }
int lsb_set_sensor_time_offset(lsb_slam* h, double t) { h->WheelOdomManager.TimeOffset = t; h->ImuManager.TimeOffset = t; return LSB_OK; }
double lsb_get_sensor_time_offset(const lsb_slam* h) { return h->ImuManager.TimeOffset; }
int lsb_get_gravity_reference(const lsb_slam* h, double g[3]) { for (int i = 0; i < 3; ++i) g[i] = h->ImuManager.GravityRef[i]; return LSB_OK; }

/* ---- map persistence: PCD wire format ----------------------------- */
static const char* const kMapFileNames[3] = {"edges.pcd", "planes.pcd", "blobs.pcd"};   // Plural(KeypointTypeNames[k]) + ".pcd" (Enums.h:39, Utilities.cxx:95-98)

int lsb_pcd_write(const char* path, const lsb_point* pts, size_t n, int pcd_format)
{
  if (!path || (n && !pts)) return LSB_E_INVALID;
  int rc = -1;
  try { rc = pcd::Write(path, pts, n, pcd_format); } catch (const std::exception& e) { g_last_error = e.what(); return -1; }
  if (rc == -1) g_last_error = std::string("cannot write ") + path;
  return rc;
}
long lsb_pcd_read(const char* path, lsb_point* out, long cap)
{
  if (!path) return LSB_E_INVALID;
  std::vector<lsb_point> pts;
  std::string why;
  try {
    if (pcd::Read(path, pts, &why) != 0) { g_last_error = why; return -1; }
  } catch (const std::exception& e) { g_last_error = e.what(); return -1; }
  if (out && cap > 0 && !pts.empty()) std::memcpy(out, pts.data(), std::min<size_t>(pts.size(), (size_t)cap) * sizeof(lsb_point));
  return (long)pts.size();
}
int lsb_save_maps_to_pcd(lsb_slam* h, const char* file_prefix, int pcd_format, int filtered)
{
  if (!file_prefix || pcd_format < 0 || pcd_format > 2) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    int worst = LSB_OK;
    std::vector<lsb_point> host;
    for (int k = 0; k < 3; ++k) {
      if (!h->P.use_keypoints[k]) continue;                       // Slam.cxx:511
      RollingGridGpu& m = h->LocalMaps[k];
      const lsb_point* d = m.pts.p;
      size_t n = m.n_entries;
      if (filtered) {                                             // GetMap(k, clean) Slam.cxx:670-677
        h->tmp_cloud.reserve(std::max<size_t>(m.n_entries, 1), h->stream);
        n = m.Get(true, h->tmp_cloud.p, h->stream, h->lc);
        d = h->tmp_cloud.p;
      }
      host.resize(n);
      CopyOut(h, d, n, host.data(), (long)n);
      const int rc = pcd::Write(std::string(file_prefix) + kMapFileNames[k], host.data(), n, pcd_format);
      if (rc == -1) { g_last_error = std::string("cannot write ") + file_prefix + kMapFileNames[k]; worst = LSB_E_INVALID; }
      // rc == -3 (empty map): nothing written, as savePointCloudToPCD (PointCloudStorage.h:91-92)
    }
    return worst;
  });
}
int lsb_load_maps_from_pcd(lsb_slam* h, const char* file_prefix, int reset_maps)
{
  if (!file_prefix) return LSB_E_INVALID;
  return guarded([&] {
    h->Quiesce(false);
    if (reset_maps) h->ClearMaps();                               // Slam.cxx:526-527
    int loaded = 0;
    std::vector<lsb_point> host;
    for (int k = 0; k < 3; ++k) {
      std::string why;
      if (pcd::Read(std::string(file_prefix) + kMapFileNames[k], host, &why) != 0) continue;   // Slam.cxx:531 a missing file is skipped
      const bool fixed = h->P.map_update == LSB_MAP_NONE || h->P.map_update == LSB_MAP_ADD_KPTS_TO_FIXED_MAP;   // Slam.cxx:536
      h->tmp_cloud.reserve(std::max<size_t>(host.size(), 1), h->stream);
      if (!host.empty()) LSB_CUDA(cudaMemcpyAsync(h->tmp_cloud.p, host.data(), host.size() * sizeof(lsb_point), cudaMemcpyHostToDevice, h->stream));
      h->LocalMaps[k].Add(h->tmp_cloud.p, host.size(), fixed, (double)std::time(nullptr), true, h->stream, h->lc);   // Slam.cxx:537
      LSB_CUDA(cudaStreamSynchronize(h->stream));
      loaded |= 1 << k;
    }
    return loaded;
  });
}

}  // extern "C"

// lsb_slam — host driver of the per-frame pipeline.  Mirrors the control flow of
// LidarSlam::Slam (reference slam_lib/src/Slam.cxx, line ranges per function);
// every cloud operation is a kernel on the handle's stream.
#include "slam.cuh"
#include "../common/lcp_core.h"
#include <cstring>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cmath>
#include <cfloat>
#include <limits>
#include <algorithm>

using namespace lsb;

// Enqueue order of the per-type work: the type with the most keypoints first, so that the longest
// chain of a fork/join section starts first (the results do not depend on the order).
static const int kTypeOrder[3] = {LSB_BLOB, LSB_PLANE, LSB_EDGE};
// Matching passes of an ICP iteration are enqueued by one thread: the edge chain (cube pass, warp pass, brute-force tail, RANSAC
// fit: four dependent latency-bound launches) is the longest after the blobs', so it goes second (CUPTI timeline of a frame:
// with the planes in between, the edges' first kernel started 54 us after the blobs' and ended the iteration 26 us late).
static const int kMatchOrder[3] = {LSB_BLOB, LSB_EDGE, LSB_PLANE};

static inline double StampToSec(uint64_t us) { return us * 1e-6; }   // Utils::PclStampToSec

lsb_slam::lsb_slam(int dev) : device(dev)
{
  LSB_CUDA(cudaSetDevice(dev));
  // the frame's own streams run at the highest priority, the ingest stream (prefetched frames) at the lowest: a kernel of
  // the current frame's ICP chain is scheduled ahead of the next frame's extraction blocks (LSB_MAIN_PRIORITY: tuning hook)
  int prio_lo = 0, prio_hi = 0;
  LSB_CUDA(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
  int prio_main = prio_hi;
  if (const char* e = std::getenv("LSB_MAIN_PRIORITY")) prio_main = std::atoi(e);
  LSB_CUDA(cudaStreamCreateWithPriority(&stream, cudaStreamNonBlocking, prio_main));
  LSB_CUDA(cudaStreamCreateWithPriority(&istream, cudaStreamNonBlocking, prio_lo));
  {  // keep freed blocks in the device's default pool: buffer growth must not hit the driver again
    cudaMemPool_t pool;
    LSB_CUDA(cudaDeviceGetDefaultMemPool(&pool, dev));
    unsigned long long keep = ~0ull;
    LSB_CUDA(cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep));
    // head-room: every new handle makes sure the pool holds LSB_POOL_PREWARM_MB (default 4096) of
    // unused memory, by taking one block from the driver and handing it back to the pool.  The maps,
    // grids and sort scratch of a sequence then grow out of the pool: a driver allocation in the
    // middle of a frame was measured at 5 - 125 ms on B200 (profiles/r01_e2e_stalls.txt), against
    // 2.3 ms for the whole frame.  0 = no reservation.
    size_t mb = 4096;
    if (const char* e = std::getenv("LSB_POOL_PREWARM_MB")) mb = (size_t)std::atol(e);
    unsigned long long reserved = 0, used = 0;
    LSB_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &reserved));
    LSB_CUDA(cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used));
    if (mb && reserved - used < ((unsigned long long)mb << 20)) {
      void* blk = nullptr;
      const cudaError_t rc = cudaMallocAsync(&blk, mb << 20, stream);
      if (rc == cudaSuccess) cudaFreeAsync(blk, stream);
      else cudaGetLastError();
      LSB_CUDA(cudaStreamSynchronize(stream));
      if (const char* e = std::getenv("LSB_TRACE")) if (e[0] == '1') {
        unsigned long long r2 = 0, u2 = 0;
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrReservedMemCurrent, &r2);
        cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &u2);
        std::fprintf(stderr, "[lsb pool] reserved %llu used %llu MiB -> top-up %zu MiB rc=%d -> reserved %llu used %llu MiB\n",
                     reserved >> 20, used >> 20, mb, (int)rc, r2 >> 20, u2 >> 20);
      }
    }
  }
  for (auto& e : ev) LSB_CUDA(cudaEventCreate(&e));
  {
    // LSB_TYPE_PRIORITIES="e,p,b": CUDA stream priority of the three type streams (0 = default, negative = higher); tuning hook
    int prio[3] = {prio_main, prio_main, prio_main};
    if (const char* e = std::getenv("LSB_TYPE_PRIORITIES")) std::sscanf(e, "%d,%d,%d", &prio[0], &prio[1], &prio[2]);
    for (int k = 0; k < 3; ++k) LSB_CUDA(cudaStreamCreateWithPriority(&tstream[k], cudaStreamNonBlocking, prio[k]));
  }
  LSB_CUDA(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
  for (auto& e : ev_join) LSB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  for (auto& e : ev_map_end) LSB_CUDA(cudaEventCreate(&e));
  for (auto& e : ev_timer) LSB_CUDA(cudaEventCreate(&e));
  for (auto& sl : staged) LSB_CUDA(cudaEventCreateWithFlags(&sl.ready, cudaEventDisableTiming));
  if (const char* e = std::getenv("LSB_DEFER_MAP")) defer_map_update = (e[0] != '0');
  lsb_default_params(&P);
  if (const char* e = std::getenv("LSB_TRACE")) trace.on = (e[0] == '1');
  if (const char* e = std::getenv("LSB_HOST_LANES")) use_lanes = (e[0] != '0');
  KeyPointsExtractors[0].reset(new ExtractorGpu());
  // Slam.cxx:153-157: resolution is snapped with the leaf size at call time (0.2 -> 10.0)
  for (auto& m : LocalMaps) m.SetVoxelResolution(10., stream, lc);
  for (auto& m : LocalMaps) m.SetGridSize(50, stream, lc);
  LocalMaps[LSB_EDGE].LeafSize = 0.30;
  LocalMaps[LSB_PLANE].LeafSize = 0.60;
  LocalMaps[LSB_BLOB].LeafSize = 0.30;
  // kNN cell edge relative to the leaf size (density of each map differs)
  LocalMaps[LSB_EDGE].knn_cell_factor = 5.0f;
  LocalMaps[LSB_PLANE].knn_cell_factor = 2.5f;
  LocalMaps[LSB_BLOB].knn_cell_factor = 2.5f;
  if (const char* e = std::getenv("LSB_KNN_CELL_FACTORS")) {   // tuning hook: "edge,plane,blob"
    float f[3];
    if (std::sscanf(e, "%f,%f,%f", &f[0], &f[1], &f[2]) == 3)
      for (int k = 0; k < 3; ++k) if (f[k] > 0.f) LocalMaps[k].knn_cell_factor = f[k];
  }
  Reset(true);
}

lsb_slam::~lsb_slam()
{
  if (worker) { worker->drain(); worker.reset(); }
  for (auto& w : tworker) if (w) { w->drain(); w.reset(); }
  cudaSetDevice(device);
  for (auto& t : tstream) if (t) cudaStreamSynchronize(t);
  if (istream) cudaStreamSynchronize(istream);
  if (stream) cudaStreamSynchronize(stream);
  for (auto& e : ev_map_end) if (e) cudaEventDestroy(e);
  for (auto& e : ev_timer) if (e) cudaEventDestroy(e);
  for (auto& sl : staged) if (sl.ready) cudaEventDestroy(sl.ready);
  if (istream) cudaStreamDestroy(istream);
  for (auto& e : ev) if (e) cudaEventDestroy(e);
  if (ev_fork) cudaEventDestroy(ev_fork);
  for (auto& e : ev_join) if (e) cudaEventDestroy(e);
  for (auto& t : tstream) if (t) cudaStreamDestroy(t);
  if (stream) cudaStreamDestroy(stream);
}

void lsb_slam::Trace::mark(const char* name, cudaStream_t s)
{
  if (!on) return;
  if (used == evs.size()) { cudaEvent_t e; LSB_CUDA(cudaEventCreate(&e)); evs.push_back(e); names.push_back(name); host_us.push_back(0.); }
  names[used] = name;
  host_us[used] = std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now().time_since_epoch()).count();
  LSB_CUDA(cudaEventRecord(evs[used++], s));
}

void lsb_slam::Trace::dump()
{
  if (!on || used < 2) { used = 0; return; }
  std::fprintf(stderr, "[lsb trace]");
  for (size_t i = 1; i < used; ++i) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, evs[i - 1], evs[i]);
    std::fprintf(stderr, " %s=%.0f(h%.0f)", names[i], ms * 1e3f, host_us[i] - host_us[i - 1]);
  }
  std::fprintf(stderr, "\n");
  used = 0;
}

void lsb_slam::Fork()
{
  if (serial_types) return;
  LSB_CUDA(cudaEventRecord(ev_fork, stream));
  for (auto& t : tstream) LSB_CUDA(cudaStreamWaitEvent(t, ev_fork, 0));
}

// The per-type helper threads cost busy host threads per handle in flight.  They pay while every handle has three cores to
// itself; beyond that (many sequences on one GPU, or many ranks on one host: torchrun exports LOCAL_WORLD_SIZE) the
// spinning helpers steal the cores the other handles' callers need (8 sequences on 16 cores: 298 frames/s with lanes,
// 856 without), so the types are then enqueued from the caller's thread.
static std::atomic<int> g_handles_in_flight{0};
static int HostThreadBudget()
{
  static const int budget = [] {
    int hw = (int)std::thread::hardware_concurrency();
    if (hw <= 0) hw = 4;
    int procs = 1;
    if (const char* e = std::getenv("LOCAL_WORLD_SIZE")) procs = std::max(1, std::atoi(e));
    return std::max(1, hw / procs);
  }();
  return budget;
}
// Host cores against busy host threads.  A handle in flight keeps up to five threads busy (the caller, one helper per keypoint
// type, one for staged inputs).  0 = room for all of them, the helpers spin between jobs (frames arrive back to back);
// 1 = at least three cores per handle: the helpers are used but sleep when idle (a wake-up costs a few microseconds per job);
// 2 = fewer: everything on the caller's thread, nothing deferred.
static int HostMode()
{
  const int h = std::max(1, g_handles_in_flight.load(std::memory_order_relaxed)), budget = HostThreadBudget();
  static const int forced = [] { const char* e = std::getenv("LSB_HOST_MODE"); return e ? std::atoi(e) : -1; }();   // tuning hook
  if (forced >= 0) return forced;
  return budget >= 5 * h ? 0 : (budget >= 3 * h ? 1 : 2);
}
struct InFlightGuard {
  InFlightGuard() { g_handles_in_flight.fetch_add(1, std::memory_order_relaxed); }
  ~InFlightGuard() { g_handles_in_flight.fetch_sub(1, std::memory_order_relaxed); }
};

void lsb_slam::ForEachType(const std::function<void(int)>& fn, const std::function<void()>& meanwhile)
{
  const int mode = HostMode();
  if (serial_types || !use_lanes || mode == 2) {
    JoinMapUpdate();             // a deferred update of these maps ends before the caller's thread touches them
    Fork();
    if (meanwhile) meanwhile();
    for (int k : kTypeOrder) fn(k);
    Join();
    return;
  }
  Fork();
  unsigned ticket[3];
  for (int k : kTypeOrder) ticket[k] = TypeWorker(k).post([&fn, k] { fn(k); }, mode == 0);
  std::exception_ptr first;
  if (meanwhile) { try { meanwhile(); } catch (...) { first = std::current_exception(); } }   // caller's own work on `stream`, beside the helpers
  for (int k : kTypeOrder) {
    try { TypeWorker(k).wait(ticket[k]); } catch (...) { if (!first) first = std::current_exception(); }
  }
  if (first) std::rethrow_exception(first);
  JoinMapUpdate();               // its jobs ran ahead of these on the same threads: only the bookkeeping (and its errors) is left
  Join();
}

void lsb_slam::Join()
{
  if (serial_types) return;
  for (int k = 0; k < 3; ++k) {
    LSB_CUDA(cudaEventRecord(ev_join[k], tstream[k]));
    LSB_CUDA(cudaStreamWaitEvent(stream, ev_join[k], 0));
  }
}

void lsb_slam::ClearMaps()
{
  const float zero[3] = {0.f, 0.f, 0.f};
  for (auto& m : LocalMaps) m.Reset(zero, stream);
}

void lsb_slam::Reset(bool resetLog)
{
  ClearMaps();
  KfLastPose = Iso3d::identity();
  KfCounter = 0;
  Tworld = PreviousTworld = Trelative = Iso3d::identity();
  WithinFrameMotion.reset();
  std::memset(&LocalizationUncertainty, 0, sizeof(LocalizationUncertainty));
  UncertaintyPending = false;
  NbCurrentFrames = 0;
  CurrentStamp = 0;
  for (int k = 0; k < 3; ++k) {
    CurrentKeypoints[k].n = 0; PreviousKeypoints[k].n = 0; CurrentWorldKeypoints[k].n = 0;
    WorldKeypointsValid[k] = false;
  }
  for (StageDebug* d : {&EgoDebug, &LocDebug}) {
    for (auto& o : d->off) o = 0;
    std::memset(d->hist, 0, sizeof(d->hist));
  }
  TotalMatchedKeypoints = 0;
  if (resetLog) { NbrFrameProcessed = 0; LogTrajectory.clear(); }
  std::memset(&Stats, 0, sizeof(Stats));
}

void lsb_slam::SetWorldTransformFromGuess(const Iso3d& T)
{
  Tworld = T;
  PreviousTworld = T;
  for (int k = 0; k < 3; ++k) CurrentKeypoints[k].n = 0;
}

Iso3d lsb_slam::GetBaseToLidarOffset(int dev) const
{
  auto it = BaseToLidarOffsets.find(dev);
  return it != BaseToLidarOffsets.end() ? it->second : Iso3d::identity();
}

ExtractorGpu& lsb_slam::Extractor(int dev)
{
  auto& p = KeyPointsExtractors[dev];
  if (!p) p.reset(new ExtractorGpu());
  return *p;
}

//-----------------------------------------------------------------------------
lsb_slam::StageCfg lsb_slam::MakeStageCfg() const
{
  StageCfg c;
  for (const auto& kv : KeyPointsExtractors) if (kv.second) c.ke_ids.push_back(kv.first);
  c.offsets = BaseToLidarOffsets;
  for (int k = 0; k < 3; ++k) c.use[k] = P.use_keypoints[k] != 0;
  return c;
}

// host -> device copy of the input frames (skipped for device-resident input) and the device id of each
void lsb_slam::UploadFrames(FrameList& pool, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq,
                            bool on_device, const StageCfg& cfg, cudaStream_t s)
{
  while (pool.size() < (size_t)n_frames) pool.emplace_back(new DFrame());
  for (int i = 0; i < n_frames; ++i) {
    DFrame& f = *pool[i];
    f.n = n[i]; f.stamp_us = stamp_us[i]; f.seq = seq ? seq[i] : 0u;
    if (on_device) f.d = frames[i];
    else {
      f.own.buf.reserve(std::max<size_t>(n[i], 1), s);
      if (n[i]) LSB_CUDA(cudaMemcpyAsync(f.own.buf.p, frames[i], n[i] * sizeof(lsb_point), cudaMemcpyHostToDevice, s));
      f.d = f.own.buf.p;
    }
    f.device_id = 0;
    if (n[i]) {
      if (on_device && cfg.ke_ids.size() == 1 && cfg.offsets.empty()) {
        // one sensor, no calibration: the device id selects nothing (Slam.cxx:764-780), no read-back
        f.device_id = cfg.ke_ids[0];
      } else if (on_device) {
        lsb_point first;
        LSB_CUDA(cudaMemcpyAsync(&first, frames[i], sizeof(lsb_point), cudaMemcpyDeviceToHost, s));
        LSB_CUDA(cudaStreamSynchronize(s));
        f.device_id = first.device_id;
      } else f.device_id = frames[i][0].device_id;
    }
  }
}

// lsb_prefetch_frames: the helper thread uploads the frames on the ingest stream and extracts their keypoints into a staging
// slot; lsb_add_frames of the same buffers / sizes / stamps picks the slot up instead of extracting.
int lsb_slam::PrefetchFrames(int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq, bool on_device)
{
  LSB_CUDA(cudaSetDevice(device));
  if (n_frames <= 0) return LSB_E_INVALID;
  StagedInput* slot = nullptr;
  for (auto& sl : staged) if (!sl.valid) { slot = &sl; break; }
  if (!slot) throw CudaError(LSB_E_CAPACITY, "lsb_prefetch_frames: three inputs are staged already; add one of them first");
  slot->valid = true;
  slot->on_device = on_device;
  slot->src.assign(frames, frames + n_frames);
  slot->n.assign(n, n + n_frames);
  slot->stamp.assign(stamp_us, stamp_us + n_frames);
  slot->seq.assign((size_t)n_frames, 0u);
  if (seq) slot->seq.assign(seq, seq + n_frames);
  const StageCfg cfg = MakeStageCfg();
  const int mode = HostMode();
  int ahead = 0;                                   // inputs staged and not yet added, this one included
  for (auto& sl : staged) ahead += sl.valid ? 1 : 0;
  const unsigned my_seq = ++stage_seq;
  slot->stage_seq = my_seq;
  // the call before this input's own: the (ahead - 1)-th AddFrames from now
  const unsigned gate = frames_started.load(std::memory_order_relaxed) + (unsigned)(ahead - 1);
  slot->ticket = Worker().post([this, slot, cfg, gate, my_seq] {
    const int nf = (int)slot->src.size();
    UploadFrames(slot->frames, nf, slot->src.data(), slot->n.data(), slot->stamp.data(), slot->seq.data(), slot->on_device, cfg, istream);
    unsigned spins = 0;
    bool nap = false;
    const auto t0 = std::chrono::steady_clock::now();
    while ((int)(frames_started.load(std::memory_order_acquire) - gate) < 0 &&
           (int)(stage_release.load(std::memory_order_acquire) - my_seq) < 0) {
      if ((++spins & 63u) == 0) std::this_thread::yield();
      // frames arrive back to back: spin; the caller took a break (more than a few frame periods): stop burning a core
      if ((spins & 1023u) == 0 && !nap) nap = std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(10);
      if (nap) std::this_thread::sleep_for(std::chrono::microseconds(100));
    }
    ExtractInto(slot->frames, (size_t)nf, slot->stamp[0], cfg, slot, slot->kp, slot->tmin, slot->tmax, istream);
    LSB_CUDA(cudaEventRecord(slot->ready, istream));
  }, mode == 0);
  return LSB_OK;
}

void lsb_slam::JoinMapUpdate()
{
  if (!map_pending) return;
  map_pending = false;
  std::exception_ptr first;
  for (int k = 0; k < 3; ++k) {
    if (!map_posted[k]) continue;
    map_posted[k] = false;
    try { TypeWorker(k).wait(map_ticket[k]); } catch (...) { if (!first) first = std::current_exception(); }
    LSB_CUDA(cudaStreamWaitEvent(stream, ev_map_end[k], 0));
  }
  if (first) std::rethrow_exception(first);     // what the update threw
}

void lsb_slam::FinishMapStats()
{
  JoinMapUpdate();
  if (!map_stats_pending) return;
  map_stats_pending = false;
  float worst = 0.f;
  for (int k = 0; k < 3; ++k) {
    Stats.n_map[k] = LocalMaps[k].Size();
    if (!map_stats_type[k]) continue;
    LSB_CUDA(cudaEventSynchronize(ev_map_end[k]));
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ev[4], ev_map_end[k]) == cudaSuccess) worst = std::max(worst, ms); else cudaGetLastError();
  }
  Stats.ms_map_update = worst;
}

void lsb_slam::Quiesce(bool drop_staged)
{
  LSB_CUDA(cudaSetDevice(device));
  std::exception_ptr first;
  try { FinishMapStats(); } catch (...) { first = std::current_exception(); }
  stage_release.store(stage_seq, std::memory_order_release);      // nobody waits behind a gate
  for (auto& sl : staged) {
    if (!sl.valid) continue;
    try { Worker().wait(sl.ticket); } catch (...) { if (!first) first = std::current_exception(); sl.valid = false; }
    if (drop_staged) sl.valid = false;
  }
  if (first) std::rethrow_exception(first);
}

void lsb_slam::SynchronizeAll()
{
  Quiesce(false);
  for (auto& t : tstream) LSB_CUDA(cudaStreamSynchronize(t));
  LSB_CUDA(cudaStreamSynchronize(istream));
  LSB_CUDA(cudaStreamSynchronize(stream));
}

//-----------------------------------------------------------------------------
int lsb_slam::AddFrames(int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq, bool on_device)
{
  LSB_CUDA(cudaSetDevice(device));
  InFlightGuard inflight;
  Warnings = 0;
  if (!CheckFrames(n_frames, n, stamp_us, seq)) return LSB_W_FRAME_IGNORED;
  // staged ahead of time (lsb_prefetch_frames)?  Same buffers, sizes and stamps; anything else staged is dropped
  StagedInput* hit = nullptr;
  for (auto& sl : staged) {
    if (!sl.valid || hit) continue;
    bool same = sl.on_device == on_device && (int)sl.src.size() == n_frames;
    for (int i = 0; same && i < n_frames; ++i) same = sl.src[i] == frames[i] && sl.n[i] == n[i] && sl.stamp[i] == stamp_us[i];
    if (same) hit = &sl;
  }
  if (!hit) {
    bool any = false;
    for (auto& sl : staged) any = any || sl.valid;
    if (any) Quiesce(true);
  }
  frames_started.fetch_add(1, std::memory_order_release);
  if (map_stats_pending && !map_pending) map_stats_pending = false;   // nobody asked for the previous frame's statistics
  const unsigned launches0 = lc.n;
  LSB_CUDA(cudaEventRecord(ev[0], stream));
  trace.mark("start", stream);
  NbCurrentFrames = (size_t)n_frames;
  {
    const bool keep = map_stats_pending;      // the previous frame's map update may still be running: its statistics are gone with this memset
    std::memset(&Stats, 0, sizeof(Stats));
    (void)keep;
  }
  for (int i = 0; i < n_frames; ++i) Stats.n_points += (uint32_t)n[i];
  if (hit) {
    stage_release.store(hit->stage_seq, std::memory_order_release);   // (added out of turn: its own gate may still be closed)
    try { Worker().wait(hit->ticket); }       // upload + extraction enqueued and their counts read back (rethrows)
    catch (...) { hit->valid = false; throw; }
    FramePool.swap(hit->frames);
    LSB_CUDA(cudaStreamWaitEvent(stream, hit->ready, 0));
  } else {
    UploadFrames(FramePool, n_frames, frames, n, stamp_us, seq, on_device, MakeStageCfg(), stream);
  }
  CurrentStamp = stamp_us[0];
  CurrentTime = StampToSec(CurrentStamp);
  for (int k = 0; k < 3; ++k) WorldKeypointsValid[k] = false;

  trace.mark("h2d", stream);
  ExtractKeypoints(hit);
  if (hit) hit->valid = false;
  trace.mark("extract", stream);
  LSB_CUDA(cudaEventRecord(ev[1], stream));
  ComputeEgoMotion();
  LSB_CUDA(cudaEventRecord(ev[2], stream));
  if (WheelOdomManager.CanBeUsed() || ImuManager.CanBeUsed()) ComputeSensorConstraints();   // :256-261
  Localization();
  LSB_CUDA(cudaEventRecord(ev[3], stream));
  // confidence estimators (:272-279)
  if (P.overlap_sampling_ratio > 0.f) EstimateOverlap();
  else OverlapEstimation = -1.f;                                                // SetOverlapSamplingRatio (:1365-1366)
  if (P.time_window_duration > 0.f) {
    const int w = check_motion_limits(LogTrajectory, Tworld, CurrentTime, P.time_window_duration, P.velocity_limits, P.acceleration_limits,
                                      NbrFrameProcessed, MotionLimits);
    if (w & 1) Warnings |= LSB_W_TIME_WINDOW;
    if (w & 2) Warnings |= LSB_W_MOTION_LIMITS;
  }
  uint32_t n_submap[3];
  for (int k = 0; k < 3; ++k) n_submap[k] = (uint32_t)LocalMaps[k].n_submap;   // sizes of the sub-maps this frame was matched against
  JoinMapUpdate();                                                             // (frames that skipped the localization)
  // the frame's end event: with a deferred map update it stands where the pose is known and the update starts
  LSB_CUDA(cudaEventRecord(ev[4], stream));
  if (P.map_update == LSB_MAP_ADD_KPTS_TO_FIXED_MAP || P.map_update == LSB_MAP_UPDATE) UpdateMapsUsingTworld();
  if (!map_pending) {
    Join();
    LSB_CUDA(cudaEventRecord(ev[4], stream));
  }
  trace.mark("map_update", stream);
  LSB_CUDA(cudaEventSynchronize(ev[4]));
  if (lcp_pending) { OverlapEstimation = h_lcp.p[0]; lcp_pending = false; }
  trace.dump();
  LogCurrentFrameState(CurrentTime);
  for (int k = 0; k < 3; ++k) {
    Stats.n_keypoints[k] = (uint32_t)CurrentKeypoints[k].n;
    Stats.n_submap[k] = n_submap[k];
    Stats.n_map[k] = LocalMaps[k].Size();          // deferred update: completed by FinishMapStats
    Stats.n_matches[k] = (uint32_t)LocDebug.hist[k][LSB_MATCH_SUCCESS];
    Stats.ego_n_matches[k] = (uint32_t)EgoDebug.hist[k][LSB_MATCH_SUCCESS];
  }
  cudaEventElapsedTime(&Stats.ms_extract, ev[0], ev[1]);
  cudaEventElapsedTime(&Stats.ms_ego, ev[1], ev[2]);
  cudaEventElapsedTime(&Stats.ms_localization, ev[2], ev[3]);
  if (map_pending) { Stats.ms_map_update = 0.f; map_stats_pending = true; }
  else cudaEventElapsedTime(&Stats.ms_map_update, ev[3], ev[4]);
  cudaEventElapsedTime(&Stats.ms_total, ev[0], ev[4]);
  Stats.map_update_deferred = map_pending ? 1u : 0u;
  Stats.prefetched = hit ? 1u : 0u;
  Stats.kernel_launches = lc.n - launches0;
  if (trace.on && mw.t_knn.enabled) {      // LSB_TRACE=1 with lsb_set_profiling(1): the individual brackets, in enqueue order
    std::vector<float> a, b, c;
    Stats.ms_knn = mw.t_knn.collect(&Stats.knn_launches, &a);
    Stats.ms_fit = mw.t_fit.collect(nullptr, &b);
    Stats.ms_lm = mw.t_lm.collect(&Stats.lm_launches, &c);
    std::fprintf(stderr, "[lsb trace] knn us:");
    for (float v : a) std::fprintf(stderr, " %.0f", v);
    std::fprintf(stderr, " | fit us:");
    for (float v : b) std::fprintf(stderr, " %.0f", v);
    std::fprintf(stderr, " | lm us:");
    for (float v : c) std::fprintf(stderr, " %.0f", v);
    std::fprintf(stderr, "  (per ICP iteration: types in the order blob, plane, edge)\n");
  } else {
    Stats.ms_knn = mw.t_knn.collect(&Stats.knn_launches);
    Stats.ms_fit = mw.t_fit.collect(nullptr);
    Stats.ms_lm = mw.t_lm.collect(&Stats.lm_launches);
  }
  // device->host: match histograms + LM states (counted by the workspace) and the small per-stage
  // counters (keypoint counts, time range, bboxes, map counters: < 64 B each, bounded here)
  Stats.d2h_bytes = mw.d2h_bytes + 64 * 16;
  mw.d2h_bytes = 0;
  Stats.h2d_bytes = 0;
  if (!on_device) for (int i = 0; i < n_frames; ++i) Stats.h2d_bytes += n[i] * sizeof(lsb_point);
  Latency = Stats.ms_total * 1e-3;
  NbrFrameProcessed++;
  return Warnings;
}

//-----------------------------------------------------------------------------
bool lsb_slam::CheckFrames(int n_frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  bool allEmpty = true;
  for (int i = 0; i < n_frames; ++i) if (n[i] > 0) allEmpty = false;
  if (n_frames <= 0 || allEmpty) return false;                 // :720-724
  if (stamp_us[0] == CurrentStamp) return false;               // :727-731 (stamp 0 after Reset, as the reference)
  for (int i = 0; i < n_frames; ++i) {                         // :734-740
    const unsigned sq = seq ? seq[i] : 0u;
    const unsigned dropped = sq - PreviousFramesSeq[i] - 1;
    if (PreviousFramesSeq[i] > 0 && dropped > 0) Warnings |= LSB_W_FRAMES_DROPPED;
    PreviousFramesSeq[i] = sq;
  }
  return true;
}

//-----------------------------------------------------------------------------
void lsb_slam::ExtractKeypoints(StagedInput* hit)
{
  for (int k = 0; k < 3; ++k) { PreviousKeypoints[k].swap(CurrentKeypoints[k]); CurrentKeypoints[k].n = 0; }   // :751
  if (hit) {
    // extracted ahead of time on the ingest stream: the staged clouds become the current ones, the extractor instances that
    // ran for them become the handle's (GetDebugArray reads the frame they hold)
    for (int k = 0; k < 3; ++k) CurrentKeypoints[k].swap(hit->kp[k]);
    KpTimeMin = hit->tmin; KpTimeMax = hit->tmax;
    std::lock_guard<std::mutex> g(ke_mu);
    for (int dev : hit->used_ids) {
      auto& mine = KeyPointsExtractors[dev];
      auto& theirs = hit->extractors[dev];
      if (!mine || !theirs) continue;
      theirs->adopt_scalars(*mine);
      mine.swap(theirs);
    }
    return;
  }
  ExtractInto(FramePool, NbCurrentFrames, CurrentStamp, MakeStageCfg(), nullptr, CurrentKeypoints, KpTimeMin, KpTimeMax, stream);
}

void lsb_slam::ExtractInto(FrameList& pool, size_t nb, uint64_t stamp0, const StageCfg& cfg, StagedInput* slot, DCloud out[3], double& tmin, double& tmax,
                           cudaStream_t s)
{
  for (int k = 0; k < 3; ++k) out[k].n = 0;
  tmin = std::numeric_limits<double>::max();
  tmax = std::numeric_limits<double>::lowest();
  if (slot) slot->used_ids.clear();
  // With several input frames the extractor buffers are reused, so keypoints are
  // appended to the aggregate right after each extraction.
  for (size_t fi = 0; fi < nb; ++fi) {
    DFrame& f = *pool[fi];
    if (f.n == 0) continue;
    int dev = f.device_id;
    if (std::find(cfg.ke_ids.begin(), cfg.ke_ids.end(), dev) == cfg.ke_ids.end()) {   // :764-780
      if (cfg.ke_ids.size() == 1) dev = cfg.ke_ids[0];
      else continue;
    }
    ExtractorGpu* kep = nullptr;
    if (slot) {
      // a staging slot runs its own extractor instance (the handle's holds the frame in flight) on the handle's current
      // parameters and azimuth estimate, and hands the estimate on
      auto& se = slot->extractors[dev];
      if (!se) se.reset(new ExtractorGpu());
      std::lock_guard<std::mutex> g(ke_mu);
      se->adopt_scalars(*KeyPointsExtractors[dev]);
      kep = se.get();
      if (std::find(slot->used_ids.begin(), slot->used_ids.end(), dev) == slot->used_ids.end()) slot->used_ids.push_back(dev);
    } else kep = KeyPointsExtractors[dev].get();
    ExtractorGpu& ke = *kep;
    ke.compute(f.d, f.n, s, lc);
    if (slot) { std::lock_guard<std::mutex> g(ke_mu); KeyPointsExtractors[dev]->adopt_scalars(ke); }
    // AggregateFrames(keypoints[k], false)  (:1512-1578)
    for (int k = 0; k < 3; ++k) {
      if (!cfg.use[k] || ke.counts[k] == 0) continue;
      DCloud& agg = out[k];
      const size_t start = agg.n, cnt = ke.counts[k];
      if (nb == 1) {
        // one input frame: the extractor's keypoint buffer BECOMES the aggregate (buffers change hands, nothing is copied;
        // the extractor gets the aggregate's old buffer for its next frame)
        agg.buf.swap(ke.kp[k]);       // (ExtractorGpu::compute sizes kp[k] before it writes)
      } else {
        agg.buf.reserve(start + cnt, s, true, start);
        LSB_CUDA(cudaMemcpyAsync(agg.buf.p + start, ke.kp[k].p, cnt * sizeof(lsb_point), cudaMemcpyDeviceToDevice, s));
      }
      agg.n = start + cnt;
      const double timeOffset = StampToSec(f.stamp_us) - StampToSec(stamp0);
      auto it = cfg.offsets.find(f.device_id);
      const Iso3d tf = it != cfg.offsets.end() ? it->second : Iso3d::identity();
      const bool ident = iso_is_approx(tf, Iso3d::identity());
      if (!ident || timeOffset != 0.0) dev_aggregate(agg.buf.p + start, cnt, timeOffset, tf, !ident, s, lc);
      // time range of the aggregate: t -> t + offset is monotone, so min / max commute with it
      const double lo = (timeOffset != 0.0) ? ke.time_min[k] + timeOffset : ke.time_min[k];
      const double hi = (timeOffset != 0.0) ? ke.time_max[k] + timeOffset : ke.time_max[k];
      tmin = std::min(tmin, lo);
      tmax = std::max(tmax, hi);
    }
  }
}

//-----------------------------------------------------------------------------
static lsb_match_params EgoParams(const lsb_params& P)
{
  lsb_match_params m;
  m.single_edge_per_ring = 1;                                    // Slam.cxx:879
  m.max_neighbors_distance = P.ego_motion_max_neighbors_distance;
  m.edge_nb_neighbors = P.ego_motion_edge_nb_neighbors;
  m.edge_min_nb_neighbors = P.ego_motion_edge_min_nb_neighbors;
  m.edge_max_model_error = P.ego_motion_edge_max_model_error;
  m.plane_nb_neighbors = P.ego_motion_plane_nb_neighbors;
  m.planarity_threshold = P.ego_motion_planarity_threshold;
  m.plane_max_model_error = P.ego_motion_plane_max_model_error;
  m.blob_nb_neighbors = 10;                                      // KeypointsMatcher.h:70 (unused by ego-motion)
  m.saturation_distance = 1.;
  return m;
}
static lsb_match_params LocParams(const lsb_params& P)
{
  lsb_match_params m;
  m.single_edge_per_ring = 0;
  m.max_neighbors_distance = P.localization_max_neighbors_distance;
  m.edge_nb_neighbors = P.localization_edge_nb_neighbors;
  m.edge_min_nb_neighbors = P.localization_edge_min_nb_neighbors;
  m.edge_max_model_error = P.localization_edge_max_model_error;
  m.plane_nb_neighbors = P.localization_plane_nb_neighbors;
  m.planarity_threshold = P.localization_planarity_threshold;
  m.plane_max_model_error = P.localization_plane_max_model_error;
  m.blob_nb_neighbors = P.localization_blob_nb_neighbors;
  m.saturation_distance = 1.;
  return m;
}

// keeps the per-keypoint status / weight of the last matching of a stage
void lsb_slam::SaveDebug(StageDebug& d)
{
  for (int k = 0; k < 4; ++k) d.off[k] = mw.off[k];
  if (&d == &LocDebug) return;   // localization results stay in the workspace until the next frame
  d.status.reserve(std::max<size_t>(mw.total, 1), stream);
  d.weight.reserve(std::max<size_t>(mw.total, 1), stream);
  if (mw.total) {
    LSB_CUDA(cudaMemcpyAsync(d.status.p, mw.status.p, mw.total, cudaMemcpyDeviceToDevice, stream));
    LSB_CUDA(cudaMemcpyAsync(d.weight.p, mw.weight.p, mw.total * sizeof(double), cudaMemcpyDeviceToDevice, stream));
  }
}

void lsb_slam::ComputeEgoMotion()
{
  // the frame-to-frame registration uses the type streams: the previous frame's deferred map update is joined first
  if (P.ego_motion == LSB_EGO_REGISTRATION || P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION_AND_REGISTRATION) JoinMapUpdate();
  Trelative = Iso3d::identity();
  std::memset(EgoDebug.hist, 0, sizeof(EgoDebug.hist));
  for (auto& o : EgoDebug.off) o = 0;
  // Linearly extrapolate the previous motion (:821-836)
  if (LogTrajectory.size() >= 2 && (P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION || P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION_AND_REGISTRATION)) {
    const double t = StampToSec(CurrentStamp);
    const double t1 = LogTrajectory[LogTrajectory.size() - 1].time;
    const double t0 = LogTrajectory[LogTrajectory.size() - 2].time;
    if (std::abs((t - t1) / (t1 - t0)) > P.max_extrapolation_ratio) Warnings |= LSB_W_EXTRAPOLATION;
    else {
      const Iso3d next = linear_interpolation(PreviousTworld, Tworld, t, t0, t1);
      Trelative = inverse(Tworld) * next;
    }
  }
  // Frame-to-frame registration (:839-964)
  if (P.ego_motion == LSB_EGO_REGISTRATION || P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION_AND_REGISTRATION) {
    Fork();
    for (int k = 0; k < 2; ++k) {
      if (PreviousKeypoints[k].n) EgoGrids[k].build(PreviousKeypoints[k].buf.p, PreviousKeypoints[k].n, 1.0f, nullptr, nullptr, TypeStream(k), lc);
      else EgoGrids[k].clear();
    }
    TotalMatchedKeypoints = 0;
    if (P.ego_motion_icp_max_iter > 0) {
      const MatchWorkspace::LmHostBlock& blk = RunIcpLoop(ICP_STAGE_EGO, EgoParams(P), P.ego_motion_icp_max_iter, P.ego_motion_lm_max_iter,
                                                          P.ego_motion_init_saturation_distance, P.ego_motion_final_saturation_distance, nullptr);
      const IcpState& st = blk.icp;
      for (int t = 0; t < 3; ++t)
        for (int q = 0; q < LSB_MATCH_N_STATUS; ++q) EgoDebug.hist[t][q] = blk.hist[t * LSB_MATCH_N_STATUS + q];
      Stats.ego_icp_iters = (uint32_t)st.icp_iters;
      Stats.ego_lm_iters = (uint32_t)st.lm_iters;
      Stats.knn_queries += (uint32_t)st.icp_iters * (uint32_t)(CurrentKeypoints[0].n + CurrentKeypoints[1].n);
      TotalMatchedKeypoints = st.total_matched;
      Warnings |= st.warnings;
      Trelative = st.Trelative;
    }
    SaveDebug(EgoDebug);
  }
}

//-----------------------------------------------------------------------------
void lsb_slam::Localization()
{
  PreviousTworld = Tworld;                     // :980-981
  Tworld = PreviousTworld * Trelative;
  std::memset(LocDebug.hist, 0, sizeof(LocDebug.hist));
  for (auto& o : LocDebug.off) o = 0;
  // :986-995; the three clouds are undistorted on their own streams at the head of the submap lanes below
  if (P.undistortion) { InitUndistortion(); RefineUndistortion(/*defer=*/true); }
  trace.mark("undistort", stream);

  // Submaps (:1003-1037): the three maps work on their own streams and host lanes; each lane waits
  // for its own read-backs (bbox of the keypoints, size of the submap)
  ForEachType([&](int k) {                     // (behind the previous frame's map update if that was deferred)
    RollingGridGpu& map = LocalMaps[k];
    cudaStream_t ts = TypeStream(k);
    const bool select = P.use_keypoints[k] && !map.IsSubMapKdTreeValid() && P.map_update != LSB_MAP_NONE && CurrentKeypoints[k].n;
    if (HasPendingRefinement && !select) dev_undistort(CurrentKeypoints[k].buf.p, CurrentKeypoints[k].n, PendingRefinement, 0.0, ts, lc);
    if (!P.use_keypoints[k] || map.IsSubMapKdTreeValid()) return;
    if (P.map_update == LSB_MAP_NONE) { map.BuildSubMapKdTree(ts, lc); return; }
    if (CurrentKeypoints[k].n) {
      // pending undistortion and the box of the cloud under the predicted pose in one launch (before ClearOldPoints, which does
      // not read the keypoints); the box stays on the device for the selection, box and sub-map size come back with one synchronisation
      dev_undistort_transformed_bbox(CurrentKeypoints[k].buf.p, CurrentKeypoints[k].n, HasPendingRefinement ? &PendingRefinement : nullptr, Tworld,
                                     map.box_partial, map.box_counter, map.d_bbox, map.h_bbox, map.NextBoxSeq(), ts, lc);
    }
    if (map.IsTimeThreshold()) map.ClearOldPoints(CurrentTime, ts, lc);
    if (CurrentKeypoints[k].n) {
      map.SubMapSelectDevice((int)(CurrentKeypoints[k].n / 2), ts, lc);
    } else {
      const float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
      map.SubMapSelect(mn, mx, 0, ts, lc);
    }
    map.SubMapFinish(ts, lc);
  }, [&] { if (P.localization_icp_max_iter > 0) IcpBegin(ICP_STAGE_LOC, P.localization_icp_max_iter); });
  HasPendingRefinement = false;
  trace.mark("submap+grid", stream);

  // ICP - LM loop (:1051-1156), device-resident: every iteration enqueued, one hand-over when the loop ends
  TotalMatchedKeypoints = 0;
  if (P.localization_icp_max_iter == 0) return;
  // wheel-odometer / gravity blocks, if the managers built one for this frame (:1123-1131)
  const LmSensorTerms sensorTerms = MakeSensorTerms(WheelOdomManager, ImuManager);
  const MatchWorkspace::LmHostBlock& blk = RunIcpLoop(ICP_STAGE_LOC, LocParams(P), P.localization_icp_max_iter, P.localization_lm_max_iter,
                                                      P.localization_init_saturation_distance, P.localization_final_saturation_distance, &sensorTerms);
  trace.mark("icp", stream);
  const IcpState& st = blk.icp;
  for (int t = 0; t < 3; ++t)
    for (int q = 0; q < LSB_MATCH_N_STATUS; ++q) LocDebug.hist[t][q] = blk.hist[t * LSB_MATCH_N_STATUS + q];
  SaveDebug(LocDebug);
  Stats.loc_icp_iters = (uint32_t)st.icp_iters;
  Stats.loc_lm_iters = (uint32_t)st.lm_iters;
  Stats.knn_queries += (uint32_t)st.icp_iters * (uint32_t)(CurrentKeypoints[0].n + CurrentKeypoints[1].n + CurrentKeypoints[2].n);
  TotalMatchedKeypoints = st.total_matched;
  Warnings |= st.warnings;
  Tworld = st.Tworld;
  Trelative = st.Trelative;
  WithinFrameMotion = st.Within;
  for (int k = 0; k < 3; ++k) WorldKeypointsValid[k] = WorldBoxReady[k] = finish_enqueued && CurrentKeypoints[k].n > 0;
  finish_enqueued = false;
  if (st.lm_valid) {                                                            // :1151-1155, evaluated on demand (Uncertainty())
    std::memcpy(UncertaintyH, blk.state.H, sizeof(UncertaintyH));
    UncertaintyTwoD = P.two_d_mode != 0;
    UncertaintyPending = true;
  }
}

//-----------------------------------------------------------------------------
// Prologue of an ICP - LM loop: the loop's state in device memory and the per-frame reset of the match workspace.  It needs
// nothing of the sub-maps, so the localization enqueues it on `stream` while the helper threads are still selecting them.
void lsb_slam::IcpBegin(int stage, unsigned max_iter)
{
  const bool loc = stage == ICP_STAGE_LOC;
  size_t nk[3] = {CurrentKeypoints[0].n, CurrentKeypoints[1].n, loc ? CurrentKeypoints[2].n : 0};
  IcpState init;
  std::memset(&init, 0, sizeof(init));
  init.Tworld = Tworld; init.PreviousTworld = PreviousTworld; init.Trelative = Trelative;
  init.Within = WithinFrameMotion;
  init.Pending = WithinFrameMotion;
  init.log_empty = LogTrajectory.empty() ? 1 : 0;
  init.prev_pose_time = LogTrajectory.empty() ? 0. : LogTrajectory.back().time;
  init.curr_pose_time = StampToSec(CurrentStamp);
  init.max_extrapolation_ratio = P.max_extrapolation_ratio;
  init.undistortion = loc ? P.undistortion : 0;
  d_icp.reserve(1, stream);
  dev_icp_init(d_icp.p, init, stream, lc);
  mw.icp_stage = stage;
  // several sequences sharing the GPU: each solve takes its share of the SMs (never less than a sixth)
  static const int share_max = [] { const char* e = std::getenv("LSB_LM_GRID_SHARE_MAX"); return e ? std::max(1, std::atoi(e)) : 6; }();   // tuning hook (6: best of 1 / 2 / 4 / 6 / 8 for 8 OS-128 sequences on one B200)
  mw.lm_grid_share = std::min(std::max(g_handles_in_flight.load(std::memory_order_relaxed), 1), share_max);
  mw.begin_frame(nk, (int)max_iter, stream);
  mw.next_solve_seq();
  icp_begun = true;
}

//-----------------------------------------------------------------------------
// One ICP - LM loop of the frame without a host round trip between its iterations (cuda/icp.cuh): the pose, the within-frame
// motion and the loop's break tests live in device memory; iteration i + 1 is enqueued behind iteration i and stands aside
// if the loop ended before it.  Per iteration and keypoint type (own streams): undistortion refinement of the previous solve,
// kNN (from the second iteration on bounded by the previous neighbour lists), model fit; then one persistent LM kernel whose
// tail (icp_control) moves the pose on.  The kernel that ends the loop hands the result block to the host.
const MatchWorkspace::LmHostBlock& lsb_slam::RunIcpLoop(int stage, const lsb_match_params& base, unsigned max_iter, unsigned lm_max_iter,
                                                        double sat_init, double sat_final, const LmSensorTerms* terms)
{
  const bool loc = stage == ICP_STAGE_LOC;
  const int ntypes = loc ? 3 : 2;                                   // ego-motion never uses blobs (:848, :910)
  size_t nk[3] = {CurrentKeypoints[0].n, CurrentKeypoints[1].n, loc ? CurrentKeypoints[2].n : 0};
  if (!icp_begun) IcpBegin(stage, max_iter);
  icp_begun = false;
  const bool refined = loc && P.undistortion == LSB_UNDISTORTION_REFINED;
  lsb_match_params mp = base;
  for (unsigned icpIter = 0; icpIter < max_iter; ++icpIter) {
    const double iterRatio = icpIter / static_cast<double>(max_iter - 1);
    mp.saturation_distance = (1 - iterRatio) * sat_init + iterRatio * sat_final;          // :905-906, :1084-1085
    Fork();
    for (int k : kMatchOrder) {
      if (k >= ntypes) continue;
      // the undistortion refinement of the previous iteration (:1146-1147): three independent clouds, one per type stream
      // (fused into the search kernel when that one runs per keypoint: reuse pass on a non-empty target)
      const KnnGrid& target = loc ? LocalMaps[k].grid : EgoGrids[k];
      const bool fuse = refined && icpIter > 0 && mw.can_fuse_refinement(target, nk[k]);
      if (refined && icpIter > 0 && !fuse) dev_undistort_state(CurrentKeypoints[k].buf.p, CurrentKeypoints[k].n, d_icp.p, 0, TypeStream(k), lc);
      mw.match_type(k, CurrentKeypoints[k].buf.p, nk[k], target, Iso3d::identity(), mp, TypeStream(k), lc, (int)icpIter, d_icp.p,
                    /*reuse=*/icpIter > 0, /*refine=*/fuse);
    }
    Join();
    IcpCtl ctl;
    ctl.stage = stage; ctl.icp_iter = (int)icpIter; ctl.icp_max_iter = (int)max_iter; ctl.min_matches = P.min_nb_matched_keypoints;
    mw.enqueue_solve(nullptr, (int)lm_max_iter, P.two_d_mode != 0, mp.saturation_distance, P.min_nb_matched_keypoints, stream, lc, terms,
                     d_icp.p, &ctl);
  }
  // the refinement of the iteration that ended the loop is applied at once (RefineUndistortion of the last iteration, :1146-1147);
  // when a map update may follow, the same launch leaves the world clouds and their boxes (input of RollingGrid::Add) behind
  finish_enqueued = false;
  if (loc && (P.map_update == LSB_MAP_ADD_KPTS_TO_FIXED_MAP || P.map_update == LSB_MAP_UPDATE)) {
    FinishArgs fa;
    for (int k = 0; k < 3; ++k) {
      DCloud& w = CurrentWorldKeypoints[k];
      w.buf.reserve(std::max<size_t>(nk[k], 1), stream);
      w.n = nk[k];
      LocalMaps[k].d_bbox.reserve(8, stream); LocalMaps[k].h_bbox.reserve(8);
      fa.cur[k] = CurrentKeypoints[k].buf.p; fa.world[k] = w.buf.p; fa.n[k] = (int)nk[k];
      fa.d_bb[k] = reinterpret_cast<int*>(LocalMaps[k].d_bbox.p); fa.h_bb[k] = reinterpret_cast<int*>(LocalMaps[k].h_bbox.p);
      fa.seq[k] = LocalMaps[k].NextBoxSeq();
      LocalMaps[k].box_by_kernel = nk[k] > 0;      // RollingGrid::Add polls for this box instead of synchronising its stream
    }
    dev_finish_keypoints(fa, d_icp.p, refined, finish_partial, finish_counter, stream, lc);
    finish_enqueued = true;
  } else if (refined) {
    for (int k = 0; k < 3; ++k) dev_undistort_state(CurrentKeypoints[k].buf.p, CurrentKeypoints[k].n, d_icp.p, 1, stream, lc);
  }
  return mw.wait_result(stream);
}

//-----------------------------------------------------------------------------
void lsb_slam::ComputeSensorConstraints()
{
  bool warn = false;
  WheelOdomManager.ComputeWheelAbsoluteConstraint(CurrentTime, &warn);
  ImuManager.ComputeGravityConstraint(CurrentTime, &warn);
  if (warn) Warnings |= LSB_W_SENSOR_TIME;
}

//-----------------------------------------------------------------------------
void lsb_slam::UpdateMapsUsingTworld()
{
  const Iso3d motion = inverse(KfLastPose) * Tworld;
  const double trans = norm(motion.t);
  const double rot = rotation_angle(motion.R);
  const double thresholdCoef = std::min(KfCounter / 10., 1.);
  unsigned nbMapKpts = 0;
  for (const auto& m : LocalMaps) nbMapKpts += m.Size();
  const bool isNewKeyFrame = nbMapKpts < P.min_nb_matched_keypoints * 10 || trans >= thresholdCoef * P.kf_distance_threshold ||
                             rot >= (thresholdCoef * P.kf_angle_threshold) / 180. * M_PI;
  if (!isNewKeyFrame) return;
  Stats.is_keyframe = 1;
  KfCounter++;
  KfLastPose = Tworld;
  const int mode = HostMode();
  if (defer_map_update && use_lanes && !serial_types && mode != 2) {
    // Deferred: the transform + bounding box of the three clouds is enqueued here (it reads this frame's keypoint buffers,
    // which the next frame's extraction recycles), the two read-back rounds of RollingGrid::Add are driven by the types' helper
    // threads on the type streams while the caller returns with the pose.  The next frame's sub-map selection is posted to the
    // same threads and streams, so it is ordered behind the update on the host and on the device.
    Fork();
    struct Job { const lsb_point* p; size_t n; bool use; } job[3];
    for (int k : kTypeOrder) {
      job[k].use = P.use_keypoints[k] != 0;
      job[k].p = nullptr; job[k].n = 0;
      if (!job[k].use) continue;
      bool boxed = false;
      const DCloud& w = WorldKeypoints(k, tstream[k], &LocalMaps[k], &boxed);
      if (!boxed) LocalMaps[k].AddBegin(w.buf.p, w.n, true, tstream[k], lc);
      job[k].p = w.buf.p; job[k].n = w.n;
    }
    const double now = CurrentTime;
    for (int k : kTypeOrder) {
      map_posted[k] = map_stats_type[k] = job[k].use;
      if (!job[k].use) continue;
      const Job jb = job[k];
      map_ticket[k] = TypeWorker(k).post([this, jb, now, k] {
        LocalMaps[k].AddSort(jb.p, jb.n, false, now, true, tstream[k], lc);
        LocalMaps[k].AddMerge(jb.n, tstream[k], lc);
        LSB_CUDA(cudaEventRecord(ev_map_end[k], tstream[k]));
      }, mode == 0);
    }
    map_pending = true;
    return;
  }
  // RollingGrid::Add of the three types, each on its stream and host lane
  ForEachType([&](int k) {
    if (!P.use_keypoints[k]) return;
    cudaStream_t ts = TypeStream(k);
    bool boxed = false;
    const DCloud& w = WorldKeypoints(k, ts, &LocalMaps[k], &boxed);      // the box RollingGrid::Add rolls by comes with the transform
    if (!boxed) LocalMaps[k].AddBegin(w.buf.p, w.n, true, ts, lc);
    LocalMaps[k].AddSort(w.buf.p, w.n, false, CurrentTime, true, ts, lc);
    LocalMaps[k].AddMerge(w.n, ts, lc);
  });
}

// GetKeypoints(k, true) -> TransformPointCloud (:690-702, :1491-1509), cached per frame
const DCloud& lsb_slam::WorldKeypoints(int k, cudaStream_t s, RollingGridGpu* box, bool* boxed)
{
  cudaStream_t stream = s ? s : this->stream;
  DCloud& w = CurrentWorldKeypoints[k];
  if (boxed) *boxed = false;
  if (WorldKeypointsValid[k] && WorldBoxReady[k] && box && boxed) { *boxed = true; WorldBoxReady[k] = false; }   // left by the loop's tail
  if (!WorldKeypointsValid[k]) {
    const DCloud& c = CurrentKeypoints[k];
    w.buf.reserve(std::max<size_t>(c.n, 1), stream);
    w.n = c.n;
    if (c.n) {
      LSB_CUDA(cudaMemcpyAsync(w.buf.p, c.buf.p, c.n * sizeof(lsb_point), cudaMemcpyDeviceToDevice, stream));
      if (box) { box->box_by_kernel = false; dev_transform_bbox_async(w.buf.p, w.n, Tworld, box->d_bbox, box->h_bbox, stream, lc); if (boxed) *boxed = true; }
      else dev_transform(w.buf.p, w.n, Tworld, stream, lc);
    }
    WorldKeypointsValid[k] = true;
  }
  return w;
}

//-----------------------------------------------------------------------------
void lsb_slam::LogCurrentFrameState(double time)
{
  LogTrajectory.push_back(LoggedPose{Tworld, time});
  if (P.logging_timeout != 0.) {
    if (P.logging_timeout > 0)
      while (time - LogTrajectory.front().time > P.logging_timeout && LogTrajectory.size() > 2) LogTrajectory.pop_front();
  } else {
    while (LogTrajectory.size() > 2) LogTrajectory.pop_front();
  }
}

//-----------------------------------------------------------------------------
Iso3d lsb_slam::InterpolateScanPose(double time)
{
  if (LogTrajectory.empty()) return Tworld;
  const double prevPoseTime = LogTrajectory.back().time;
  const double currPoseTime = StampToSec(CurrentStamp);
  if (std::abs(time / (currPoseTime - prevPoseTime)) > P.max_extrapolation_ratio) { Warnings |= LSB_W_EXTRAPOLATION; return Tworld; }
  return linear_interpolation(PreviousTworld, Tworld, currPoseTime + time, prevPoseTime, currPoseTime);
}

void lsb_slam::InitUndistortion()
{
  // time range over ALL keypoint clouds (:1293-1300): reduced by the extraction kernels and read
  // back with the keypoint counts (ExtractKeypoints), no extra pass / synchronisation here
  const double first = KpTimeMin, last = KpTimeMax;
  WithinFrameMotion.set_times(first, last);
  WithinFrameMotion.set_transforms(Iso3d::identity(), Iso3d::identity());
  if (WithinFrameMotion.t1 - WithinFrameMotion.t0 < 1e-6) { Warnings |= LSB_W_TIME_FIELD; WithinFrameMotion.set_times(0., 0.); }
  else if (WithinFrameMotion.t1 - WithinFrameMotion.t0 > 10.) Warnings |= LSB_W_TIME_FIELD;
}

void lsb_slam::RefineUndistortion(bool defer)
{
  const Iso3d prevBegin = WithinFrameMotion.H0(), prevEnd = WithinFrameMotion.H1();
  const Iso3d wBegin = InterpolateScanPose(WithinFrameMotion.t0);
  const Iso3d wEnd = InterpolateScanPose(WithinFrameMotion.t1);
  const Iso3d baseToWorld = inverse(Tworld);
  const Iso3d newBegin = baseToWorld * wBegin, newEnd = baseToWorld * wEnd;
  WithinFrameMotion.set_transforms(newBegin, newEnd);
  Interpolator interp = WithinFrameMotion;
  interp.set_transforms(newBegin * inverse(prevBegin), newEnd * inverse(prevEnd));
  for (int k = 0; k < 3; ++k) WorldKeypointsValid[k] = false;
  if (defer) {          // applied per type, on the type's own stream, at the head of the next matching pass
    PendingRefinement = interp;
    HasPendingRefinement = true;
    return;
  }
  for (int k = 0; k < 3; ++k) dev_undistort(CurrentKeypoints[k].buf.p, CurrentKeypoints[k].n, interp, 0.0, stream, lc);
}

//-----------------------------------------------------------------------------
// Slam::EstimateOverlap (:1370-1388): LCP estimator of ConfidenceEstimators.cxx:27-65 on the device.
// A sampled point's nearest sub-map point is searched in every map in use (k = 1, capped at the
// distance beyond which the probability is exactly 0.f), on the maps' own streams.
void lsb_slam::EstimateOverlap()
{
  const size_t total = RegisteredFrame(tmp_cloud);
  const float ratio = std::min(std::max(P.overlap_sampling_ratio, 0.f), 1.f);
  const int nb = (int)((float)total * ratio);
  bool use[3];
  int nuse = 0;
  for (int k = 0; k < 3; ++k) { use[k] = P.use_keypoints[k] && LocalMaps[k].IsSubMapKdTreeValid(); nuse += use[k] ? 1 : 0; }
  if (nb == 0 || nuse == 0) { OverlapEstimation = -1.f; return; }
  lcp_q.reserve(3 * (size_t)nb, stream);
  d_lcp.reserve(1, stream); h_lcp.reserve(1);
  dev_lcp_gather(tmp_cloud.p, nb, ratio, lcp_q.p, stream, lc);
  for (int k = 0; k < 3; ++k) if (use[k]) { lcp_d2[k].reserve((size_t)nb, stream); lcp_idx[k].reserve((size_t)nb, stream); }
  Fork();
  const float* d2[3] = {nullptr, nullptr, nullptr};
  float thr[3] = {1.f, 1.f, 1.f};
  for (int k : kTypeOrder) {
    if (!use[k]) continue;
    thr[k] = lcp_sq_threshold(LocalMaps[k].LeafSize);
    const float cap = std::sqrt(lcp_sq_cutoff(thr[k])) * 1.0001f + 1e-3f;
    LocalMaps[k].grid.search(lcp_q.p, (size_t)nb, 1, cap, lcp_idx[k].p, lcp_d2[k].p, nullptr, /*entry_major=*/true, TypeStream(k), lc);
    d2[k] = lcp_d2[k].p;
    Stats.knn_queries += (uint32_t)nb;
  }
  Join();
  dev_lcp_reduce(d2, thr, nb, lcp_chunks, d_lcp.p, stream, lc);
  LSB_CUDA(cudaMemcpyAsync(h_lcp.p, d_lcp.p, sizeof(float), cudaMemcpyDeviceToHost, stream));
  lcp_pending = true;
}

//-----------------------------------------------------------------------------
// AggregateFrames(CurrentFrames, true)  (:1512-1578)
size_t lsb_slam::RegisteredFrame(DevBuf<lsb_point>& out)
{
  size_t total = 0;
  for (size_t i = 0; i < NbCurrentFrames; ++i) total += FramePool[i]->n;
  out.reserve(std::max<size_t>(total, 1), stream);
  size_t start = 0;
  for (size_t i = 0; i < NbCurrentFrames; ++i) {
    DFrame& f = *FramePool[i];
    if (f.n == 0) continue;
    LSB_CUDA(cudaMemcpyAsync(out.p + start, f.d, f.n * sizeof(lsb_point), cudaMemcpyDeviceToDevice, stream));
    const double timeOffset = StampToSec(f.stamp_us) - StampToSec(CurrentStamp);
    const Iso3d baseToLidar = GetBaseToLidarOffset(f.device_id);
    if (P.undistortion) {
      Interpolator interp = WithinFrameMotion;
      interp.set_transforms(Tworld * WithinFrameMotion.H0() * baseToLidar, Tworld * WithinFrameMotion.H1() * baseToLidar);
      // point.time += timeOffset, then interpolator(point.time)
      if (timeOffset != 0.0) dev_aggregate(out.p + start, f.n, timeOffset, Iso3d::identity(), false, stream, lc);
      dev_undistort(out.p + start, f.n, interp, 0.0, stream, lc);
    } else {
      const Iso3d tf = Tworld * baseToLidar;
      const bool ident = iso_is_approx(tf, Iso3d::identity());
      if (!ident || timeOffset != 0.0) dev_aggregate(out.p + start, f.n, timeOffset, tf, !ident, stream, lc);
    }
    start += f.n;
  }
  return total;
}

// lsb_slam: the per-sequence state machine behind the C ABI = LidarSlam::Slam's
// per-frame pipeline (reference slam_lib/src/Slam.cxx:230-344 and the steps it
// calls), with every cloud resident in device memory.  Pose bookkeeping (a few
// 4x4 products per frame) stays on the host in FP64 (common/linalg.h).
#pragma once
#include <deque>
#include <map>
#include <memory>
#include <vector>
#include "../cuda/dev_util.h"
#include "../cuda/extract.cuh"
#include "../cuda/knn.cuh"
#include "../cuda/match.cuh"
#include "../cuda/map.cuh"
#include "../cuda/convert.cuh"
#include "../common/linalg.h"
#include "../common/lm_covariance.h"
#include "../common/motion_limits.h"
#include "../common/sensor_constraints.h"
#include "worker.h"
#include <mutex>

struct DCloud {
  lsb::DevBuf<lsb_point> buf;
  size_t n = 0;
  void swap(DCloud& o) { buf.swap(o.buf); size_t t = n; n = o.n; o.n = t; }
};

struct lsb_slam {
  int device = 0;
  cudaStream_t stream = nullptr;
  // one stream per keypoint type: sub-map extraction, kNN + model fit and the map update of the three
  // types are independent until the LM reduction (KeypointsMatcher / RollingGrid instances per type)
  cudaStream_t tstream[3] = {nullptr, nullptr, nullptr};
  cudaEvent_t ev_fork = nullptr, ev_join[3] = {nullptr, nullptr, nullptr};
  bool serial_types = false;       // profiling: every type on `stream`, kernels timed without overlap
  cudaStream_t TypeStream(int k) const { return serial_types ? stream : tstream[k]; }
  void Fork();                     // type streams wait for `stream`
  void Join();                     // `stream` waits for the type streams
  // Fork, fn(k) for the three types, Join.  Each type has its own helper thread (host/worker.h) that enqueues the type's launch
  // sequences and waits for its small read-backs, so that the three sequences do not queue behind one another on one host
  // thread; the same thread runs the type's deferred map update, so a type's jobs stay in order without a join in between.
  // serial_types / LSB_HOST_LANES=0 / many handles sharing the host cores: everything on the caller's thread.
  std::unique_ptr<lsb::AsyncWorker> tworker[3];
  lsb::AsyncWorker& TypeWorker(int k) { if (!tworker[k]) tworker[k].reset(new lsb::AsyncWorker(device)); return *tworker[k]; }
  bool use_lanes = true;
  void ForEachType(const std::function<void(int)>& fn, const std::function<void()>& meanwhile = {});   // meanwhile: run by the caller beside the helpers
  lsb::LaunchCounter lc;
  lsb_params P;

  std::map<int, std::unique_ptr<lsb::ExtractorGpu>> KeyPointsExtractors;   // Slam.h:500
  std::map<int, lsb::Iso3d> BaseToLidarOffsets;                            // Slam.h:505
  lsb::RollingGridGpu LocalMaps[3];                                        // Slam.h:553

  lsb::Iso3d Tworld, PreviousTworld, Trelative;                            // Slam.h:461-467
  lsb::Interpolator WithinFrameMotion;                                     // Slam.h:481
  struct LoggedPose { lsb::Iso3d T; double time; };
  std::deque<LoggedPose> LogTrajectory;                                    // Slam.h:487
  lsb::Iso3d KfLastPose;
  int KfCounter = 0;
  unsigned NbrFrameProcessed = 0;
  double CurrentTime = 0.;
  uint64_t CurrentStamp = 0;
  std::map<int, unsigned> PreviousFramesSeq;
  double Latency = 0.;

  // CurrentRawKeypoints and CurrentUndistortedKeypoints alias each other in the
  // reference (Slam.cxx:984): one device storage.
  DCloud CurrentKeypoints[3], PreviousKeypoints[3], CurrentWorldKeypoints[3];
  bool WorldKeypointsValid[3] = {false, false, false};
  bool WorldBoxReady[3] = {false, false, false};     // LocalMaps[k].h_bbox holds the box of CurrentWorldKeypoints[k] (k_finish_keypoints)
  bool finish_enqueued = false;                      // the loop's tail produced the world clouds of this frame
  lsb::DevBuf<int> finish_partial;
  lsb::DevBuf<unsigned> finish_counter;
  double KpTimeMin = 0., KpTimeMax = 0.;   // `time` range of CurrentKeypoints (read back with the keypoint counts)

  // device copies of the input frames (the reference retains the input clouds, Slam.cxx:237)
  struct DFrame { DCloud own; const lsb_point* d = nullptr; size_t n = 0; uint64_t stamp_us = 0; uint32_t seq = 0; int device_id = 0; };
  typedef std::vector<std::unique_ptr<DFrame>> FrameList;
  FrameList FramePool;
  size_t NbCurrentFrames = 0;

  // ---- work the caller does not wait for (host/worker.h) --------------------------------------------------------------
  // (1) Deferred map update: AddFrames returns once the pose is known; RollingGrid::Add of the three maps (two read-back
  //     rounds each) is driven by the helper thread on the type streams and joined before the next frame's sub-maps are
  //     selected, or by whatever entry point looks at the maps first.
  // (2) Prefetch (lsb_prefetch_frames): upload + keypoint extraction of a frame handed in ahead of time run on the ingest
  //     stream while the current frame is in its ICP loop (which leaves most SMs idle); AddFrames of that frame finds its
  //     keypoints ready.  Three slots: frames t + 1 and t + 2 can both be staged before frame t is consumed (the upload of t + 2 then has a whole frame period).
  std::unique_ptr<lsb::AsyncWorker> worker;
  lsb::AsyncWorker& Worker() { if (!worker) worker.reset(new lsb::AsyncWorker(device)); return *worker; }
  bool defer_map_update = true;          // LSB_DEFER_MAP=0 / lsb_set_pipelining(h, 0): the map update ends inside AddFrames
  bool map_pending = false;
  unsigned map_ticket[3] = {0, 0, 0};
  bool map_posted[3] = {false, false, false};
  cudaEvent_t ev_map_end[3] = {nullptr, nullptr, nullptr};
  bool map_stats_pending = false;        // ms_map_update / n_map of the last frame still to be filled in
  bool map_stats_type[3] = {false, false, false};
  void JoinMapUpdate();                  // host side of the deferred update over, `stream` ordered behind the type streams
  void FinishMapStats();                 // waits for the deferred kernels and completes Stats (lsb_get_frame_stats)
  void Quiesce(bool drop_staged);        // every helper job over (entry points other than AddFrames / prefetch / pose getters)
  void SynchronizeAll();                 // Quiesce + every stream idle

  struct StageCfg {                      // what the extraction reads from the handle, captured when the job is posted
    std::vector<int> ke_ids;
    std::map<int, lsb::Iso3d> offsets;
    bool use[3] = {true, true, true};
  };
  struct StagedInput {
    bool valid = false, on_device = false;
    unsigned ticket = 0, stage_seq = 0;
    std::vector<const lsb_point*> src;
    std::vector<size_t> n;
    std::vector<uint64_t> stamp;
    std::vector<uint32_t> seq;
    FrameList frames;
    DCloud kp[3];
    double tmin = 0., tmax = 0.;
    std::map<int, std::unique_ptr<lsb::ExtractorGpu>> extractors;
    std::vector<int> used_ids;           // extractors that ran for this input
    cudaEvent_t ready = nullptr;         // recorded on the ingest stream behind the last kernel of the job
  } staged[3];
  cudaStream_t istream = nullptr;        // ingest stream (lowest priority: the ICP of the current frame goes first)
  // An input staged TWO calls ahead is uploaded at once but extracted only when the call before its own begins: its extraction
  // then runs beside that frame's map update and sub-map selection (short launches that leave the SMs idle), as it does for
  // device-resident inputs, instead of trailing the upload into the first matching pass.  frames_started counts AddFrames calls;
  // stage_release forces every staged job up to that sequence number through (set by whoever is about to wait for one).
  std::atomic<unsigned> frames_started{0}, stage_release{0};
  unsigned stage_seq = 0;
  std::mutex ke_mu;                      // extractor instances / their scalar state (parameters, azimuth estimate) change hands
  int PrefetchFrames(int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq, bool on_device);
  StageCfg MakeStageCfg() const;
  void UploadFrames(FrameList& pool, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq,
                    bool on_device, const StageCfg& cfg, cudaStream_t s);
  // SpinningSensorKeypointExtractor per frame + AggregateFrames(keypoints, false) into out[3]; exts == nullptr: the handle's own extractors
  void ExtractInto(FrameList& pool, size_t nb, uint64_t stamp0, const StageCfg& cfg, StagedInput* slot, DCloud out[3], double& tmin, double& tmax,
                   cudaStream_t s);
  // whole-loop timing (lsb_timer_begin / lsb_timer_end)
  cudaEvent_t ev_timer[2] = {nullptr, nullptr};

  lsb::MatchWorkspace mw;
  lsb::DevBuf<lsb::IcpState> d_icp;   // state of the device-resident ICP loop (cuda/icp.cuh)
  // enqueues every iteration of an ICP - LM loop, waits for the hand-over of the iteration that ends it
  const lsb::MatchWorkspace::LmHostBlock& RunIcpLoop(int stage, const lsb_match_params& base, unsigned max_iter, unsigned lm_max_iter,
                                                     double sat_init, double sat_final, const lsb::LmSensorTerms* terms);
  void IcpBegin(int stage, unsigned max_iter);     // the loop's prologue (may be enqueued ahead of RunIcpLoop)
  bool icp_begun = false;
  lsb::KnnGrid EgoGrids[2];
  // match results kept for GetDebugArray (Slam.cxx:635-657)
  struct StageDebug { lsb::DevBuf<uint8_t> status; lsb::DevBuf<double> weight; size_t off[4] = {0, 0, 0, 0}; int hist[3][LSB_MATCH_N_STATUS] = {}; };
  StageDebug EgoDebug, LocDebug;
  // LocalOptimizer::EstimateRegistrationError (LocalOptimizer.cxx:112-140) is evaluated when somebody asks for it: the normal
  // matrix of the last solve is kept, the 6x6 eigen-decompositions stay off the frame's critical path
  mutable lsb_registration_error LocalizationUncertainty;
  mutable bool UncertaintyPending = false;
  double UncertaintyH[21];
  bool UncertaintyTwoD = false;
  const lsb_registration_error& Uncertainty() const
  {
    if (UncertaintyPending) { lsb::registration_error_from_normal_matrix(UncertaintyH, UncertaintyTwoD, &LocalizationUncertainty); UncertaintyPending = false; }
    return LocalizationUncertainty;
  }
  unsigned TotalMatchedKeypoints = 0;
  float OverlapEstimation = -1.f;
  lsb::MotionLimitsState MotionLimits;                                     // Slam.h:660-663
  lsb::WheelOdometryManager WheelOdomManager;                              // Slam.h:572
  lsb::ImuManager ImuManager;                                              // Slam.h:578
  lsb_frame_stats Stats;
  int Warnings = 0;

  // scratch
  lsb::ConvertWorkspace conv;                                              // driver-cloud adaptors
  lsb::DevBuf<lsb_point> tmp_cloud;
  lsb::DevBuf<float> tmp_f;
  lsb::DevBuf<uint8_t> tmp_u8;
  cudaEvent_t ev[6] = {};
  // LSB_TRACE=1: named spans of the frame on `stream` (CUDA events), printed to stderr per frame
  struct Trace {
    bool on = false;
    std::vector<cudaEvent_t> evs;
    std::vector<const char*> names;
    std::vector<double> host_us;      // host clock at the mark (enqueue time line)
    size_t used = 0;
    void mark(const char* name, cudaStream_t s);
    void dump();
  } trace;

  explicit lsb_slam(int dev);
  ~lsb_slam();
  void Reset(bool resetLog);
  void ClearMaps();
  int AddFrames(int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq, bool on_device);
  void SetWorldTransformFromGuess(const lsb::Iso3d& T);
  lsb::Iso3d GetBaseToLidarOffset(int dev) const;
  lsb::ExtractorGpu& Extractor(int dev);
  // device clouds produced on demand
  // box != nullptr: the bounding box of the cloud is reduced into box->d_bbox / h_bbox while it is transformed (*boxed = true);
  // a cloud that was already cached comes back without a box (*boxed = false)
  const DCloud& WorldKeypoints(int k, cudaStream_t s = nullptr, lsb::RollingGridGpu* box = nullptr, bool* boxed = nullptr);
  size_t RegisteredFrame(lsb::DevBuf<lsb_point>& out);

 private:
  bool CheckFrames(int n_frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq);   // Slam.cxx:709-743
  void ExtractKeypoints(StagedInput* hit);               // :746-810
  void ComputeEgoMotion();                               // :813-972
  void ComputeSensorConstraints();                       // :346-352
  void Localization();                                   // :975-1175
  void UpdateMapsUsingTworld();                          // :1178-1222
  void LogCurrentFrameState(double time);                // :1225-1264
  lsb::Iso3d InterpolateScanPose(double time);           // :1271-1285
  void InitUndistortion();                               // :1288-1319
  void RefineUndistortion(bool defer = false);           // :1322-1352
  lsb::Interpolator PendingRefinement;                   // refinement enqueued with the next matching pass (per type stream)
  bool HasPendingRefinement = false;
  void SaveDebug(StageDebug& d);
  void EstimateOverlap();                                // :1370-1388 + ConfidenceEstimators.cxx:27-65
  // overlap estimator scratch
  lsb::DevBuf<float> lcp_q, lcp_d2[3], lcp_chunks, d_lcp;
  lsb::DevBuf<int32_t> lcp_idx[3];
  lsb::PinnedBuf<float> h_lcp;
  bool lcp_pending = false;        // a value is on its way to h_lcp (read after the frame's final sync)
};

namespace lsb {
// tail of the localization loop for the three keypoint clouds (slam_kernels.cu: k_finish_keypoints)
struct FinishArgs {
  lsb_point* cur[3];
  lsb_point* world[3];
  int n[3];
  int b0[4];            // first block of each type (filled by dev_finish_keypoints)
  int* d_bb[3];         // box of the world cloud, ordered-int encoding: device copy and pinned host copy (RollingGridGpu::d_bbox / h_bbox)
  int* h_bb[3];
  int seq[3];           // sequence word written behind each box (h_bb[k][6])
};
void dev_finish_keypoints(FinishArgs a, const IcpState* st, bool refine, DevBuf<int>& partial, DevBuf<unsigned>& counter, cudaStream_t s,
                          LaunchCounter& lc);
// head of a type's sub-map selection: pending undistortion (nullptr: none) + box of the cloud under T, box published in d_bbox and
// in h_bbox (pinned, readable after the stream's next synchronisation); one launch
void dev_undistort_transformed_bbox(lsb_point* d_pts, size_t n, const Interpolator* pending, const Iso3d& T, DevBuf<int>& partial,
                                    DevBuf<unsigned>& counter, DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox, int seq, cudaStream_t s,
                                    LaunchCounter& lc);   // seq: written behind the box in h_bbox[6] (RollingGridGpu::box_seq)
// device helpers implemented in slam_kernels.cu
void dev_transform(lsb_point* d_pts, size_t n, const Iso3d& T, cudaStream_t s, LaunchCounter& lc);                     // Utils::TransformPoint
// dev_transform + the bounding box of the result (ordered-int encoding in d_bbox, copy to h_bbox enqueued): no second pass
void dev_transform_bbox_async(lsb_point* d_pts, size_t n, const Iso3d& T, DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox, cudaStream_t s,
                              LaunchCounter& lc);
void dev_aggregate(lsb_point* d_pts, size_t n, double timeOffset, const Iso3d& T, bool applyT, cudaStream_t s, LaunchCounter& lc);
void dev_undistort(lsb_point* d_pts, size_t n, const Interpolator& it, double timeOffset, cudaStream_t s, LaunchCounter& lc);   // Slam.cxx:1345-1350
// the same with the interpolator the device-resident ICP loop holds (cuda/icp.cuh); mode 0: head of an iteration, 1: after the loop
void dev_undistort_state(lsb_point* d_pts, size_t n, const IcpState* st, int mode, cudaStream_t s, LaunchCounter& lc);
void dev_icp_init(IcpState* d_state, const IcpState& init, cudaStream_t s, LaunchCounter& lc);
// measured FP64 / FP32 peaks of the device (cuda/peaks.cu)
void measure_fp_peaks(int device, double out_tflops[4]);
// LCP overlap estimator (cuda/lcp.cu)
void dev_lcp_gather(const lsb_point* d_cloud, int nb, float ratio, float* d_q_xyz, cudaStream_t s, LaunchCounter& lc);
void dev_lcp_reduce(const float* const d2[3], const float thr[3], int nb, DevBuf<float>& chunk_sum, float* d_out, cudaStream_t s, LaunchCounter& lc);
}  // namespace lsb

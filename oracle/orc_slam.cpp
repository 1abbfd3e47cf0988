// CPU ORACLE (test infrastructure) — LidarSlam::Slam pipeline driver restatement.
// Reference: slam_lib/src/Slam.cxx (line ranges per function).
#include "orc_slam.h"
#include "../lidarslam_b200/csrc/common/lcp_core.h"
#include <chrono>
#include <cstring>

namespace orc {

static inline double StampToSec(uint64_t us) { return us * 1e-6; }   // Utils::PclStampToSec

Slam::Slam()
{
  lsb_default_params(&P);
  KeyPointsExtractors[0] = Extractor();
  // Slam.cxx:153-157 (resolution is snapped with the leaf size at call time: 0.2 -> 10.0)
  SetVoxelGridResolution(10.);
  SetVoxelGridSize(50);
  SetVoxelGridLeafSize(LSB_EDGE, 0.30);
  SetVoxelGridLeafSize(LSB_PLANE, 0.60);
  SetVoxelGridLeafSize(LSB_BLOB, 0.30);
  Reset();
}

void Slam::Reset(bool resetLog)
{
  ClearMaps();
  KfLastPose = Iso3d::identity();
  KfCounter = 0;
  Tworld = PreviousTworld = Trelative = Iso3d::identity();
  WithinFrameMotion.reset();
  std::memset(&LocalizationUncertainty, 0, sizeof(LocalizationUncertainty));
  CurrentFrames.clear();
  CurrentStamp = 0;
  for (int k = 0; k < 3; ++k) {
    CurrentKeypoints[k].clear(); PreviousKeypoints[k].clear(); CurrentWorldKeypoints[k].clear();
    EgoMotionMatchingResults[k] = MatchingResults();
    LocalizationMatchingResults[k] = MatchingResults();
  }
  if (resetLog) { NbrFrameProcessed = 0; LogTrajectory.clear(); }
  std::memset(&Stats, 0, sizeof(Stats));
}

void Slam::SetWorldTransformFromGuess(const Iso3d& T)
{
  Tworld = T;
  PreviousTworld = T;
  for (int k = 0; k < 3; ++k) CurrentKeypoints[k].clear();
}

Iso3d Slam::GetBaseToLidarOffset(int dev) const
{
  auto it = BaseToLidarOffsets.find(dev);
  return it != BaseToLidarOffsets.end() ? it->second : Iso3d::identity();
}

//-----------------------------------------------------------------------------
int Slam::AddFrames(const std::vector<Frame>& frames)
{
  const auto t0 = std::chrono::steady_clock::now();
  Warnings = 0;
  if (!CheckFrames(frames)) return LSB_W_FRAME_IGNORED;
  CurrentFrames = frames;
  CurrentStamp = frames[0].stamp_us;
  CurrentTime = StampToSec(CurrentStamp);
  std::memset(&Stats, 0, sizeof(Stats));
  for (const Frame& f : frames) Stats.n_points += (uint32_t)f.n;

  auto tick = [] { return std::chrono::steady_clock::now(); };
  auto ms = [](std::chrono::steady_clock::time_point a, std::chrono::steady_clock::time_point b) { return std::chrono::duration<float, std::milli>(b - a).count(); };
  auto a = tick();
  ExtractKeypoints();
  auto b = tick(); Stats.ms_extract = ms(a, b);
  ComputeEgoMotion();
  a = tick(); Stats.ms_ego = ms(b, a);
  if (WheelOdomManager.CanBeUsed() || ImuManager.CanBeUsed()) {   // Slam.cxx:256-261, 346-352
    bool warn = false;
    WheelOdomManager.ComputeWheelAbsoluteConstraint(CurrentTime, &warn);
    ImuManager.ComputeGravityConstraint(CurrentTime, &warn);
    if (warn) Warnings |= LSB_W_SENSOR_TIME;
  }
  Localization();
  b = tick(); Stats.ms_localization = ms(a, b);
  if (P.overlap_sampling_ratio > 0) EstimateOverlap();       // :272-276
  else OverlapEstimation = -1.f;                             // SetOverlapSamplingRatio (:1365-1366)
  if (P.time_window_duration > 0.f) {                        // :277-278
    const int w = lsb::check_motion_limits(LogTrajectory, Tworld, CurrentTime, P.time_window_duration, P.velocity_limits, P.acceleration_limits,
                                           NbrFrameProcessed, MotionLimits);
    if (w & 1) Warnings |= LSB_W_TIME_WINDOW;
    if (w & 2) Warnings |= LSB_W_MOTION_LIMITS;
  }
  if (P.map_update == LSB_MAP_ADD_KPTS_TO_FIXED_MAP || P.map_update == LSB_MAP_UPDATE) UpdateMapsUsingTworld();
  a = tick(); Stats.ms_map_update = ms(b, a);
  LogCurrentFrameState(CurrentTime);
  for (int k = 0; k < 3; ++k) {
    Stats.n_keypoints[k] = (uint32_t)CurrentKeypoints[k].size();
    Stats.n_submap[k] = (uint32_t)LocalMaps[k].SubMap.size();
    Stats.n_map[k] = LocalMaps[k].Size();
    Stats.n_matches[k] = LocalizationMatchingResults[k].NbMatches();
    Stats.ego_n_matches[k] = EgoMotionMatchingResults[k].NbMatches();
  }
  Latency = std::chrono::duration<double>(tick() - t0).count();
  Stats.ms_total = (float)(Latency * 1e3);
  NbrFrameProcessed++;
  return Warnings;
}

//-----------------------------------------------------------------------------
bool Slam::CheckFrames(const std::vector<Frame>& frames)
{
  bool allEmpty = true;
  for (const Frame& f : frames) if (f.n > 0) allEmpty = false;
  if (frames.empty() || allEmpty) return false;
  if (frames[0].stamp_us == CurrentStamp) return false;   // :727-731
  for (size_t i = 0; i < frames.size(); ++i) {
    const unsigned dropped = frames[i].seq - PreviousFramesSeq[(int)i] - 1;
    if (PreviousFramesSeq[(int)i] > 0 && dropped > 0) Warnings |= LSB_W_FRAMES_DROPPED;
    PreviousFramesSeq[(int)i] = frames[i].seq;
  }
  return true;
}

//-----------------------------------------------------------------------------
void Slam::ExtractKeypoints()
{
  for (int k = 0; k < 3; ++k) { PreviousKeypoints[k] = std::move(CurrentKeypoints[k]); CurrentKeypoints[k].clear(); }
  struct Extracted { Cloud kp[3]; uint64_t stamp; };
  std::vector<Extracted> ext;
  for (const Frame& f : CurrentFrames) {
    if (f.n == 0) continue;
    int dev = f.pts[0].device_id;
    if (!KeyPointsExtractors.count(dev)) {   // :764-780
      if (KeyPointsExtractors.size() == 1) dev = 0;
      else continue;
    }
    Extractor& ke = KeyPointsExtractors[dev];
    ke.NbThreads = P.nb_threads;
    ke.ComputeKeyPoints(f.pts, f.n);
    Extracted e;
    for (int k = 0; k < 3; ++k) e.kp[k] = std::move(ke.Keypoints[k]);
    e.stamp = f.stamp_us;
    ext.push_back(std::move(e));
  }
  // AggregateFrames(keypoints[k], false)  (:1512-1578)
  for (int k = 0; k < 3; ++k) {
    if (!P.use_keypoints[k]) continue;
    Cloud& agg = CurrentKeypoints[k];
    for (const Extracted& e : ext) {
      const Cloud& c = e.kp[k];
      if (c.empty()) continue;
      const size_t start = agg.size();
      agg.insert(agg.end(), c.begin(), c.end());
      const double timeOffset = StampToSec(e.stamp) - StampToSec(CurrentStamp);
      const Iso3d tf = GetBaseToLidarOffset(c.front().device_id);
      const bool ident = lsb::iso_is_approx(tf, Iso3d::identity());
#pragma omp parallel for num_threads(P.nb_threads)
      for (long i = (long)start; i < (long)agg.size(); ++i) {
        agg[i].time += timeOffset;
        if (!ident) lsb::transform_point_f(tf, agg[i].x, agg[i].y, agg[i].z);
      }
    }
  }
}

//-----------------------------------------------------------------------------
static lsb_match_params EgoParams(const lsb_params& P)
{
  lsb_match_params m;
  m.single_edge_per_ring = 1;
  m.max_neighbors_distance = P.ego_motion_max_neighbors_distance;
  m.edge_nb_neighbors = P.ego_motion_edge_nb_neighbors;
  m.edge_min_nb_neighbors = P.ego_motion_edge_min_nb_neighbors;
  m.edge_max_model_error = P.ego_motion_edge_max_model_error;
  m.plane_nb_neighbors = P.ego_motion_plane_nb_neighbors;
  m.planarity_threshold = P.ego_motion_planarity_threshold;
  m.plane_max_model_error = P.ego_motion_plane_max_model_error;
  m.blob_nb_neighbors = 10;   // KeypointsMatcher.h:70 default (unused by ego-motion)
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

void Slam::ComputeEgoMotion()
{
  Trelative = Iso3d::identity();
  if (LogTrajectory.size() >= 2 && (P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION || P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION_AND_REGISTRATION)) {
    const double t = StampToSec(CurrentStamp);
    const double t1 = LogTrajectory[LogTrajectory.size() - 1].time;
    const double t0 = LogTrajectory[LogTrajectory.size() - 2].time;
    if (std::abs((t - t1) / (t1 - t0)) > P.max_extrapolation_ratio) Warnings |= LSB_W_EXTRAPOLATION;
    else {
      const Iso3d next = lsb::linear_interpolation(PreviousTworld, Tworld, t, t0, t1);
      Trelative = lsb::inverse(Tworld) * next;
    }
  }
  if (P.ego_motion == LSB_EGO_REGISTRATION || P.ego_motion == LSB_EGO_MOTION_EXTRAPOLATION_AND_REGISTRATION) {
    KdTree trees[2];
    for (int k = 0; k < 2; ++k) trees[k].Reset(&PreviousKeypoints[k]);
    TotalMatchedKeypoints = 0;
    lsb_match_params mp = EgoParams(P);
    for (unsigned icpIter = 0; icpIter < P.ego_motion_icp_max_iter; ++icpIter) {
      const double iterRatio = icpIter / static_cast<double>(P.ego_motion_icp_max_iter - 1);
      mp.saturation_distance = (1 - iterRatio) * P.ego_motion_init_saturation_distance + iterRatio * P.ego_motion_final_saturation_distance;
      Matcher matcher(mp, Trelative, P.nb_threads);
      for (int k = 0; k < 2; ++k) {
        EgoMotionMatchingResults[k] = matcher.BuildMatchResiduals(CurrentKeypoints[k], trees[k], k);
        Stats.knn_queries += (uint32_t)CurrentKeypoints[k].size();
      }
      Stats.ego_icp_iters++;
      TotalMatchedKeypoints = 0;
      for (int k = 0; k < 2; ++k) TotalMatchedKeypoints += EgoMotionMatchingResults[k].NbMatches();
      if (TotalMatchedKeypoints < P.min_nb_matched_keypoints) { Warnings |= LSB_W_EGO_SKIPPED; break; }
      Optimizer opt;
      opt.TwoDMode = P.two_d_mode != 0;
      opt.SetPosePrior(Trelative);
      opt.LMMaxIter = P.ego_motion_lm_max_iter;
      opt.NbThreads = P.nb_threads;
      opt.Sat = mp.saturation_distance;
      for (int k = 0; k < 2; ++k) opt.AddResiduals(EgoMotionMatchingResults[k]);
      const int nSuccessful = opt.Solve();
      Stats.ego_lm_iters += (uint32_t)opt.State.iter;
      Trelative = opt.GetOptimizedPose();
      if (nSuccessful == 1) break;
    }
  }
}

//-----------------------------------------------------------------------------
void Slam::Localization()
{
  PreviousTworld = Tworld;
  Tworld = PreviousTworld * Trelative;
  if (P.undistortion) { InitUndistortion(); RefineUndistortion(); }

  for (int k = 0; k < 3; ++k) {
    if (!P.use_keypoints[k] || LocalMaps[k].IsSubMapKdTreeValid()) continue;
    if (P.map_update == LSB_MAP_NONE) { LocalMaps[k].BuildSubMapKdTree(); continue; }
    if (LocalMaps[k].IsTimeThreshold()) LocalMaps[k].ClearOldPoints(CurrentTime);
    // pcl::transformPointCloud (PCL <= 1.10 explicit form) + pcl::getMinMax3D  (:1026-1029)
    const Cloud& c = CurrentKeypoints[k];
    const float FM = std::numeric_limits<float>::max();
    Arr3f mn{FM, FM, FM}, mx{-FM, -FM, -FM};
    for (const Point& p : c) {
      const double x = p.x, y = p.y, z = p.z;
      const float wx = (float)(Tworld.R.m[0] * x + Tworld.R.m[1] * y + Tworld.R.m[2] * z + Tworld.t.x);
      const float wy = (float)(Tworld.R.m[3] * x + Tworld.R.m[4] * y + Tworld.R.m[5] * z + Tworld.t.y);
      const float wz = (float)(Tworld.R.m[6] * x + Tworld.R.m[7] * y + Tworld.R.m[8] * z + Tworld.t.z);
      mn.x = std::min(mn.x, wx); mn.y = std::min(mn.y, wy); mn.z = std::min(mn.z, wz);
      mx.x = std::max(mx.x, wx); mx.y = std::max(mx.y, wy); mx.z = std::max(mx.z, wz);
    }
    LocalMaps[k].BuildSubMapKdTree(mn, mx, (int)(c.size() / 2));
  }

  TotalMatchedKeypoints = 0;
  lsb_match_params mp = LocParams(P);
  for (unsigned icpIter = 0; icpIter < P.localization_icp_max_iter; ++icpIter) {
    const double iterRatio = icpIter / static_cast<double>(P.localization_icp_max_iter - 1);
    mp.saturation_distance = (1 - iterRatio) * P.localization_init_saturation_distance + iterRatio * P.localization_final_saturation_distance;
    Matcher matcher(mp, Tworld, P.nb_threads);
    for (int k = 0; k < 3; ++k) {
      LocalizationMatchingResults[k] = matcher.BuildMatchResiduals(CurrentKeypoints[k], LocalMaps[k].Tree, k);
      Stats.knn_queries += (uint32_t)CurrentKeypoints[k].size();
    }
    Stats.loc_icp_iters++;
    TotalMatchedKeypoints = 0;
    for (int k = 0; k < 3; ++k) TotalMatchedKeypoints += LocalizationMatchingResults[k].NbMatches();
    if (TotalMatchedKeypoints < P.min_nb_matched_keypoints) {
      Trelative = Iso3d::identity();
      Tworld = PreviousTworld;
      if (P.undistortion) WithinFrameMotion.set_transforms(Iso3d::identity(), Iso3d::identity());
      Warnings |= LSB_W_LOC_SKIPPED;
      break;
    }
    Optimizer opt;
    opt.TwoDMode = P.two_d_mode != 0;
    opt.SetPosePrior(Tworld);
    opt.LMMaxIter = P.localization_lm_max_iter;
    opt.NbThreads = P.nb_threads;
    opt.Sat = mp.saturation_distance;
    for (int k = 0; k < 3; ++k) opt.AddResiduals(LocalizationMatchingResults[k]);
    opt.SensorTerms = lsb::MakeSensorTerms(WheelOdomManager, ImuManager);   // Slam.cxx:1123-1131
    const int nSuccessful = opt.Solve();
    Stats.loc_lm_iters += (uint32_t)opt.State.iter;
    Tworld = opt.GetOptimizedPose();
    Trelative = lsb::inverse(PreviousTworld) * Tworld;
    if (P.undistortion == LSB_UNDISTORTION_REFINED) RefineUndistortion();
    if (nSuccessful == 1 || icpIter == P.localization_icp_max_iter - 1) {
      LocalizationUncertainty = opt.EstimateRegistrationError();
      break;
    }
  }
}

//-----------------------------------------------------------------------------
void Slam::UpdateMapsUsingTworld()
{
  const Iso3d motion = lsb::inverse(KfLastPose) * Tworld;
  const double trans = lsb::norm(motion.t);
  const double rot = lsb::rotation_angle(motion.R);
  const double thresholdCoef = std::min(KfCounter / 10., 1.);
  unsigned nbMapKpts = 0;
  for (const auto& m : LocalMaps) nbMapKpts += m.Size();
  const bool isNewKeyFrame = nbMapKpts < P.min_nb_matched_keypoints * 10 || trans >= thresholdCoef * P.kf_distance_threshold ||
                             rot >= (thresholdCoef * P.kf_angle_threshold) / 180. * M_PI;
  if (!isNewKeyFrame) return;
  Stats.is_keyframe = 1;
  KfCounter++;
  KfLastPose = Tworld;
  for (int k = 0; k < 3; ++k) {   // GetKeypoints(k, true) -> TransformPointCloud (:1491-1509)
    CurrentWorldKeypoints[k] = CurrentKeypoints[k];
    Cloud& c = CurrentWorldKeypoints[k];
#pragma omp parallel for num_threads(P.nb_threads)
    for (long i = 0; i < (long)c.size(); ++i) lsb::transform_point_f(Tworld, c[i].x, c[i].y, c[i].z);
  }
#pragma omp parallel for num_threads(std::min(P.nb_threads, 3))
  for (int k = 0; k < 3; ++k)
    if (P.use_keypoints[k]) LocalMaps[k].Add(CurrentWorldKeypoints[k], false, CurrentTime);
}

//-----------------------------------------------------------------------------
void Slam::LogCurrentFrameState(double time)
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
Iso3d Slam::InterpolateScanPose(double time)
{
  if (LogTrajectory.empty()) return Tworld;
  const double prevPoseTime = LogTrajectory.back().time;
  const double currPoseTime = StampToSec(CurrentStamp);
  if (std::abs(time / (currPoseTime - prevPoseTime)) > P.max_extrapolation_ratio) { Warnings |= LSB_W_EXTRAPOLATION; return Tworld; }
  return lsb::linear_interpolation(PreviousTworld, Tworld, currPoseTime + time, prevPoseTime, currPoseTime);
}

void Slam::InitUndistortion()
{
  double first = std::numeric_limits<double>::max(), last = std::numeric_limits<double>::lowest();
  for (int k = 0; k < 3; ++k)
    for (const Point& p : CurrentKeypoints[k]) { first = std::min(first, p.time); last = std::max(last, p.time); }
  WithinFrameMotion.set_times(first, last);
  WithinFrameMotion.set_transforms(Iso3d::identity(), Iso3d::identity());
  if (WithinFrameMotion.t1 - WithinFrameMotion.t0 < 1e-6) { Warnings |= LSB_W_TIME_FIELD; WithinFrameMotion.set_times(0., 0.); }
  else if (WithinFrameMotion.t1 - WithinFrameMotion.t0 > 10.) Warnings |= LSB_W_TIME_FIELD;
}

void Slam::RefineUndistortion()
{
  const Iso3d prevBegin = WithinFrameMotion.H0(), prevEnd = WithinFrameMotion.H1();
  const Iso3d wBegin = InterpolateScanPose(WithinFrameMotion.t0);
  const Iso3d wEnd = InterpolateScanPose(WithinFrameMotion.t1);
  const Iso3d baseToWorld = lsb::inverse(Tworld);
  const Iso3d newBegin = baseToWorld * wBegin, newEnd = baseToWorld * wEnd;
  WithinFrameMotion.set_transforms(newBegin, newEnd);
  lsb::Interpolator interp = WithinFrameMotion;
  interp.set_transforms(newBegin * lsb::inverse(prevBegin), newEnd * lsb::inverse(prevEnd));
  for (int k = 0; k < 3; ++k) {
    Cloud& c = CurrentKeypoints[k];
#pragma omp parallel for num_threads(P.nb_threads)
    for (long i = 0; i < (long)c.size(); ++i) {
      const Iso3d T = interp.at(c[i].time);
      lsb::transform_point_f(T, c[i].x, c[i].y, c[i].z);
    }
  }
}

//-----------------------------------------------------------------------------
Cloud Slam::GetRegisteredFrame()
{
  Cloud agg;
  for (const Frame& f : CurrentFrames) {
    if (f.n == 0) continue;
    const size_t start = agg.size();
    agg.insert(agg.end(), f.pts, f.pts + f.n);
    const double timeOffset = StampToSec(f.stamp_us) - StampToSec(CurrentStamp);
    const Iso3d baseToLidar = GetBaseToLidarOffset(f.pts[0].device_id);
    if (P.undistortion) {
      lsb::Interpolator interp = WithinFrameMotion;
      interp.set_transforms(Tworld * WithinFrameMotion.H0() * baseToLidar, Tworld * WithinFrameMotion.H1() * baseToLidar);
      for (size_t i = start; i < agg.size(); ++i) {
        agg[i].time += timeOffset;
        const Iso3d T = interp.at(agg[i].time);
        lsb::transform_point_f(T, agg[i].x, agg[i].y, agg[i].z);
      }
    } else {
      const Iso3d tf = Tworld * baseToLidar;
      const bool ident = lsb::iso_is_approx(tf, Iso3d::identity());
      for (size_t i = start; i < agg.size(); ++i) {
        agg[i].time += timeOffset;
        if (!ident) lsb::transform_point_f(tf, agg[i].x, agg[i].y, agg[i].z);
      }
    }
  }
  return agg;
}

// Slam::EstimateOverlap (:1370-1388) + Confidence::LCPEstimator (ConfidenceEstimators.cxx:27-65)
void Slam::EstimateOverlap()
{
  const Cloud agg = GetRegisteredFrame();
  std::vector<int> maps;
  for (int k = 0; k < 3; ++k)
    if (P.use_keypoints[k] && LocalMaps[k].IsSubMapKdTreeValid()) maps.push_back(k);
  const float ratio = P.overlap_sampling_ratio;
  const int nbPoints = (int)(agg.size() * ratio);
  if (nbPoints == 0 || maps.empty()) { OverlapEstimation = -1.f; return; }
  std::vector<float> proba((size_t)nbPoints, 0.f);
#pragma omp parallel for num_threads(P.nb_threads > 0 ? P.nb_threads : 1) schedule(static)
  for (int n = 0; n < nbPoints; ++n) {
    const Point& pt = agg[lsb::lcp_sample_index(n, ratio)];
    const float q[3] = {pt.x, pt.y, pt.z};
    float best = 0.f;
    for (int k : maps) {
      int idx;
      float d2;
      if (LocalMaps[k].Tree.Knn(q, 1, &idx, &d2)) {
        const float pr = lsb::lcp_probability(d2, lsb::lcp_sq_threshold(LocalMaps[k].LeafSize));
        if (pr > best) best = pr;
      }
    }
    proba[n] = best;
  }
  // canonical radix-32 tree sum (common/lcp_core.h)
  auto fold32 = [](float* v) { for (int off = 16; off >= 1; off >>= 1) for (int l = 0; l < off; ++l) v[l] += v[l + off]; };
  const size_t nChunks = ((size_t)nbPoints + LSB_LCP_CHUNK - 1) / LSB_LCP_CHUNK;
  const size_t nBlocks = (nChunks + 31) / 32;
  std::vector<float> chunk(nBlocks * 32, 0.f);
  for (size_t c = 0; c < nChunks; ++c) {
    float group[32];
    for (int g = 0; g < 32; ++g) {
      float lane[32];
      for (int l = 0; l < 32; ++l) {
        const size_t i = c * LSB_LCP_CHUNK + (size_t)g * 32 + l;
        lane[l] = (i < (size_t)nbPoints) ? proba[i] : 0.f;
      }
      fold32(lane);
      group[g] = lane[0];
    }
    fold32(group);
    chunk[c] = group[0];
  }
  float lcp = 0.f;
  for (size_t b = 0; b < nBlocks; ++b) {
    float v[32];
    for (int l = 0; l < 32; ++l) v[l] = chunk[b * 32 + l];
    fold32(v);
    lcp = (b == 0) ? v[0] : lcp + v[0];
  }
  OverlapEstimation = lcp / nbPoints;
}

}  // namespace orc

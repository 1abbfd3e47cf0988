// CPU ORACLE (test infrastructure) — LidarSlam::Slam pipeline driver restatement
// (reference slam_lib/src/Slam.cxx, defaults slam_lib/include/LidarSlam/Slam.h:403-694).
#pragma once
#include "../lidarslam_b200/csrc/common/motion_limits.h"
#include "../lidarslam_b200/csrc/common/sensor_constraints.h"
#include "orc_extractor.h"
#include "orc_matcher.h"
#include "orc_optimizer.h"
#include "orc_rolling_grid.h"
#include <deque>
#include <map>

namespace orc {

struct Frame { const Point* pts; size_t n; uint64_t stamp_us; uint32_t seq; };

struct Slam {
  lsb_params P;
  std::map<int, Extractor> KeyPointsExtractors;         // Slam.h:500
  std::map<int, Iso3d> BaseToLidarOffsets;              // Slam.h:505
  RollingGrid LocalMaps[3];                             // Slam.h:553

  Iso3d Tworld, PreviousTworld, Trelative;              // Slam.h:461-467
  lsb::Interpolator WithinFrameMotion;                  // Slam.h:481
  struct LoggedPose { Iso3d T; double time; };
  std::deque<LoggedPose> LogTrajectory;                 // Slam.h:487
  Iso3d KfLastPose;                                     // Slam.h:531
  int KfCounter = 0;
  unsigned NbrFrameProcessed = 0;
  double CurrentTime = 0.;
  uint64_t CurrentStamp = 0;                            // CurrentFrames[0]->header.stamp
  std::map<int, unsigned> PreviousFramesSeq;
  double Latency = 0.;

  // CurrentRawKeypoints and CurrentUndistortedKeypoints alias each other in the
  // reference (Slam.cxx:984 copies shared_ptrs): one storage here.
  Cloud CurrentKeypoints[3], PreviousKeypoints[3], CurrentWorldKeypoints[3];
  std::vector<Frame> CurrentFrames;
  std::vector<Cloud> CurrentFramesCopy;

  MatchingResults EgoMotionMatchingResults[3], LocalizationMatchingResults[3];
  lsb_registration_error LocalizationUncertainty;
  unsigned TotalMatchedKeypoints = 0;
  float OverlapEstimation = -1.f;
  lsb::MotionLimitsState MotionLimits;
  lsb::WheelOdometryManager WheelOdomManager;   // Slam.h:572
  lsb::ImuManager ImuManager;                   // Slam.h:578
  std::vector<lsb_point> Converted;             // last frame produced by orc_convert_* (lidar_conversions nodes)
  lsb_frame_stats Stats;
  int Warnings = 0;

  Slam();
  void Reset(bool resetLog = true);                     // Slam.cxx:164-210
  int AddFrames(const std::vector<Frame>& frames);      // Slam.cxx:230-344
  void ClearMaps() { for (auto& m : LocalMaps) m.Reset(); }
  void SetWorldTransformFromGuess(const Iso3d& T);      // Slam.cxx:490-501
  Cloud GetRegisteredFrame();                           // Slam.cxx:660-667 / 1512-1578 (world)
  void SetVoxelGridLeafSize(int k, double s) { LocalMaps[k].LeafSize = s; }
  void SetVoxelGridSize(int s) { for (auto& m : LocalMaps) m.SetGridSize(s); }
  void SetVoxelGridResolution(double r) { for (auto& m : LocalMaps) m.SetVoxelResolution(r); }

 private:
  bool CheckFrames(const std::vector<Frame>& frames);   // :709-743
  void ExtractKeypoints();                               // :746-810
  void ComputeEgoMotion();                               // :813-972
  void Localization();                                   // :975-1175
  void UpdateMapsUsingTworld();                          // :1178-1222
  void LogCurrentFrameState(double time);                // :1225-1264
  Iso3d InterpolateScanPose(double time);                // :1271-1285
  void InitUndistortion();                               // :1288-1319
  void RefineUndistortion();                             // :1322-1352
  void EstimateOverlap();                                // :1370-1388 + ConfidenceEstimators.cxx:27-65
  Iso3d GetBaseToLidarOffset(int dev) const;             // :1625-1628
};

}  // namespace orc

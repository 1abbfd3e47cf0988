// LidarSlam::Slam — header-only C++ host interface ABOVE the C ABI of liblidarslam_b200
// (include/lidarslam_b200.h): the same class, method names, argument meaning and error behaviour
// as the reference's public interface (slam_lib/include/LidarSlam/Slam.h:98-394), so that callers
// written against it (ROS node LidarSlamNode.cxx:158-178, 519-622; ParaView filter
// vtkSlam.cxx:217-322) compile against this header and link against the B200 library.
//
// Types.  The reference's signatures use pcl::PointCloud<LidarPoint>, Eigen::Isometry3d and
// LidarSlam::Transform.  PCL and Eigen are not part of this repository: by default this header
// defines minimal stand-ins with the members the wrappers touch (`points`, `header.stamp/seq/
// frame_id`, `size()`, `empty()`; a 4x4 row-major isometry).  Define LIDARSLAM_B200_WITH_PCL before
// including it where PCL + Eigen exist and the real types are used instead (LidarPoint and
// lsb_point share one 32-byte layout, LidarPoint.h:31-64, so clouds are passed without a copy).
//
// Error behaviour follows the reference: no exceptions from the per-frame calls, no status codes;
// soft failures print the reference's warnings on stderr (Utilities.h:62-63).  The constructor
// throws if no CUDA device is present: there is no CPU fallback.
#pragma once
#include <array>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>
#include "../lidarslam_b200.h"

#ifdef LIDARSLAM_B200_WITH_PCL
#include <Eigen/Geometry>
#include <pcl/point_cloud.h>
#include <LidarSlam/LidarPoint.h>
#include <LidarSlam/Transform.h>
#endif

namespace LidarSlam
{

// Enums.h:30-151
enum Keypoint { EDGE = 0, PLANE = 1, BLOB = 2, nKeypointTypes };
static const std::vector<Keypoint> KeypointTypes = {EDGE, PLANE, BLOB};
static const std::map<Keypoint, std::string> KeypointTypeNames = {{EDGE, "edge"}, {PLANE, "plane"}, {BLOB, "blob"}};
enum UndistortionMode { NONE = 0, ONCE = 1, REFINED = 2 };
enum class EgoMotionMode { NONE = 0, MOTION_EXTRAPOLATION = 1, REGISTRATION = 2, MOTION_EXTRAPOLATION_AND_REGISTRATION = 3 };
enum class MappingMode { NONE = 0, ADD_KPTS_TO_FIXED_MAP = 1, UPDATE = 2 };
enum class SamplingMode { FIRST = 0, LAST = 1, MAX_INTENSITY = 2, CENTER_POINT = 3, CENTROID = 4 };
enum PCDFormat { ASCII = 0, BINARY = 1, BINARY_COMPRESSED = 2 };            // PointCloudStorage.h:60-65

#ifndef LIDARSLAM_B200_WITH_PCL
using LidarPoint = lsb_point;                                                // LidarPoint.h:31-64, same 32 bytes

// the members of pcl::PointCloud<LidarPoint> the SLAM interface and its wrappers use
struct LidarPointCloud {
  using Ptr = std::shared_ptr<LidarPointCloud>;
  using ConstPtr = std::shared_ptr<const LidarPointCloud>;
  struct Header { std::uint64_t stamp = 0; std::uint32_t seq = 0; std::string frame_id; } header;   // stamp in microseconds
  std::vector<LidarPoint> points;
  std::size_t size() const { return points.size(); }
  bool empty() const { return points.empty(); }
  void resize(std::size_t n) { points.resize(n); }
  void reserve(std::size_t n) { points.reserve(n); }
  void push_back(const LidarPoint& p) { points.push_back(p); }
  LidarPoint& operator[](std::size_t i) { return points[i]; }
  const LidarPoint& operator[](std::size_t i) const { return points[i]; }
};

// Eigen::Isometry3d stand-in: 4x4 row-major homogeneous matrix
struct Isometry3d {
  double m[16];
  Isometry3d() { std::memset(m, 0, sizeof(m)); m[0] = m[5] = m[10] = m[15] = 1.; }
  static Isometry3d Identity() { return Isometry3d(); }
  double& operator()(int r, int c) { return m[4 * r + c]; }
  double operator()(int r, int c) const { return m[4 * r + c]; }
  const double* data() const { return m; }
  double* data() { return m; }
};

// Transform.h:25-88
struct Transform {
  Isometry3d transform;
  double time = 0.;
  std::string frameid;
  Transform() = default;
  Transform(const Isometry3d& t, double stamp = 0., const std::string& frame = "") : transform(t), time(stamp), frameid(frame) {}
  static Transform Identity() { return Transform(); }
  double x() const { return transform(0, 3); }
  double y() const { return transform(1, 3); }
  double z() const { return transform(2, 3); }
  void SetIsometry(const Isometry3d& t) { transform = t; }
  Isometry3d GetIsometry() const { return transform; }
};
namespace detail {
inline void ToRowMajor(const Isometry3d& t, double out[16]) { std::memcpy(out, t.m, sizeof(t.m)); }
inline Isometry3d FromRowMajor(const double in[16]) { Isometry3d t; std::memcpy(t.m, in, sizeof(t.m)); return t; }
}  // namespace detail
#else
using LidarPointCloud = pcl::PointCloud<LidarPoint>;
using Isometry3d = Eigen::Isometry3d;
namespace detail {
inline void ToRowMajor(const Eigen::Isometry3d& t, double out[16])
{
  for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out[4 * r + c] = t.matrix()(r, c);
}
inline Eigen::Isometry3d FromRowMajor(const double in[16])
{
  Eigen::Isometry3d t = Eigen::Isometry3d::Identity();
  for (int r = 0; r < 3; ++r) for (int c = 0; c < 4; ++c) t.matrix()(r, c) = in[4 * r + c];
  return t;
}
}  // namespace detail
#endif

namespace SensorConstraints {                                               // SensorConstraints.h:13-23
struct WheelOdomMeasurement { double Time = 0., Distance = 0.; };
struct GravityMeasurement { double Time = 0.; double Acceleration[3] = {0., 0., 0.}; };
}  // namespace SensorConstraints

// Parameters of SpinningSensorKeypointExtractor (SSKE.h:44-74, 126-167): the object the wrappers configure and install per
// device with Slam::SetKeyPointsExtractor.  The extraction itself runs inside the library.
class SpinningSensorKeypointExtractor {
 public:
  SpinningSensorKeypointExtractor() { lsb_default_ke_params(&P); }
#define LSB_KE_PARAM(Name, field, type) \
  type Get##Name() const { return (type)P.field; } \
  void Set##Name(type v) { P.field = v; }
  LSB_KE_PARAM(NeighborWidth, neighbor_width, int)
  LSB_KE_PARAM(MinDistanceToSensor, min_distance_to_sensor, float)
  LSB_KE_PARAM(MinBeamSurfaceAngle, min_beam_surface_angle, float)
  LSB_KE_PARAM(PlaneSinAngleThreshold, plane_sin_angle_threshold, float)
  LSB_KE_PARAM(EdgeSinAngleThreshold, edge_sin_angle_threshold, float)
  LSB_KE_PARAM(DistToLineThreshold, dist_to_line_threshold, float)
  LSB_KE_PARAM(EdgeDepthGapThreshold, edge_depth_gap_threshold, float)
  LSB_KE_PARAM(EdgeSaliencyThreshold, edge_saliency_threshold, float)
  LSB_KE_PARAM(EdgeIntensityGapThreshold, edge_intensity_gap_threshold, float)
  LSB_KE_PARAM(AzimuthalResolution, azimuthal_resolution, float)
#undef LSB_KE_PARAM
  void SetNbThreads(int) {}                                                 // kept for API parity: the GPU path has no thread count
  lsb_ke_params P;
};

#define LSB_PRINT_WARNING(s) std::fprintf(stderr, "\033[1;33m[WARNING] %s\033[0m\n", s)     // Utilities.h:62
#define LSB_PRINT_ERROR(s) std::fprintf(stderr, "\033[1;31m[ERROR] %s\033[0m\n", s)         // Utilities.h:63

class Slam
{
public:
  using Point = LidarPoint;
  using PointCloud = LidarPointCloud;
  using KeypointExtractorPtr = std::shared_ptr<SpinningSensorKeypointExtractor>;

  // ---- initialisation (Slam.cxx:143-210) --------------------------------------
  explicit Slam(int cudaDevice = 0)
  {
    if (lsb_create(&this->Gpu, cudaDevice) != LSB_OK) throw std::runtime_error(std::string("LidarSlam::Slam: ") + lsb_last_error());
    lsb_get_params(this->Gpu, &this->P);
    this->KeyPointsExtractors[0] = std::make_shared<SpinningSensorKeypointExtractor>();
  }
  ~Slam() { lsb_destroy(this->Gpu); }
  Slam(const Slam&) = delete;
  Slam& operator=(const Slam&) = delete;
  void Reset(bool resetLog = true) { lsb_reset(this->Gpu, resetLog ? 1 : 0); }

  // ---- main SLAM use (Slam.cxx:230-344) ---------------------------------------
  void AddFrame(const PointCloud::Ptr& pc) { this->AddFrames({pc}); }
  void AddFrames(const std::vector<PointCloud::Ptr>& frames)
  {
    std::vector<const lsb_point*> ptr;
    std::vector<std::size_t> n;
    std::vector<std::uint64_t> stamp;
    std::vector<std::uint32_t> seq;
    for (const auto& f : frames) {
      ptr.push_back(f && !f->empty() ? reinterpret_cast<const lsb_point*>(f->points.data()) : nullptr);   // same 32-byte layout
      n.push_back(f ? f->size() : 0);
      stamp.push_back(f ? (std::uint64_t)f->header.stamp : 0);
      seq.push_back(f ? (std::uint32_t)f->header.seq : 0);
    }
    this->PushExtractors();
    const int rc = lsb_add_frames(this->Gpu, (int)frames.size(), ptr.data(), n.data(), stamp.data(), seq.data());
    this->LastStatus = rc;
    if (rc < 0) { LSB_PRINT_ERROR(lsb_last_error()); return; }
    if (rc & LSB_W_FRAME_IGNORED) LSB_PRINT_ERROR("SLAM entry is an empty pointcloud or has the same timestamp as the previous one : frames ignored.");   // :720-731
    if (rc & LSB_W_FRAMES_DROPPED) LSB_PRINT_WARNING("SLAM dropped some frames.");                                                               // :734-740
    if (rc & LSB_W_TIME_FIELD) LSB_PRINT_WARNING("'time' field is not properly set and undistortion will be skipped/approximated.");            // :1307-1318
    if (rc & LSB_W_EGO_SKIPPED) LSB_PRINT_WARNING("Not enough keypoints, EgoMotion skipped for this frame.");                                   // :919-923
    if (rc & LSB_W_LOC_SKIPPED) LSB_PRINT_ERROR("Not enough keypoints, Localization skipped for this frame.");                                  // :1098-1107
    if (rc & LSB_W_EXTRAPOLATION) LSB_PRINT_WARNING("Unable to extrapolate scan pose from previous motion.");                                   // :829-830, :1278-1282
    if (rc & LSB_W_SENSOR_TIME) LSB_PRINT_WARNING("No external sensor measure corresponds to the current frame acquisition (times don't match)");  // SensorConstraints.cxx:19
    if (rc & LSB_W_MOTION_LIMITS) LSB_PRINT_WARNING("The pose does not comply with the motion limits.");                                         // :1482-1483
    if (!(rc & LSB_W_FRAME_IGNORED)) { this->CurrentFrames = frames; ++this->NbrFrameProcessed; }                                               // :237, :342
    this->PullExtractors();
  }

  // Not part of the reference interface (include/lidarslam_b200.h, "pipelining"): hands the NEXT input in ahead of time, e.g.
  // from a bag replay or a driver queue.  Its upload and keypoint extraction run while the current frame is in its ICP loop;
  // AddFrames called later with the same clouds picks the staged keypoints up.  The clouds are retained until then.
  void PrefetchFrames(const std::vector<PointCloud::Ptr>& frames)
  {
    std::vector<const lsb_point*> ptr;
    std::vector<std::size_t> n;
    std::vector<std::uint64_t> stamp;
    std::vector<std::uint32_t> seq;
    for (const auto& f : frames) {
      ptr.push_back(f && !f->empty() ? reinterpret_cast<const lsb_point*>(f->points.data()) : nullptr);
      n.push_back(f ? f->size() : 0);
      stamp.push_back(f ? (std::uint64_t)f->header.stamp : 0);
      seq.push_back(f ? (std::uint32_t)f->header.seq : 0);
    }
    this->PushExtractors();
    if (lsb_prefetch_frames(this->Gpu, (int)frames.size(), ptr.data(), n.data(), stamp.data(), seq.data()) < 0) { LSB_PRINT_ERROR(lsb_last_error()); return; }
    this->StagedFrames.push_back(frames);
    while (this->StagedFrames.size() > 3) this->StagedFrames.erase(this->StagedFrames.begin());
  }
  // 0: every AddFrames ends with its map update (the reference's timing); 1 (default): the map update is joined by the next frame
  void SetPipelining(bool deferMapUpdate) { lsb_set_pipelining(this->Gpu, deferMapUpdate ? 1 : 0); }

  Transform GetWorldTransform() const                                        // Slam.cxx:550-553
  {
    double T[16], t = 0.;
    lsb_get_world_transform(this->Gpu, T, &t);
    return Transform(detail::FromRowMajor(T), t, this->WorldFrameId);
  }
  Transform GetLatencyCompensatedWorldTransform() const                      // Slam.cxx:556-591
  {
    double T[16], t = 0.;
    lsb_get_latency_compensated_world_transform(this->Gpu, T, &t);
    return Transform(detail::FromRowMajor(T), t, this->WorldFrameId);
  }
  std::array<double, 36> GetTransformCovariance() const                      // Slam.cxx:594-597
  {
    std::array<double, 36> c;
    lsb_get_covariance(this->Gpu, c.data());
    return c;
  }
  std::vector<Transform> GetTrajectory() const                               // Slam.cxx:600-608
  {
    const long n = lsb_get_trajectory(this->Gpu, nullptr, nullptr, 0);
    std::vector<double> T(16 * (std::size_t)std::max(n, 0L)), t((std::size_t)std::max(n, 0L));
    if (n > 0) lsb_get_trajectory(this->Gpu, T.data(), t.data(), n);
    std::vector<Transform> out;
    for (long i = 0; i < n; ++i) out.emplace_back(detail::FromRowMajor(&T[16 * i]), t[i], this->WorldFrameId);
    return out;
  }

  PointCloud::Ptr GetMap(Keypoint k, bool clean = false) const              // Slam.cxx:670-677
  {
    return this->Cloud([&](lsb_point* o, long cap) { return lsb_get_map(this->Gpu, k, clean ? 1 : 0, o, cap); }, this->WorldFrameId);
  }
  PointCloud::Ptr GetTargetSubMap(Keypoint k) const                          // Slam.cxx:680-687
  {
    return this->Cloud([&](lsb_point* o, long cap) { return lsb_get_submap(this->Gpu, k, o, cap); }, this->WorldFrameId);
  }
  PointCloud::Ptr GetKeypoints(Keypoint k, bool worldCoordinates = false)    // Slam.cxx:690-702
  {
    return this->Cloud([&](lsb_point* o, long cap) { return lsb_get_keypoints(this->Gpu, k, worldCoordinates ? 1 : 0, o, cap); },
                       worldCoordinates ? this->WorldFrameId : this->BaseFrameId);
  }
  PointCloud::Ptr GetRegisteredFrame()                                       // Slam.cxx:660-667
  {
    return this->Cloud([&](lsb_point* o, long cap) { return lsb_get_registered_frame(this->Gpu, o, cap); }, this->WorldFrameId);
  }
  unsigned int GetNbrFrameProcessed() const { return this->NbrFrameProcessed; }

  // Slam.cxx:611-657 (the keys the wrappers publish)
  std::unordered_map<std::string, double> GetDebugInformation() const
  {
    lsb_frame_stats s;
    lsb_registration_error e;
    lsb_get_frame_stats(this->Gpu, &s);
    lsb_get_registration_error(this->Gpu, &e);
    std::unordered_map<std::string, double> m;
    m["EgoMotion: edges used"] = s.ego_n_matches[EDGE];
    m["EgoMotion: planes used"] = s.ego_n_matches[PLANE];
    m["Localization: edges used"] = s.n_matches[EDGE];
    m["Localization: planes used"] = s.n_matches[PLANE];
    m["Localization: blobs used"] = s.n_matches[BLOB];
    m["Localization: position error"] = e.position_error;
    m["Localization: orientation error"] = e.orientation_error;
    m["Confidence: overlap"] = lsb_get_overlap(this->Gpu);
    m["Confidence: comply motion limits"] = lsb_get_comply_motion_limits(this->Gpu);
    return m;
  }
  std::unordered_map<std::string, std::vector<double>> GetDebugArray() const
  {
    std::unordered_map<std::string, std::vector<double>> m;
    static const char* const stage[2] = {"EgoMotion", "Localization"};
    static const char* const kind[3] = {"edge", "plane", "blob"};
    for (int st = 0; st < 2; ++st)
      for (int k = 0; k < 3; ++k) {
        const long n = lsb_get_match_debug(this->Gpu, st, k, nullptr, nullptr, 0);
        if (n <= 0) continue;
        std::vector<std::uint8_t> status((std::size_t)n);
        std::vector<double> weight((std::size_t)n);
        lsb_get_match_debug(this->Gpu, st, k, status.data(), weight.data(), n);
        m[std::string(stage[st]) + ": " + kind[k] + " matches"] = std::vector<double>(status.begin(), status.end());
        m[std::string(stage[st]) + ": " + kind[k] + " weights"] = weight;
      }
    return m;
  }

  void SetWorldTransformFromGuess(const Transform& poseGuess)                // Slam.cxx:490-501
  {
    double T[16];
    detail::ToRowMajor(poseGuess.GetIsometry(), T);
    lsb_set_world_transform_from_guess(this->Gpu, T);
  }
  void SaveMapsToPCD(const std::string& filePrefix, PCDFormat pcdFormat = PCDFormat::BINARY_COMPRESSED, bool submap = true) const   // :504-517
  {
    lsb_save_maps_to_pcd(this->Gpu, filePrefix.c_str(), (int)pcdFormat, submap ? 1 : 0);
  }
  void LoadMapsFromPCD(const std::string& filePrefix, bool resetMaps = true) // Slam.cxx:520-543
  {
    lsb_load_maps_from_pcd(this->Gpu, filePrefix.c_str(), resetMaps ? 1 : 0);
  }

  // ---- general parameters (Slam.h:201-245) ------------------------------------
#define LSB_PARAM(Name, field, type) \
  type Get##Name() const { return (type)this->P.field; } \
  void Set##Name(type v) { this->P.field = static_cast<decltype(this->P.field)>(v); this->PushParams(); }
  LSB_PARAM(NbThreads, nb_threads, int)
  LSB_PARAM(Verbosity, verbosity, int)
  void SetUseBlobs(bool ub) { this->P.use_keypoints[BLOB] = ub ? 1 : 0; this->PushParams(); }
  bool GetUseBlobs() const { return this->P.use_keypoints[BLOB] != 0; }
  LSB_PARAM(EgoMotion, ego_motion, EgoMotionMode)
  LSB_PARAM(Undistortion, undistortion, UndistortionMode)
  LSB_PARAM(LoggingTimeout, logging_timeout, double)
  void SetBaseFrameId(const std::string& s) { this->BaseFrameId = s; }
  std::string GetBaseFrameId() const { return this->BaseFrameId; }
  void SetWorldFrameId(const std::string& s) { this->WorldFrameId = s; }
  std::string GetWorldFrameId() const { return this->WorldFrameId; }

  // ---- keypoints extraction (Slam.h:227-245, Slam.cxx:1601-1632) ----------------
  std::map<std::uint8_t, KeypointExtractorPtr> GetKeyPointsExtractors() const { return this->KeyPointsExtractors; }
  void SetKeyPointsExtractors(const std::map<std::uint8_t, KeypointExtractorPtr>& e) { this->KeyPointsExtractors = e; }
  KeypointExtractorPtr GetKeyPointsExtractor(std::uint8_t deviceId = 0) const
  {
    auto it = this->KeyPointsExtractors.find(deviceId);
    return it != this->KeyPointsExtractors.end() ? it->second : KeypointExtractorPtr();
  }
  void SetKeyPointsExtractor(KeypointExtractorPtr extractor, std::uint8_t deviceId = 0) { this->KeyPointsExtractors[deviceId] = extractor; }
  Isometry3d GetBaseToLidarOffset(std::uint8_t deviceId = 0) const
  {
    auto it = this->BaseToLidarOffsets.find(deviceId);
    return it != this->BaseToLidarOffsets.end() ? it->second : Isometry3d::Identity();
  }
  void SetBaseToLidarOffset(const Isometry3d& transform, std::uint8_t deviceId = 0)
  {
    this->BaseToLidarOffsets[deviceId] = transform;
    double T[16];
    detail::ToRowMajor(transform, T);
    lsb_set_base_to_lidar(this->Gpu, deviceId, T);
  }

  // ---- optimization parameters (Slam.h:251-328) ---------------------------------
  LSB_PARAM(TwoDMode, two_d_mode, bool)
  LSB_PARAM(EgoMotionLMMaxIter, ego_motion_lm_max_iter, unsigned int)
  LSB_PARAM(EgoMotionICPMaxIter, ego_motion_icp_max_iter, unsigned int)
  LSB_PARAM(EgoMotionMaxNeighborsDistance, ego_motion_max_neighbors_distance, double)
  LSB_PARAM(EgoMotionEdgeNbNeighbors, ego_motion_edge_nb_neighbors, unsigned int)
  LSB_PARAM(EgoMotionEdgeMinNbNeighbors, ego_motion_edge_min_nb_neighbors, unsigned int)
  LSB_PARAM(EgoMotionPlaneNbNeighbors, ego_motion_plane_nb_neighbors, unsigned int)
  LSB_PARAM(EgoMotionPlanarityThreshold, ego_motion_planarity_threshold, double)
  LSB_PARAM(EgoMotionEdgeMaxModelError, ego_motion_edge_max_model_error, double)
  LSB_PARAM(EgoMotionPlaneMaxModelError, ego_motion_plane_max_model_error, double)
  LSB_PARAM(EgoMotionInitSaturationDistance, ego_motion_init_saturation_distance, double)
  LSB_PARAM(EgoMotionFinalSaturationDistance, ego_motion_final_saturation_distance, double)
  LSB_PARAM(LocalizationLMMaxIter, localization_lm_max_iter, unsigned int)
  LSB_PARAM(LocalizationICPMaxIter, localization_icp_max_iter, unsigned int)
  LSB_PARAM(LocalizationMaxNeighborsDistance, localization_max_neighbors_distance, double)
  LSB_PARAM(LocalizationEdgeNbNeighbors, localization_edge_nb_neighbors, unsigned int)
  LSB_PARAM(LocalizationEdgeMinNbNeighbors, localization_edge_min_nb_neighbors, unsigned int)
  LSB_PARAM(LocalizationPlaneNbNeighbors, localization_plane_nb_neighbors, unsigned int)
  LSB_PARAM(LocalizationPlanarityThreshold, localization_planarity_threshold, double)
  LSB_PARAM(LocalizationEdgeMaxModelError, localization_edge_max_model_error, double)
  LSB_PARAM(LocalizationPlaneMaxModelError, localization_plane_max_model_error, double)
  LSB_PARAM(LocalizationBlobNbNeighbors, localization_blob_nb_neighbors, unsigned int)
  LSB_PARAM(LocalizationInitSaturationDistance, localization_init_saturation_distance, double)
  LSB_PARAM(LocalizationFinalSaturationDistance, localization_final_saturation_distance, double)

  // ---- external sensors (Slam.h:331-344) --------------------------------------------
  void SetWheelOdomWeight(double weight) { lsb_set_wheel_odom_weight(this->Gpu, weight); }
  double GetWheelOdomWeight() const { return lsb_get_wheel_odom_weight(this->Gpu); }
  void SetGravityWeight(double weight) { lsb_set_gravity_weight(this->Gpu, weight); }
  double GetGravityWeight() const { return lsb_get_gravity_weight(this->Gpu); }
  void SetSensorTimeOffset(double timeOffset) { lsb_set_sensor_time_offset(this->Gpu, timeOffset); }
  double GetSensorTimeOffset() const { return lsb_get_sensor_time_offset(this->Gpu); }
  void AddGravityMeasurement(const SensorConstraints::GravityMeasurement& gm)
  {
    lsb_gravity_measurement m = {gm.Time, {gm.Acceleration[0], gm.Acceleration[1], gm.Acceleration[2]}};
    lsb_add_gravity_measurement(this->Gpu, &m);
  }
  void AddWheelOdomMeasurement(const SensorConstraints::WheelOdomMeasurement& om)
  {
    lsb_wheel_odom_measurement m = {om.Time, om.Distance};
    lsb_add_wheel_odom_measurement(this->Gpu, &m);
  }
  void ClearSensorMeasurements() { lsb_clear_sensor_measurements(this->Gpu); }

  // ---- key frames and maps (Slam.h:350-372) ---------------------------------------
  LSB_PARAM(KfDistanceThreshold, kf_distance_threshold, double)
  LSB_PARAM(KfAngleThreshold, kf_angle_threshold, double)
  LSB_PARAM(MapUpdate, map_update, MappingMode)
  double GetVoxelGridDecayingThreshold() { return this->VoxelGridDecayingThreshold; }
  void SetVoxelGridDecayingThreshold(double decay) { this->VoxelGridDecayingThreshold = decay; lsb_set_voxel_grid_decaying_threshold(this->Gpu, decay); }
  SamplingMode GetVoxelGridSamplingMode(Keypoint k) { return this->VoxelGridSamplingMode[k]; }
  void SetVoxelGridSamplingMode(Keypoint k, SamplingMode sm) { this->VoxelGridSamplingMode[k] = sm; lsb_set_voxel_grid_sampling_mode(this->Gpu, k, (int)sm); }
  void ClearMaps() { lsb_clear_maps(this->Gpu); }
  void SetVoxelGridLeafSize(Keypoint k, double size) { lsb_set_voxel_grid_leaf_size(this->Gpu, k, size); }
  void SetVoxelGridSize(int size) { lsb_set_voxel_grid_size(this->Gpu, size); }
  void SetVoxelGridResolution(double resolution) { lsb_set_voxel_grid_resolution(this->Gpu, resolution); }
  void SetVoxelGridMinFramesPerVoxel(unsigned int minFrames) { lsb_set_voxel_grid_min_frames_per_voxel(this->Gpu, minFrames); }

  // ---- confidence estimation (Slam.h:378-394) ---------------------------------------
  LSB_PARAM(OverlapSamplingRatio, overlap_sampling_ratio, float)
  float GetOverlapEstimation() const { return lsb_get_overlap(this->Gpu); }
  int GetTotalMatchedKeypoints() const { return (int)lsb_get_total_matched_keypoints(this->Gpu); }
  std::array<float, 2> GetAccelerationLimits() const { return {this->P.acceleration_limits[0], this->P.acceleration_limits[1]}; }
  void SetAccelerationLimits(const std::array<float, 2>& a) { this->P.acceleration_limits[0] = a[0]; this->P.acceleration_limits[1] = a[1]; this->PushParams(); }
  std::array<float, 2> GetVelocityLimits() const { return {this->P.velocity_limits[0], this->P.velocity_limits[1]}; }
  void SetVelocityLimits(const std::array<float, 2>& v) { this->P.velocity_limits[0] = v[0]; this->P.velocity_limits[1] = v[1]; this->PushParams(); }
  LSB_PARAM(TimeWindowDuration, time_window_duration, float)
  bool GetComplyMotionLimits() const { return lsb_get_comply_motion_limits(this->Gpu) != 0; }
#undef LSB_PARAM

  // not part of the reference interface: the status of the last AddFrames (LSB_W_* flags, < 0 = LSB_E_*) and the handle
  int GetLastStatus() const { return this->LastStatus; }
  lsb_slam* GetHandle() const { return this->Gpu; }

private:
  lsb_slam* Gpu = nullptr;
  lsb_params P;
  int LastStatus = 0;
  unsigned int NbrFrameProcessed = 0;
  std::string BaseFrameId = "base", WorldFrameId = "world";                // Slam.h:456-457
  std::vector<PointCloud::Ptr> CurrentFrames;                                // kept like the reference (Slam.cxx:237)
  std::vector<std::vector<PointCloud::Ptr>> StagedFrames;                    // inputs handed to PrefetchFrames, alive until they are added
  std::map<std::uint8_t, KeypointExtractorPtr> KeyPointsExtractors;          // Slam.h:500
  std::map<std::uint8_t, Isometry3d> BaseToLidarOffsets;                     // Slam.h:505
  double VoxelGridDecayingThreshold = -1.;
  SamplingMode VoxelGridSamplingMode[3] = {SamplingMode::MAX_INTENSITY, SamplingMode::MAX_INTENSITY, SamplingMode::MAX_INTENSITY};

  void PushParams() { lsb_set_params(this->Gpu, &this->P); }
  // The wrappers set parameters directly on the extractor objects they installed (LidarSlamNode.cxx:778-817), so the blocks
  // are compared with what the library last saw before each frame and forwarded only when a setter changed something.
  std::map<std::uint8_t, lsb_ke_params> PushedExtractors;
  void PushExtractors()
  {
    for (const auto& kv : this->KeyPointsExtractors) {
      if (!kv.second) continue;
      auto it = this->PushedExtractors.find(kv.first);
      if (it != this->PushedExtractors.end() && std::memcmp(&it->second, &kv.second->P, sizeof(lsb_ke_params)) == 0) continue;
      lsb_set_extractor_params(this->Gpu, kv.first, &kv.second->P);
      this->PushedExtractors[kv.first] = kv.second->P;
    }
  }
  // The azimuthal resolution estimated from the first frame lives in the extractor object in the reference
  // (SpinningSensorKeypointExtractor.cxx:169) and is kept from then on: mirror it so that GetAzimuthalResolution() reports
  // it and a later parameter push carries it instead of resetting it.
  void PullExtractors()
  {
    for (const auto& kv : this->KeyPointsExtractors) {
      lsb_ke_params lib;
      if (!kv.second || lsb_get_extractor_params(this->Gpu, kv.first, &lib) != LSB_OK) continue;
      auto it = this->PushedExtractors.find(kv.first);
      if (it == this->PushedExtractors.end()) continue;
      if (std::memcmp(&it->second, &kv.second->P, sizeof(lsb_ke_params)) == 0) kv.second->P.azimuthal_resolution = lib.azimuthal_resolution;
      it->second.azimuthal_resolution = lib.azimuthal_resolution;
    }
  }
  template <typename F>
  PointCloud::Ptr Cloud(F&& get, const std::string& frameId) const
  {
    PointCloud::Ptr pc(new PointCloud);
    const long n = get(nullptr, 0);
    if (n > 0) {
      pc->resize((std::size_t)n);
      get(reinterpret_cast<lsb_point*>(pc->points.data()), n);
    }
    if (!this->CurrentFrames.empty() && this->CurrentFrames[0]) pc->header.stamp = this->CurrentFrames[0]->header.stamp;   // Utils::BuildPclHeader
    pc->header.frame_id = frameId;
    pc->header.seq = this->NbrFrameProcessed;
    return pc;
  }
};

#undef LSB_PRINT_WARNING
#undef LSB_PRINT_ERROR

}  // end of LidarSlam namespace

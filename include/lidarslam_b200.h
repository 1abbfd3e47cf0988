/* lidarslam_b200 — C ABI of the B200-native LidarSlam per-frame hot path.
 *
 * Drop-in boundary: a reference maintainer keeps LidarSlam::Slam's public C++
 * API (reference slam_lib/include/LidarSlam/Slam.h:98-394) and forwards its
 * bodies to these entry points (see INTEGRATION.md and
 * include/LidarSlam/Slam.h, the header-only C++ mirror of that class).  Every entry point cites the
 * reference interface it replaces.  Plain pointers and sizes only; all buffers
 * passed in and out are HOST memory unless the name says `_dev`.
 *
 * Return codes: 0 = ok, <0 = hard error (LSB_E_*), >0 = soft-skip flags that
 * mirror the reference's PRINT_WARNING / PRINT_ERROR paths (LSB_W_*).
 * No global state: one handle per SLAM sequence, handles are independent and
 * may be driven concurrently from different host threads.
 */
#ifndef LIDARSLAM_B200_H
#define LIDARSLAM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LSB_API __attribute__((visibility("default")))

/* ---- error / warning codes ------------------------------------------------ */
#define LSB_OK 0
#define LSB_E_CUDA (-1)        /* CUDA runtime error (see lsb_last_error) */
#define LSB_E_INVALID (-2)     /* invalid argument */
#define LSB_E_CAPACITY (-3)    /* a fixed device capacity was exceeded */
#define LSB_E_NO_DEVICE (-4)   /* no CUDA device: there is NO CPU fallback */
#define LSB_W_FRAME_IGNORED 1      /* Slam.cxx:720-731 empty input or duplicate stamp */
#define LSB_W_LOC_SKIPPED 2        /* Slam.cxx:1098-1107 not enough matches */
#define LSB_W_EGO_SKIPPED 4        /* Slam.cxx:919-923 */
#define LSB_W_EXTRAPOLATION 8      /* Slam.cxx:829-830 / 1278-1282 */
#define LSB_W_TIME_FIELD 16        /* Slam.cxx:1307-1318 */
#define LSB_W_FRAMES_DROPPED 32    /* Slam.cxx:734-740 */
#define LSB_W_TIME_WINDOW 64       /* Slam.cxx:1421-1432 velocity estimated on another window than required */
#define LSB_W_MOTION_LIMITS 128    /* Slam.cxx:1482-1483 pose does not comply with the motion limits */
#define LSB_W_SENSOR_TIME 256      /* SensorConstraints.cxx:19,75,121 no wheel-odometer / IMU measure around the frame stamp */

/* ---- data contracts -------------------------------------------------------- */

/* Bit-identical to LidarSlam::LidarPoint (LidarPoint.h:31-64): a
 * pcl::PointCloud<LidarPoint>::points.data() can be passed without a copy. */
typedef struct lsb_point {
  float x, y, z, w;      /* w = data[3] = 1.0f */
  double time;           /* seconds, offset to the frame header stamp */
  float intensity;
  uint16_t laser_id;     /* 0 = lowest ring, increasing upward (SSKE.h:84) */
  uint8_t device_id;
  uint8_t label;         /* in maps: 1 = fixed point (RollingGrid.cxx:218) */
} lsb_point;

/* Enums.h:30-39 */
enum { LSB_EDGE = 0, LSB_PLANE = 1, LSB_BLOB = 2, LSB_N_KEYPOINT_TYPES = 3 };
/* Enums.h:46-70 UndistortionMode */
enum { LSB_UNDISTORTION_NONE = 0, LSB_UNDISTORTION_ONCE = 1, LSB_UNDISTORTION_REFINED = 2 };
/* Enums.h:74-100 EgoMotionMode */
enum { LSB_EGO_NONE = 0, LSB_EGO_MOTION_EXTRAPOLATION = 1, LSB_EGO_REGISTRATION = 2, LSB_EGO_MOTION_EXTRAPOLATION_AND_REGISTRATION = 3 };
/* Enums.h:104-122 MappingMode */
enum { LSB_MAP_NONE = 0, LSB_MAP_ADD_KPTS_TO_FIXED_MAP = 1, LSB_MAP_UPDATE = 2 };
/* Enums.h:128-151 SamplingMode */
enum { LSB_SAMPLING_FIRST = 0, LSB_SAMPLING_LAST = 1, LSB_SAMPLING_MAX_INTENSITY = 2, LSB_SAMPLING_CENTER_POINT = 3, LSB_SAMPLING_CENTROID = 4 };
/* KeypointsMatcher.h:82-94 MatchStatus */
enum { LSB_MATCH_SUCCESS = 0, LSB_MATCH_BAD_MODEL_PARAMETRIZATION, LSB_MATCH_NOT_ENOUGH_NEIGHBORS, LSB_MATCH_NEIGHBORS_TOO_FAR,
       LSB_MATCH_BAD_PCA_STRUCTURE, LSB_MATCH_INVALID_NUMERICAL, LSB_MATCH_MSE_TOO_LARGE, LSB_MATCH_UNKNOWN, LSB_MATCH_N_STATUS };

/* POD mirror of SpinningSensorKeypointExtractor's parameters (SSKE.h:126-167). */
typedef struct lsb_ke_params {
  int32_t neighbor_width;              /* 4 */
  float min_distance_to_sensor;        /* 1.5 m */
  float min_beam_surface_angle;        /* 10 deg */
  float plane_sin_angle_threshold;     /* 0.5 */
  float edge_sin_angle_threshold;      /* 0.86 */
  float dist_to_line_threshold;        /* 0.20 m */
  float edge_depth_gap_threshold;      /* 0.15 m */
  float edge_saliency_threshold;       /* 1.5 m */
  float edge_intensity_gap_threshold;  /* 50 */
  float azimuthal_resolution;          /* rad; outside [1e-6, pi/4] => estimated from the next frame */
} lsb_ke_params;

/* POD mirror of KeypointsMatcher::Parameters (KeypointsMatcher.h:43-76). */
typedef struct lsb_match_params {
  int32_t single_edge_per_ring;
  double max_neighbors_distance;
  uint32_t edge_nb_neighbors, edge_min_nb_neighbors;
  double edge_max_model_error;
  uint32_t plane_nb_neighbors;
  double planarity_threshold, plane_max_model_error;
  uint32_t blob_nb_neighbors;
  double saturation_distance;
} lsb_match_params;

/* POD mirror of the plain Set/Get parameters of LidarSlam::Slam (Slam.h:403-694).
 * Voxel-grid parameters have setter semantics in the reference (re-binning,
 * resolution snapping) and therefore have their own entry points below. */
typedef struct lsb_params {
  int32_t nb_threads;                  /* kept for API parity; ignored on the GPU */
  int32_t use_keypoints[3];            /* {1,1,0} */
  int32_t ego_motion;                  /* LSB_EGO_MOTION_EXTRAPOLATION */
  int32_t undistortion;                /* LSB_UNDISTORTION_REFINED */
  int32_t verbosity;                   /* accepted and IGNORED (the reference prints phase timers, Slam.cxx:246-295): phase times are in lsb_frame_stats */
  double logging_timeout;              /* 0 */
  int32_t two_d_mode;                  /* 0 */
  uint32_t ego_motion_icp_max_iter;    /* 4 */
  uint32_t localization_icp_max_iter;  /* 3 */
  uint32_t ego_motion_lm_max_iter;     /* 15 */
  uint32_t localization_lm_max_iter;   /* 15 */
  double ego_motion_max_neighbors_distance;     /* 5 */
  double localization_max_neighbors_distance;   /* 5 */
  uint32_t ego_motion_edge_nb_neighbors;        /* 8 */
  uint32_t ego_motion_edge_min_nb_neighbors;    /* 3 */
  double ego_motion_edge_max_model_error;       /* 0.2 */
  uint32_t localization_edge_nb_neighbors;      /* 10 */
  uint32_t localization_edge_min_nb_neighbors;  /* 4 */
  double localization_edge_max_model_error;     /* 0.2 */
  uint32_t ego_motion_plane_nb_neighbors;       /* 5 */
  double ego_motion_planarity_threshold;        /* 0.04 */
  double ego_motion_plane_max_model_error;      /* 0.2 */
  uint32_t localization_plane_nb_neighbors;     /* 5 */
  double localization_planarity_threshold;      /* 0.04 */
  double localization_plane_max_model_error;    /* 0.2 */
  uint32_t localization_blob_nb_neighbors;      /* 10 */
  double ego_motion_init_saturation_distance;   /* 5 */
  double ego_motion_final_saturation_distance;  /* 1 */
  double localization_init_saturation_distance; /* 2 */
  double localization_final_saturation_distance;/* 0.5 */
  double kf_distance_threshold;        /* 0.5 m */
  double kf_angle_threshold;           /* 5 deg */
  int32_t map_update;                  /* LSB_MAP_UPDATE */
  double max_extrapolation_ratio;      /* 3 */
  uint32_t min_nb_matched_keypoints;   /* 20 */
  float overlap_sampling_ratio;        /* 0 (disabled) */
  /* Slam::CheckMotionLimits (Slam.cxx:1391-1484, Slam.h:685-694) */
  float time_window_duration;          /* 0 (disabled), seconds */
  float velocity_limits[2];            /* FLT_MAX: linear m/s, angular deg/s */
  float acceleration_limits[2];        /* FLT_MAX: linear m/s2, angular deg/s2 */
} lsb_params;

/* LocalOptimizer::RegistrationError (LocalOptimizer.h:34-49) */
typedef struct lsb_registration_error {
  double covariance[36];               /* column-major 6x6, order X Y Z rX rY rZ */
  double position_error;               /* m */
  double position_error_direction[3];
  double orientation_error;            /* degrees */
  double orientation_error_direction[3];
} lsb_registration_error;

/* Per-frame counters the bench turns into roofline numbers. */
typedef struct lsb_frame_stats {
  uint32_t n_points;                   /* input points of the frame(s) */
  uint32_t n_keypoints[3];
  uint32_t n_submap[3];
  uint32_t ego_icp_iters, loc_icp_iters;
  uint32_t ego_lm_iters, loc_lm_iters; /* LM iterations (cost/Jacobian evaluations - 1 per solve) */
  uint32_t n_matches[3];               /* localization, last ICP iteration */
  uint32_t ego_n_matches[3];
  uint32_t knn_queries;                /* kNN queries issued this frame */
  uint32_t is_keyframe;
  uint32_t n_map[3];
  uint32_t kernel_launches;            /* our kernels launched for this frame */
  float ms_extract, ms_ego, ms_localization, ms_map_update, ms_total; /* CUDA-event times */
  float ms_knn;                        /* CUDA-event time of the kNN kernel launches (needs lsb_set_profiling) */
  uint32_t knn_launches;               /* kNN kernel launches this frame */
  float ms_fit, ms_lm;                 /* model-fit kernels / LM evaluation kernels (needs lsb_set_profiling) */
  uint32_t lm_launches;
  uint64_t d2h_bytes, h2d_bytes;       /* bytes copied device->host / host->device inside the call */
  uint32_t map_update_deferred;        /* 1: the call returned with the pose while the map update went on (ms_total ends where it starts;
                                          ms_map_update and n_map are completed by lsb_get_frame_stats, which waits for it) */
  uint32_t prefetched;                 /* 1: the input had been staged by lsb_prefetch_frames (ms_extract ~ 0) */
  uint32_t launches_total;             /* our kernels launched by this handle since it was created (staged and deferred work included) */
} lsb_frame_stats;

typedef struct lsb_slam lsb_slam;      /* opaque: one CUDA stream + all device buffers of one sequence */

/* ---- lifetime -------------------------------------------------------------- */
LSB_API int lsb_device_count(void);
LSB_API const char* lsb_last_error(void);                       /* thread-local message of the last LSB_E_* */
LSB_API int lsb_create(lsb_slam** out, int cuda_device);        /* Slam::Slam()  Slam.cxx:143-161 */
LSB_API void lsb_destroy(lsb_slam* h);
LSB_API int lsb_reset(lsb_slam* h, int reset_log);              /* Slam::Reset   Slam.cxx:164-210 */

/* ---- parameters ------------------------------------------------------------ */
LSB_API void lsb_default_params(lsb_params* p);                 /* defaults of Slam.h:403-694 */
LSB_API void lsb_default_ke_params(lsb_ke_params* p);           /* defaults of SSKE.h:126-167 */
LSB_API int lsb_set_params(lsb_slam* h, const lsb_params* p);   /* the SetMacro setters Slam.h:201-394 */
LSB_API int lsb_get_params(const lsb_slam* h, lsb_params* p);
LSB_API int lsb_set_extractor_params(lsb_slam* h, int device_id, const lsb_ke_params* p); /* Slam::SetKeyPointsExtractor Slam.cxx:1619-1622 */
LSB_API int lsb_get_extractor_params(const lsb_slam* h, int device_id, lsb_ke_params* p);
LSB_API int lsb_set_base_to_lidar(lsb_slam* h, int device_id, const double T_rowmajor[16]); /* Slam::SetBaseToLidarOffset Slam.cxx:1629-1632 */
LSB_API int lsb_set_voxel_grid_leaf_size(lsb_slam* h, int keypoint_type, double size);      /* Slam.cxx:1671-1674 */
LSB_API int lsb_set_voxel_grid_size(lsb_slam* h, int size);                                 /* Slam.cxx:1677-1681, RollingGrid.cxx:59-70 */
LSB_API int lsb_set_voxel_grid_resolution(lsb_slam* h, double resolution);                  /* Slam.cxx:1684-1688, RollingGrid.cxx:73-88 */
LSB_API int lsb_set_voxel_grid_min_frames_per_voxel(lsb_slam* h, unsigned int n);           /* Slam.cxx:1691-1695 */
LSB_API int lsb_set_voxel_grid_decaying_threshold(lsb_slam* h, double seconds);             /* Slam.cxx:1652-1656 */
LSB_API int lsb_set_voxel_grid_sampling_mode(lsb_slam* h, int keypoint_type, int mode);     /* Slam.cxx:1665-1668 */

/* ---- the per-frame entry: Slam::AddFrames (Slam.cxx:230-344) ----------------
 * frames[i] points to n[i] HOST points of LiDAR device frames[i][0].device_id;
 * frames[0] is the main frame (its stamp is the pose time).  Synchronous:
 * returns after map update and logging, like the reference. */
LSB_API int lsb_add_frames(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n,
                           const uint64_t* stamp_us, const uint32_t* seq);

/* Same entry with the frames already resident in DEVICE memory of the handle's GPU
 * (frames[i] are device pointers; they must stay valid until the next call, as the
 * reference retains its input clouds, Slam.cxx:237).  Used to time the path with
 * inputs in HBM. */
LSB_API int lsb_add_frames_dev(lsb_slam* h, int n_frames, const lsb_point* const* dev_frames, const size_t* n,
                               const uint64_t* stamp_us, const uint32_t* seq);

/* ---- pipelining (no counterpart in the reference: Slam::AddFrames is synchronous, Slam.cxx:230-344) ----------------------------
 * A trajectory is sequential, but two pieces of a frame do not depend on its neighbours' results and are taken off the caller's
 * critical path:
 *  - the map update of frame t (Slam::UpdateMapsUsingTworld, Slam.cxx:1178-1222) is deferred: lsb_add_frames returns once the
 *    pose is known, the update runs behind it and is joined before the next frame's sub-maps are selected or by the first entry
 *    point that looks at the maps / keypoints / statistics.  Results are identical; lsb_set_pipelining(h, 0) restores the
 *    fully synchronous call.
 *  - lsb_prefetch_frames hands the NEXT input in ahead of time (a bag replay, a driver queue): its upload and keypoint extraction
 *    (Slam::ExtractKeypoints, Slam.cxx:746-810, which reads nothing but the frame and the extractor parameters) run on an ingest
 *    stream while the current frame is in its ICP loop.  lsb_add_frames / lsb_add_frames_dev called later with the same buffer
 *    pointers, sizes and stamps picks the staged keypoints up; any other input drops what is staged.  At most three inputs are
 *    staged at a time (LSB_E_CAPACITY beyond).  The buffers must stay valid and unchanged until the matching add call returns.
 *    Parameter setters drop staged inputs (they were extracted with the old parameters). */
LSB_API int lsb_prefetch_frames(lsb_slam* h, int n_frames, const lsb_point* const* frames, const size_t* n,
                                const uint64_t* stamp_us, const uint32_t* seq);
LSB_API int lsb_prefetch_frames_dev(lsb_slam* h, int n_frames, const lsb_point* const* dev_frames, const size_t* n,
                                    const uint64_t* stamp_us, const uint32_t* seq);
LSB_API int lsb_set_pipelining(lsb_slam* h, int defer_map_update);
LSB_API int lsb_synchronize(lsb_slam* h);            /* every deferred / staged job over, every stream of the handle idle */
/* Whole-loop timing with CUDA events on the handle's stream: begin = lsb_synchronize + event; end = deferred and staged work
 * joined into the stream + event + wait; *ms = time between the two (benchmarks: K calls bracketed as one region). */
LSB_API int lsb_timer_begin(lsb_slam* h);
LSB_API int lsb_timer_end(lsb_slam* h, float* ms);

/* Memory helpers for callers without their own CUDA allocator: pinned host
 * buffers (fast H2D for lsb_add_frames) and device buffers (lsb_add_frames_dev). */
LSB_API void* lsb_host_alloc(size_t bytes);
LSB_API void lsb_host_free(void* p);
LSB_API void* lsb_dev_alloc(lsb_slam* h, size_t bytes);
LSB_API void lsb_dev_free(lsb_slam* h, void* p);
LSB_API int lsb_dev_upload(lsb_slam* h, void* dev_dst, const void* host_src, size_t bytes);
LSB_API int lsb_dev_flush_l2(lsb_slam* h);   /* writes a 256 MiB scratch buffer (benchmark hygiene) */

/* ---- outputs ---------------------------------------------------------------- */
LSB_API int lsb_get_world_transform(const lsb_slam* h, double T_rowmajor[16], double* time);   /* Slam::GetWorldTransform Slam.cxx:550-553 */
LSB_API int lsb_get_latency_compensated_world_transform(const lsb_slam* h, double T_rowmajor[16], double* time); /* Slam.cxx:556-591 */
LSB_API int lsb_get_relative_transform(const lsb_slam* h, double T_rowmajor[16]);              /* Trelative (Slam.h:461) */
LSB_API int lsb_get_covariance(const lsb_slam* h, double cov36[36]);                           /* Slam::GetTransformCovariance Slam.cxx:594-597 */
LSB_API int lsb_get_registration_error(const lsb_slam* h, lsb_registration_error* e);          /* LocalOptimizer.cxx:112-140 */
LSB_API long lsb_get_trajectory(const lsb_slam* h, double* T_rowmajor16_each, double* times, long cap); /* Slam::GetTrajectory */
/* clouds: return the count; copy at most cap points when out != NULL */
LSB_API long lsb_get_keypoints(lsb_slam* h, int type, int world, lsb_point* out, long cap);    /* Slam::GetKeypoints Slam.cxx:690-702 */
LSB_API long lsb_get_map(lsb_slam* h, int type, int clean, lsb_point* out, long cap);          /* Slam::GetMap Slam.cxx:670-677 */
LSB_API long lsb_get_submap(lsb_slam* h, int type, lsb_point* out, long cap);                  /* Slam::GetTargetSubMap Slam.cxx:680-687 */
LSB_API long lsb_get_registered_frame(lsb_slam* h, lsb_point* out, long cap);                  /* Slam::GetRegisteredFrame Slam.cxx:660-667 */
/* Slam::GetDebugArray (Slam.cxx:635-657): stage 0 = EgoMotion, 1 = Localization */
LSB_API long lsb_get_match_debug(lsb_slam* h, int stage, int type, uint8_t* status, double* weight, long cap);
/* SpinningSensorKeypointExtractor::GetDebugArray (SSKE.cxx:640-680), input order.
 * field: 0 sin_angle 1 saliency 2 depth_gap 3 intensity_gap 4-6 {edge,plane,blob}_keypoint 7-9 {edge,plane,blob}_validity */
LSB_API long lsb_get_ke_debug(lsb_slam* h, int device_id, int field, float* out, long cap);
LSB_API int lsb_get_frame_stats(const lsb_slam* h, lsb_frame_stats* s);
/* Measurement aid (no reference counterpart): arithmetic peaks of the device from a dependent-chain microbenchmark, in
 * TFLOP/s: [0] FP64 FMA, [1] FP64 multiply + add not fused (what this library issues: it is built without FMA contraction),
 * [2] FP32 FMA, [3] FP32 multiply + add not fused.  bench.py divides by these instead of nominal figures. */
LSB_API int lsb_measure_fp_peaks(int cuda_device, double out_tflops[4]);
/* When on, CUDA events bracket every kNN / fit / LM kernel launch and
 * lsb_frame_stats.ms_knn / ms_fit / ms_lm are filled; the three keypoint types, which
 * otherwise run on their own streams, are serialised on the handle's stream so that the
 * durations are those of the kernels alone (slower frames: not for throughput runs). */
LSB_API int lsb_set_profiling(lsb_slam* h, int on);
/* how many times the extractor of device_id estimated its azimuthal resolution from a frame (SSKE.cxx:165-170, 593-637):
 * 1 after the first frame unless the caller set a usable value (then 0) or reset it */
LSB_API unsigned int lsb_get_azimuth_estimations(const lsb_slam* h, int device_id);
LSB_API float lsb_get_overlap(const lsb_slam* h);                                              /* Slam::GetOverlapEstimation */
LSB_API int lsb_get_comply_motion_limits(const lsb_slam* h);                                   /* Slam::GetComplyMotionLimits (Slam.h:394) */
LSB_API unsigned int lsb_get_total_matched_keypoints(const lsb_slam* h);                       /* Slam.h TotalMatchedKeypoints */

/* ---- state control ---------------------------------------------------------- */
LSB_API int lsb_set_world_transform_from_guess(lsb_slam* h, const double T_rowmajor[16]);      /* Slam.cxx:490-501 */
LSB_API int lsb_clear_maps(lsb_slam* h);                                                        /* Slam.cxx:1639-1643 */
LSB_API int lsb_map_load_points(lsb_slam* h, int type, const lsb_point* pts, long n, int fixed); /* body of Slam::LoadMapsFromPCD Slam.cxx:520-543 */

/* ---- external sensor constraints of the localization solve ------------------ */
typedef struct lsb_wheel_odom_measurement { double time, distance; } lsb_wheel_odom_measurement;            /* SensorConstraints.h:13-17 */
typedef struct lsb_gravity_measurement { double time; double acceleration[3]; } lsb_gravity_measurement;    /* SensorConstraints.h:19-23 */
LSB_API int lsb_add_wheel_odom_measurement(lsb_slam* h, const lsb_wheel_odom_measurement* m);   /* Slam::AddWheelOdomMeasurement Slam.cxx:1589-1592 */
LSB_API int lsb_add_gravity_measurement(lsb_slam* h, const lsb_gravity_measurement* m);         /* Slam::AddGravityMeasurement Slam.cxx:1584-1587 */
LSB_API int lsb_clear_sensor_measurements(lsb_slam* h);                                          /* Slam.cxx:1594-1598 */
LSB_API int lsb_set_wheel_odom_weight(lsb_slam* h, double weight);                               /* Slam.h:331 */
LSB_API double lsb_get_wheel_odom_weight(const lsb_slam* h);                                     /* Slam.h:332 */
LSB_API int lsb_set_gravity_weight(lsb_slam* h, double weight);                                  /* Slam.h:334 */
LSB_API double lsb_get_gravity_weight(const lsb_slam* h);                                        /* Slam.h:335 */
LSB_API int lsb_set_sensor_time_offset(lsb_slam* h, double seconds);                             /* Slam.cxx:223-227 */
LSB_API double lsb_get_sensor_time_offset(const lsb_slam* h);                                    /* Slam.h:339 */
LSB_API int lsb_get_gravity_reference(const lsb_slam* h, double g[3]);                           /* ImuManager::GetGravityRef SensorConstraints.h:103 */

/* ---- input adaptors on the device: raw driver clouds -> LidarPoint frames -------- */
/* velodyne_pcl::PointXYZIRT (VelodyneToLidarNode.h:37), 32 B */
typedef struct lsb_velodyne_point { float x, y, z, _pad0; float intensity; uint16_t ring; uint16_t _pad1; float time; float _pad2; } lsb_velodyne_point;
/* pcl::PointXYZI (RobosenseToLidarNode.h:40), 32 B */
typedef struct lsb_xyzi_point { float x, y, z, _pad0; float intensity; float _pad1[3]; } lsb_xyzi_point;
/* node parameters (VelodyneToLidarNode.h:63-75, RobosenseToLidarNode.h:62-77) */
typedef struct lsb_conversion_params {
  const int32_t* laser_id_mapping;   /* NULL / 0: identity (Robosense with 16 rings: the RS16 table) */
  int32_t n_laser_id_mapping;
  int32_t device_id;                 /* 0 */
  double rpm;                        /* 600 */
  int32_t timestamp_first_packet;    /* 0: header.stamp is the last packet */
} lsb_conversion_params;
/* VelodyneToLidarNode::Callback (VelodyneToLidarNode.cxx:50-111): laser-id remap, device id, and - when the time field is
 * not set (last - first <= 1e-8) - point times from the azimuth advancement (Utilities.h:58-108).  The converted frame
 * stays on the device for lsb_add_converted_frame; `out` (host, may be NULL) receives at most `cap` points.
 * Returns the number of points, LSB_W_FRAME_IGNORED-like 0 for an empty cloud. */
LSB_API long lsb_convert_velodyne(lsb_slam* h, const lsb_velodyne_point* pts, size_t n, const lsb_conversion_params* p, lsb_point* out, long cap);
/* RobosenseToLidarNode::Callback (RobosenseToLidarNode.cxx:58-137): organised cloud (width x height, ring-major); drops
 * non-finite points and dual returns identical to the previous kept point; laser id from the row; time from the column. */
LSB_API long lsb_convert_robosense(lsb_slam* h, const lsb_xyzi_point* pts, uint32_t width, uint32_t height, const lsb_conversion_params* p,
                                   lsb_point* out, long cap);
/* Slam::AddFrame on the frame left on the device by the last lsb_convert_* call (no host round trip). */
LSB_API int lsb_add_converted_frame(lsb_slam* h, uint64_t stamp_us, uint32_t seq);

/* ---- map persistence: PCD wire format (pcl::io::savePCDFile* / loadPCDFile restated, no PCL) ---- */
enum { LSB_PCD_ASCII = 0, LSB_PCD_BINARY = 1, LSB_PCD_BINARY_COMPRESSED = 2 };                    /* PointCloudStorage.h:60-65 PCDFormat */
/* Slam::SaveMapsToPCD (Slam.cxx:504-517): writes <prefix>{edges,planes,blobs}.pcd for the keypoint types in use;
 * an empty map writes nothing.  LSB_E_INVALID if a file cannot be written. */
LSB_API int lsb_save_maps_to_pcd(lsb_slam* h, const char* file_prefix, int pcd_format, int filtered);
/* Slam::LoadMapsFromPCD (Slam.cxx:520-543): a missing / unreadable file is skipped; points are fixed when map_update is
 * NONE or ADD_KPTS_TO_FIXED_MAP; their time is the wall clock.  Returns a bit mask of the types loaded (1 << type). */
LSB_API int lsb_load_maps_from_pcd(lsb_slam* h, const char* file_prefix, int reset_maps);
/* savePointCloudToPCD (PointCloudStorage.h:85-115): 0 ok, -1 I/O error, -3 empty cloud, -4 unknown format. */
LSB_API int lsb_pcd_write(const char* path, const lsb_point* pts, size_t n, int pcd_format);
/* pcl::io::loadPCDFile<LidarPoint> (Slam.cxx:531): returns the number of points in the file (-1: unreadable) and copies
 * at most `cap` of them; call with cap = 0 to size the buffer.  Needs no device. */
LSB_API long lsb_pcd_read(const char* path, lsb_point* out, long cap);

/* ---- kernel-level entry points (parity tests and benchmarks) -----------------
 * Each runs ONE stage of the path on caller-supplied host data through the same
 * kernels lsb_add_frames uses. */

/* SpinningSensorKeypointExtractor::ComputeKeyPoints (SSKE.cxx:118-136) on one frame.
 * label_out[i] (input order): bit0 EDGE, bit1 PLANE, bit2 BLOB.  counts[3] = keypoints per type.
 * Keypoints are then readable with lsb_get_extracted_keypoints. */
/* Stage entry for tests: the stable single-launch radix sort of (key, value) pairs the map update orders a cloud's leaf keys
 * with (csrc/cuda/sort32.cu; stands where the reference walks its nested unordered_maps, RollingGrid.cxx:188-313).  Host
 * arrays, sorted in place by the low key_bits bits of the key; n <= 262144. */
LSB_API int lsb_sort_pairs32(lsb_slam* h, uint32_t* keys, uint32_t* values, size_t n, int key_bits);
LSB_API int lsb_extract_keypoints(lsb_slam* h, const lsb_point* pts, size_t n, uint8_t* label_out, uint32_t counts[3]);
LSB_API long lsb_get_extracted_keypoints(lsb_slam* h, int type, lsb_point* out, long cap);

/* KDTreePCLAdaptor::Reset + KnnSearch (KDTreePCLAdaptor.h:57-105): exact kNN of
 * nq float queries in a cloud of m points.  idx_out/d2_out are [nq*k]; entries
 * beyond the number found (m < k) are -1 / +inf.  Order: ascending (d2, index). */
LSB_API int lsb_knn_search(lsb_slam* h, const lsb_point* cloud, size_t m, const float* queries_xyz, size_t nq, int k,
                           int32_t* idx_out, float* d2_out);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  Same result as lsb_knn_search
 * (exact, ascending (d2, index), -1 / +inf beyond the number found), but ALLOCATES NOTHING: cloud, queries and outputs
 * are DEVICE pointers on `cuda_device` supplied by the caller, the work is enqueued on `cuda_stream` (a cudaStream_t;
 * NULL = the default stream) and NOT synchronised, and the kernel's scratch (per-segment partial lists + tickets) is a
 * file-scope static __device__ array of 4 MiB + 8 KiB that every GPU loading the library holds until the process exits.
 * Limits (beyond them: LSB_E_CAPACITY with a message in lsb_last_error): nq <= LSB_NOALLOC_MAX_QUERIES, k <= LSB_NOALLOC_MAX_K.
 * All calls on a device share that array: calls that overlap on the GPU (different streams without a dependency) RACE.
 * The first call on a device looks the array's address up (cudaGetSymbolAddress): make it before starting a capture. */
#define LSB_NOALLOC_MAX_QUERIES 2048
#define LSB_NOALLOC_MAX_K 16
LSB_API int lsb_knn_search_noalloc(int cuda_device, const lsb_point* d_cloud, size_t m, const float* d_queries_xyz, size_t nq, int k,
                                   int32_t* d_idx_out, float* d_d2_out, void* cuda_stream);

/* KeypointsMatcher::BuildMatchResiduals x3 + LocalOptimizer::Solve
 * (KeypointsMatcher.cxx:33-74, LocalOptimizer.cxx:74-102) for ONE ICP iteration:
 * keypoints[t] (BASE frame) matched against target[t] with the pose prior, then LM.
 * Outputs: status/weight per keypoint (index-aligned, types concatenated EDGE|PLANE|BLOB),
 * optimised pose, number of accepted LM steps, and the final J^T J (21 upper + 6 + cost). */
LSB_API int lsb_match_and_solve(lsb_slam* h, const lsb_point* const keypoints[3], const size_t nk[3],
                                const lsb_point* const target[3], const size_t nt[3],
                                const lsb_match_params* mp, int lm_max_iter, const double prior_rowmajor[16],
                                uint8_t* status_out, double* weight_out, double pose_out_rowmajor[16],
                                int* lm_accepted_steps, int* lm_iterations, double* final_cost,
                                lsb_registration_error* err_out);

/* RollingGrid::Add (RollingGrid.cxx:160-318) on map `type`, then the map is
 * readable with lsb_get_map. */
LSB_API int lsb_map_add(lsb_slam* h, int type, const lsb_point* pts, size_t n, int fixed, double current_time);

/* Undistortion / rigid transform (Slam.cxx:1337-1351, Utilities.h:274-278):
 * applies interp(H0,H1,t0,t1)(p.time) to every point in place. */
LSB_API int lsb_undistort(lsb_slam* h, lsb_point* pts, size_t n, const double H0_rowmajor[16], const double H1_rowmajor[16],
                          double t0, double t1);

#ifdef __cplusplus
}
#endif
#endif /* LIDARSLAM_B200_H */

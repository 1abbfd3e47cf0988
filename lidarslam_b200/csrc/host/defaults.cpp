// Default parameter values of the reference, shared by the product library and
// the oracle (they are interface constants, not algorithms).
//   lsb_default_params     <- slam_lib/include/LidarSlam/Slam.h:403-694
//   lsb_default_ke_params  <- slam_lib/include/LidarSlam/SpinningSensorKeypointExtractor.h:126-167
#include <cfloat>
#include <cstring>
#include "../../../include/lidarslam_b200.h"

extern "C" {

void lsb_default_params(lsb_params* p)
{
  std::memset(p, 0, sizeof(*p));
  p->nb_threads = 1;
  p->use_keypoints[LSB_EDGE] = 1; p->use_keypoints[LSB_PLANE] = 1; p->use_keypoints[LSB_BLOB] = 0;
  p->ego_motion = LSB_EGO_MOTION_EXTRAPOLATION;
  p->undistortion = LSB_UNDISTORTION_REFINED;
  p->verbosity = 0;
  p->logging_timeout = 0.;
  p->two_d_mode = 0;
  p->ego_motion_icp_max_iter = 4;
  p->localization_icp_max_iter = 3;
  p->ego_motion_lm_max_iter = 15;
  p->localization_lm_max_iter = 15;
  p->ego_motion_max_neighbors_distance = 5.;
  p->localization_max_neighbors_distance = 5.;
  p->ego_motion_edge_nb_neighbors = 8;
  p->ego_motion_edge_min_nb_neighbors = 3;
  p->ego_motion_edge_max_model_error = 0.2;
  p->localization_edge_nb_neighbors = 10;
  p->localization_edge_min_nb_neighbors = 4;
  p->localization_edge_max_model_error = 0.2;
  p->ego_motion_plane_nb_neighbors = 5;
  p->ego_motion_planarity_threshold = 0.04;
  p->ego_motion_plane_max_model_error = 0.2;
  p->localization_plane_nb_neighbors = 5;
  p->localization_planarity_threshold = 0.04;
  p->localization_plane_max_model_error = 0.2;
  p->localization_blob_nb_neighbors = 10;
  p->ego_motion_init_saturation_distance = 5.;
  p->ego_motion_final_saturation_distance = 1.;
  p->localization_init_saturation_distance = 2.;
  p->localization_final_saturation_distance = 0.5;
  p->kf_distance_threshold = 0.5;
  p->kf_angle_threshold = 5.;
  p->map_update = LSB_MAP_UPDATE;
  p->max_extrapolation_ratio = 3.;
  p->min_nb_matched_keypoints = 20;
  p->overlap_sampling_ratio = 0.f;
  p->time_window_duration = 0.f;                                   // Slam.h:694
  p->velocity_limits[0] = p->velocity_limits[1] = FLT_MAX;         // Slam.h:685
  p->acceleration_limits[0] = p->acceleration_limits[1] = FLT_MAX; // Slam.h:687
}

void lsb_default_ke_params(lsb_ke_params* p)
{
  p->neighbor_width = 4;
  p->min_distance_to_sensor = 1.5f;
  p->min_beam_surface_angle = 10.f;
  p->plane_sin_angle_threshold = 0.5f;
  p->edge_sin_angle_threshold = 0.86f;
  p->dist_to_line_threshold = 0.20f;
  p->edge_depth_gap_threshold = 0.15f;
  p->edge_saliency_threshold = 1.5f;
  p->edge_intensity_gap_threshold = 50.f;
  p->azimuthal_resolution = 0.f;
}

}  // extern "C"

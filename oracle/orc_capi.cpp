// CPU ORACLE (test infrastructure) — C entry points for ctypes, mirroring the
// lsb_* C ABI (include/lidarslam_b200.h) one to one with an orc_ prefix so the
// parity tests call both sides the same way.
#include "orc_slam.h"
#include <cstring>
#include <omp.h>

using namespace orc;

static Iso3d FromRowMajor(const double* T)
{
  Iso3d r;
  for (int i = 0; i < 3; ++i) {
    for (int j = 0; j < 3; ++j) r.R(i, j) = T[i * 4 + j];
  }
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
static long CopyCloud(const Cloud& c, lsb_point* out, long cap)
{
  if (out) std::memcpy(out, c.data(), sizeof(lsb_point) * (size_t)std::min<long>(cap, (long)c.size()));
  return (long)c.size();
}

extern "C" {

#define ORC_API __attribute__((visibility("default")))

ORC_API int orc_max_threads() { return omp_get_max_threads(); }
ORC_API Slam* orc_create() { return new Slam(); }
ORC_API void orc_destroy(Slam* h) { delete h; }
ORC_API int orc_reset(Slam* h, int resetLog) { h->Reset(resetLog != 0); return 0; }
ORC_API int orc_set_params(Slam* h, const lsb_params* p) { h->P = *p; return 0; }
ORC_API int orc_get_params(const Slam* h, lsb_params* p) { *p = h->P; return 0; }
ORC_API int orc_set_extractor_params(Slam* h, int dev, const lsb_ke_params* p)
{
  // same rule as lsb_set_extractor_params: an out-of-range incoming resolution keeps a usable estimate (SSKE.cxx:165-170)
  lsb_ke_params& P = h->KeyPointsExtractors[dev].P;
  const float est = P.azimuthal_resolution;
  P = *p;
  const bool unset = !(p->azimuthal_resolution >= 1e-6f && p->azimuthal_resolution <= 0.78539816339744830962f);
  if (unset && est >= 1e-6f && est <= 0.78539816339744830962f) P.azimuthal_resolution = est;
  return 0;
}
ORC_API int orc_get_extractor_params(Slam* h, int dev, lsb_ke_params* p)
{
  if (!h->KeyPointsExtractors.count(dev)) return LSB_E_INVALID;
  *p = h->KeyPointsExtractors[dev].P;
  return 0;
}
ORC_API int orc_set_base_to_lidar(Slam* h, int dev, const double* T) { h->BaseToLidarOffsets[dev] = FromRowMajor(T); return 0; }
ORC_API int orc_set_voxel_grid_leaf_size(Slam* h, int k, double s) { h->SetVoxelGridLeafSize(k, s); return 0; }
ORC_API int orc_set_voxel_grid_size(Slam* h, int s) { h->SetVoxelGridSize(s); return 0; }
ORC_API int orc_set_voxel_grid_resolution(Slam* h, double r) { h->SetVoxelGridResolution(r); return 0; }
ORC_API int orc_set_voxel_grid_min_frames_per_voxel(Slam* h, unsigned n) { for (auto& m : h->LocalMaps) m.MinFramesPerVoxel = n; return 0; }
ORC_API int orc_set_voxel_grid_decaying_threshold(Slam* h, double s) { for (auto& m : h->LocalMaps) m.DecayingThreshold = s; return 0; }
ORC_API int orc_set_voxel_grid_sampling_mode(Slam* h, int k, int mode) { h->LocalMaps[k].Sampling = mode; return 0; }

ORC_API int orc_add_frames(Slam* h, int n_frames, const lsb_point* const* frames, const size_t* n, const uint64_t* stamp_us, const uint32_t* seq)
{
  std::vector<Frame> f;
  for (int i = 0; i < n_frames; ++i) f.push_back(Frame{frames[i], n[i], stamp_us[i], seq ? seq[i] : 0u});
  return h->AddFrames(f);
}

ORC_API int orc_get_world_transform(const Slam* h, double* T, double* time)
{
  if (h->LogTrajectory.empty()) { ToRowMajor(Iso3d::identity(), T); if (time) *time = 0.; return 0; }
  ToRowMajor(h->LogTrajectory.back().T, T);
  if (time) *time = h->LogTrajectory.back().time;
  return 0;
}
ORC_API long orc_get_trajectory(const Slam* h, double* Ts, double* times, long cap)   // Slam::GetTrajectory
{
  const long n = (long)h->LogTrajectory.size();
  for (long i = 0; i < n && i < cap; ++i) {
    if (Ts) ToRowMajor(h->LogTrajectory[i].T, Ts + 16 * i);
    if (times) times[i] = h->LogTrajectory[i].time;
  }
  return n;
}
ORC_API int orc_get_relative_transform(const Slam* h, double* T) { ToRowMajor(h->Trelative, T); return 0; }
ORC_API int orc_get_registration_error(const Slam* h, lsb_registration_error* e) { *e = h->LocalizationUncertainty; return 0; }
ORC_API int orc_get_covariance(const Slam* h, double* c) { std::memcpy(c, h->LocalizationUncertainty.covariance, 36 * sizeof(double)); return 0; }
ORC_API long orc_get_keypoints(Slam* h, int type, int world, lsb_point* out, long cap)
{
  if (!world) return CopyCloud(h->CurrentKeypoints[type], out, cap);
  Cloud c = h->CurrentKeypoints[type];
  for (auto& p : c) lsb::transform_point_f(h->Tworld, p.x, p.y, p.z);
  return CopyCloud(c, out, cap);
}
ORC_API long orc_get_map(Slam* h, int type, int clean, lsb_point* out, long cap) { return CopyCloud(h->LocalMaps[type].Get(clean != 0), out, cap); }
ORC_API long orc_get_submap(Slam* h, int type, lsb_point* out, long cap) { return CopyCloud(h->LocalMaps[type].SubMap, out, cap); }
ORC_API long orc_get_registered_frame(Slam* h, lsb_point* out, long cap) { return CopyCloud(h->GetRegisteredFrame(), out, cap); }
ORC_API long orc_get_match_debug(Slam* h, int stage, int type, uint8_t* status, double* weight, long cap)
{
  const MatchingResults& m = stage == 0 ? h->EgoMotionMatchingResults[type] : h->LocalizationMatchingResults[type];
  const long n = (long)m.Rejections.size();
  for (long i = 0; i < std::min(n, cap); ++i) { if (status) status[i] = m.Rejections[i]; if (weight) weight[i] = m.Weights[i]; }
  return n;
}
ORC_API long orc_get_ke_debug(Slam* h, int dev, int field, float* out, long cap)
{
  if (!h->KeyPointsExtractors.count(dev)) return LSB_E_INVALID;
  const std::vector<float> v = h->KeyPointsExtractors[dev].DebugArray(field);
  if (out) std::memcpy(out, v.data(), sizeof(float) * (size_t)std::min<long>(cap, (long)v.size()));
  return (long)v.size();
}
ORC_API int orc_get_frame_stats(const Slam* h, lsb_frame_stats* s) { *s = h->Stats; return 0; }
ORC_API unsigned orc_get_total_matched_keypoints(const Slam* h) { return h->TotalMatchedKeypoints; }
ORC_API float orc_get_overlap(const Slam* h) { return h->OverlapEstimation; }
// external sensor constraints (Slam.cxx:1584-1598, Slam.h:331-339)
ORC_API int orc_add_wheel_odom_measurement(Slam* h, const lsb_wheel_odom_measurement* m)
{
  lsb::WheelOdomMeasurement w;
  w.Time = m->time; w.Distance = m->distance;
  h->WheelOdomManager.AddMeasurement(w);
  return 0;
}
ORC_API int orc_add_gravity_measurement(Slam* h, const lsb_gravity_measurement* m)
{
  lsb::GravityMeasurement g;
  g.Time = m->time;
  for (int i = 0; i < 3; ++i) g.Acceleration[i] = m->acceleration[i];
  h->ImuManager.AddMeasurement(g);
  return 0;
}
ORC_API int orc_clear_sensor_measurements(Slam* h) { h->WheelOdomManager.Reset(); h->ImuManager.Reset(); return 0; }
ORC_API int orc_set_wheel_odom_weight(Slam* h, double w) { h->WheelOdomManager.Weight = w; return 0; }
ORC_API int orc_set_gravity_weight(Slam* h, double w) { h->ImuManager.Weight = w; return 0; }
ORC_API int orc_set_sensor_time_offset(Slam* h, double t) { h->WheelOdomManager.TimeOffset = t; h->ImuManager.TimeOffset = t; return 0; }
ORC_API int orc_get_gravity_reference(const Slam* h, double* g) { for (int i = 0; i < 3; ++i) g[i] = h->ImuManager.GravityRef[i]; return 0; }
ORC_API int orc_get_comply_motion_limits(const Slam* h) { return h->MotionLimits.Comply ? 1 : 0; }
ORC_API int orc_set_world_transform_from_guess(Slam* h, const double* T) { h->SetWorldTransformFromGuess(FromRowMajor(T)); return 0; }
ORC_API int orc_clear_maps(Slam* h) { h->ClearMaps(); return 0; }
ORC_API int orc_map_load_points(Slam* h, int type, const lsb_point* pts, long n, int fixed)
{
  Cloud c(pts, pts + n);
  h->LocalMaps[type].Add(c, fixed != 0, 0.);
  return 0;
}

// ---- kernel-level entry points ------------------------------------------------
ORC_API int orc_extract_keypoints(Slam* h, const lsb_point* pts, size_t n, uint8_t* label_out, uint32_t* counts)
{
  Extractor& ke = h->KeyPointsExtractors[0];
  ke.NbThreads = h->P.nb_threads;
  ke.ComputeKeyPoints(pts, n);
  const std::vector<uint8_t> l = ke.LabelsInputOrder();
  if (label_out) std::memcpy(label_out, l.data(), n);
  for (int k = 0; k < 3; ++k) counts[k] = (uint32_t)ke.Keypoints[k].size();
  return 0;
}
ORC_API long orc_get_extracted_keypoints(Slam* h, int type, lsb_point* out, long cap) { return CopyCloud(h->KeyPointsExtractors[0].Keypoints[type], out, cap); }

ORC_API int orc_knn_search(Slam* h, const lsb_point* cloud, size_t m, const float* q, size_t nq, int k, int32_t* idx_out, float* d2_out)
{
  Cloud c(cloud, cloud + m);
  KdTree t;
  t.Reset(&c);
#pragma omp parallel for num_threads(h->P.nb_threads) schedule(guided, 8)
  for (long i = 0; i < (long)nq; ++i) {
    int* id = idx_out + i * k;
    float* d = d2_out + i * k;
    const int n = t.Knn(q + 3 * i, k, id, d);
    for (int j = n; j < k; ++j) { id[j] = -1; d[j] = std::numeric_limits<float>::infinity(); }
  }
  return 0;
}

ORC_API int orc_match_and_solve(Slam* h, const lsb_point* const* keypoints, const size_t* nk, const lsb_point* const* target, const size_t* nt,
                                const lsb_match_params* mp, int lm_max_iter, const double* prior, uint8_t* status_out, double* weight_out,
                                double* pose_out, int* lm_accepted_steps, int* lm_iterations, double* final_cost, lsb_registration_error* err_out)
{
  const Iso3d T = FromRowMajor(prior);
  Matcher matcher(*mp, T, h->P.nb_threads);
  Optimizer opt;
  opt.TwoDMode = h->P.two_d_mode != 0;
  opt.SetPosePrior(T);
  opt.LMMaxIter = (unsigned)lm_max_iter;
  opt.NbThreads = h->P.nb_threads;
  opt.Sat = mp->saturation_distance;
  size_t off = 0;
  unsigned total = 0;
  for (int k = 0; k < 3; ++k) {
    Cloud kp(keypoints[k], keypoints[k] + nk[k]), tg(target[k], target[k] + nt[k]);
    KdTree tree;
    tree.Reset(&tg);
    MatchingResults r = matcher.BuildMatchResiduals(kp, tree, k);
    for (size_t i = 0; i < nk[k]; ++i) { status_out[off + i] = r.Rejections[i]; weight_out[off + i] = r.Weights[i]; }
    off += nk[k];
    total += r.NbMatches();
    opt.AddResiduals(r);
  }
  if (total == 0) {
    ToRowMajor(T, pose_out);
    *lm_accepted_steps = 0; *lm_iterations = 0; *final_cost = 0.;
    if (err_out) std::memset(err_out, 0, sizeof(*err_out));
    return LSB_W_LOC_SKIPPED;
  }
  opt.Solve();
  ToRowMajor(opt.GetOptimizedPose(), pose_out);
  *lm_accepted_steps = opt.State.num_successful;
  *lm_iterations = opt.State.iter;
  *final_cost = opt.State.cost;
  if (err_out) *err_out = opt.EstimateRegistrationError();
  return 0;
}

ORC_API int orc_map_add(Slam* h, int type, const lsb_point* pts, size_t n, int fixed, double current_time)
{
  Cloud c(pts, pts + n);
  h->LocalMaps[type].Add(c, fixed != 0, current_time);
  return 0;
}

ORC_API int orc_undistort(Slam*, lsb_point* pts, size_t n, const double* H0, const double* H1, double t0, double t1)
{
  lsb::Interpolator it;
  it.t0 = t0; it.t1 = t1;
  it.set_transforms(FromRowMajor(H0), FromRowMajor(H1));
  for (size_t i = 0; i < n; ++i) {
    const Iso3d T = it.at(pts[i].time);
    lsb::transform_point_f(T, pts[i].x, pts[i].y, pts[i].z);
  }
  return 0;
}

// Exposes the shared deterministic math for tests/test_detmath.py
ORC_API void orc_detmath(int fn, const double* a, const double* b, double* out, long n)
{
  for (long i = 0; i < n; ++i) {
    switch (fn) {
      case 0: out[i] = lsb::dm::sin(a[i]); break;
      case 1: out[i] = lsb::dm::cos(a[i]); break;
      case 2: out[i] = lsb::dm::atan2(a[i], b[i]); break;
      case 3: out[i] = lsb::dm::acos(a[i]); break;
      case 4: out[i] = lsb::dm::asin(a[i]); break;
      case 5: out[i] = lsb::dm::exp(a[i]); break;
      case 6: out[i] = (double)lsb::dm::expf((float)a[i]); break;
      default: out[i] = 0.;
    }
  }
}
// eigen33 on n symmetric matrices (row-major 9 doubles each): evals ascending + evecs row-major
ORC_API void orc_eigen33(const double* m9, double* evals3, double* evecs9, long n, int as_float)
{
  for (long i = 0; i < n; ++i) {
    if (as_float) {
      Mat3f M, V; float ev[3];
      for (int j = 0; j < 9; ++j) M.m[j] = (float)m9[i * 9 + j];
      lsb::eigen33(M, V, ev);
      for (int j = 0; j < 9; ++j) evecs9[i * 9 + j] = V.m[j];
      for (int j = 0; j < 3; ++j) evals3[i * 3 + j] = ev[j];
    } else {
      Mat3d M, V; double ev[3];
      for (int j = 0; j < 9; ++j) M.m[j] = m9[i * 9 + j];
      lsb::eigen33(M, V, ev);
      for (int j = 0; j < 9; ++j) evecs9[i * 9 + j] = V.m[j];
      for (int j = 0; j < 3; ++j) evals3[i * 3 + j] = ev[j];
    }
  }
}

// The 28 terms of lm_eval_record for n records (A row-major 9, P 3, X 3, w) at pose x[6]
// (tests/test_oracle_math.py checks them against a numpy restatement of the reference residual
// r = sqrt(w rho') A (R X + t - P) with a finite-difference Jacobian).
ORC_API void orc_lm_terms(const double* A9, const double* P3, const double* X3, const double* w, long n, const double* x6, double sat,
                          double* out28)
{
  lsb::LmPoint pt;
  lsb::lm_make_point(x6, pt);
  for (long i = 0; i < n; ++i) {
    lsb::LmRecord r;
    lsb::lm_record_set(r, A9 + 9 * i, P3 + 3 * i, X3 + 3 * i, w[i]);
    lsb::lm_eval_record(r, pt, sat, out28 + 28 * i);
  }
}
// LocalOptimizer::Solve on raw residual records; returns the number of LM iterations.
// sensor14 (nullable): wheel {use, prev xyz, distance, weight}, gravity {use, ref xyz, current xyz, weight}
ORC_API int orc_lm_solve_sensors(const double* A9, const double* P3, const double* X3, const double* w, long n, const double* x0, double sat,
                                 int max_iter, int two_d, int threads, const double* sensor14, double* x_out, double* cost_out,
                                 int* accepted_out, int* term_out)
{
  Optimizer opt;
  if (sensor14) {
    lsb::LmSensorTerms& c = opt.SensorTerms;
    c.use_wheel = sensor14[0] != 0.;
    for (int i = 0; i < 3; ++i) c.wheel_prev[i] = sensor14[1 + i];
    c.wheel_dist = sensor14[4]; c.wheel_weight = sensor14[5];
    c.use_gravity = sensor14[6] != 0.;
    for (int i = 0; i < 3; ++i) { c.g_ref[i] = sensor14[7 + i]; c.g_cur[i] = sensor14[10 + i]; }
    c.gravity_weight = sensor14[13];
  }
  opt.TwoDMode = two_d != 0;
  opt.LMMaxIter = (unsigned)max_iter;
  opt.NbThreads = threads > 0 ? threads : 1;
  opt.Sat = sat;
  for (int i = 0; i < 6; ++i) opt.PoseArray[i] = x0[i];
  for (long i = 0; i < n; ++i) {
    lsb::LmRecord r;
    lsb::lm_record_set(r, A9 + 9 * i, P3 + 3 * i, X3 + 3 * i, w[i]);
    opt.Records.push_back(r);
  }
  opt.Solve();
  for (int i = 0; i < 6; ++i) x_out[i] = opt.State.x[i];
  *cost_out = opt.State.cost;
  *accepted_out = opt.State.num_successful;
  *term_out = opt.State.term;
  return opt.State.iter;
}
// the same with the per-evaluation trace (rows of 5 doubles, see Optimizer::Trace) and the normal matrix at the solution
ORC_API int orc_lm_solve_trace(const double* A9, const double* P3, const double* X3, const double* w, long n, const double* x0, double sat,
                               int max_iter, int two_d, int threads, double* x_out, double* cost_out, int* accepted_out, int* term_out,
                               double* trace_out, int trace_cap, int* trace_rows, double* H21_out)
{
  Optimizer opt;
  opt.TwoDMode = two_d != 0;
  opt.LMMaxIter = (unsigned)max_iter;
  opt.NbThreads = threads > 0 ? threads : 1;
  opt.Sat = sat;
  for (int i = 0; i < 6; ++i) opt.PoseArray[i] = x0[i];
  for (long i = 0; i < n; ++i) {
    lsb::LmRecord r;
    lsb::lm_record_set(r, A9 + 9 * i, P3 + 3 * i, X3 + 3 * i, w[i]);
    opt.Records.push_back(r);
  }
  opt.Solve();
  for (int i = 0; i < 6; ++i) x_out[i] = opt.State.x[i];
  *cost_out = opt.State.cost;
  *accepted_out = opt.State.num_successful;
  *term_out = opt.State.term;
  const int rows = std::min<int>((int)opt.Trace.size(), trace_cap);
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < 5; ++c) trace_out[5 * r + c] = opt.Trace[r][c];
  *trace_rows = (int)opt.Trace.size();
  if (H21_out) for (int i = 0; i < 21; ++i) H21_out[i] = opt.State.H[i];
  return opt.State.iter;
}
// LocalOptimizer::EstimateRegistrationError on a given normal matrix (upper triangle, row-major, 21 entries)
ORC_API int orc_registration_error(const double* H21, int two_d, lsb_registration_error* out)
{
  lsb::registration_error_from_normal_matrix(H21, two_d, out);
  return 0;
}
ORC_API int orc_lm_solve(const double* A9, const double* P3, const double* X3, const double* w, long n, const double* x0, double sat,
                         int max_iter, int two_d, int threads, double* x_out, double* cost_out, int* accepted_out, int* term_out)
{
  return orc_lm_solve_sensors(A9, P3, X3, w, n, x0, sat, max_iter, two_d, threads, nullptr, x_out, cost_out, accepted_out, term_out);
}

}  // extern "C"

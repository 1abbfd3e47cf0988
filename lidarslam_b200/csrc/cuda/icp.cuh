// Device-resident ICP loop state.  The ICP - LM loops of Slam::ComputeEgoMotion
// (reference slam_lib/src/Slam.cxx:876-954) and Slam::Localization (:1071-1156) run match -> solve -> pose
// update -> undistortion refinement for up to ICPMaxIter iterations; every decision between two iterations
// (enough matches?, the pose as XYZRPY, the new within-frame motion, the "no LM step was taken" break) is
// taken here on the device by one thread at the tail of the persistent LM kernel (icp_control), so the host
// enqueues all iterations back to back and reads one result block when the loop is over.  The kernels of an
// iteration that comes after the break see `done` and return at once.
#pragma once
#include "../common/linalg.h"
#include "../common/lm_core.h"
#include "../../../include/lidarslam_b200.h"

namespace lsb {

enum { ICP_STAGE_EGO = 0, ICP_STAGE_LOC = 1 };

struct IcpState {
  Iso3d Tworld, PreviousTworld, Trelative;
  Interpolator Within;       // Slam::WithinFrameMotion
  Interpolator Pending;      // the refinement still to be applied to the keypoint clouds (composition of RefineUndistortion)
  // InterpolateScanPose (:1271-1285) inputs
  int log_empty;
  double prev_pose_time, curr_pose_time, max_extrapolation_ratio;
  int undistortion;          // lsb_params.undistortion
  // results
  int done;                  // the loop is over: later iterations do nothing
  int final_refine;          // Pending has to be applied once more after the loop (last iteration, not deferred)
  int warnings;              // LSB_W_* raised inside the loop
  int icp_iters;             // iterations executed (Stats.*_icp_iters)
  int lm_iters;              // LM iterations summed over the solves that counted (Stats.*_lm_iters)
  int last_iter;             // index of the last executed iteration (its match histograms are the ones kept)
  int lm_valid;              // last_lm holds the state of the breaking iteration's solve (registration error input)
  unsigned total_matched;    // TotalMatchedKeypoints of the last executed iteration
  LmState last_lm;
};

// per-solve constants (kernel argument)
struct IcpCtl {
  int stage;                 // ICP_STAGE_*
  int icp_iter, icp_max_iter;
  unsigned min_matches;
};

// Slam::InterpolateScanPose (:1271-1285)
LSB_HD Iso3d icp_interpolate_scan_pose(IcpState& s, double time)
{
  if (s.log_empty) return s.Tworld;
  const double dt = s.curr_pose_time - s.prev_pose_time;
  const double r = time / dt;
  if (((r < 0.0) ? -r : r) > s.max_extrapolation_ratio) { s.warnings |= LSB_W_EXTRAPOLATION; return s.Tworld; }
  return linear_interpolation(s.PreviousTworld, s.Tworld, s.curr_pose_time + time, s.prev_pose_time, s.curr_pose_time);
}

// Slam::RefineUndistortion (:1322-1352): new within-frame motion from the current pose; Pending = the transform that
// takes the keypoints from the previous undistortion to the new one
LSB_HD void icp_refine_undistortion(IcpState& s)
{
  const Iso3d prevBegin = s.Within.H0(), prevEnd = s.Within.H1();
  const Iso3d wBegin = icp_interpolate_scan_pose(s, s.Within.t0);
  const Iso3d wEnd = icp_interpolate_scan_pose(s, s.Within.t1);
  const Iso3d baseToWorld = inverse(s.Tworld);
  const Iso3d newBegin = baseToWorld * wBegin, newEnd = baseToWorld * wEnd;
  s.Within.set_transforms(newBegin, newEnd);
  s.Pending = s.Within;
  s.Pending.set_transforms(newBegin * inverse(prevBegin), newEnd * inverse(prevEnd));
}

// What the host loop did between a solve and the next matching pass.  lm = final state of the solve of this iteration,
// matched = matches of this iteration (sum of the SUCCESS bins).
LSB_HD void icp_control(IcpState& s, const IcpCtl& c, const LmState& lm, unsigned matched)
{
  s.icp_iters += 1;
  s.last_iter = c.icp_iter;
  s.total_matched = matched;
  if (matched < c.min_matches) {
    if (c.stage == ICP_STAGE_EGO) s.warnings |= LSB_W_EGO_SKIPPED;                      // :919-923
    else {                                                                              // :1098-1107
      s.Trelative = Iso3d::identity();
      s.Tworld = s.PreviousTworld;
      if (s.undistortion) s.Within.set_transforms(Iso3d::identity(), Iso3d::identity());
      s.warnings |= LSB_W_LOC_SKIPPED;
    }
    s.done = 1;
    return;
  }
  s.lm_iters += lm.iter;
  const bool last = (lm.num_successful + 1 == 1) || c.icp_iter == c.icp_max_iter - 1;  // summary.num_successful_steps == 1 (:950, :1151)
  if (c.stage == ICP_STAGE_EGO) {
    s.Trelative = xyzrpy_to_iso(lm.x);                                                  // :946-947
    if (last) s.done = 1;                                                               // :950, or the loop ran out
    return;
  }
  s.Tworld = xyzrpy_to_iso(lm.x);                                                       // :1140-1141
  s.Trelative = inverse(s.PreviousTworld) * s.Tworld;
  if (s.undistortion == LSB_UNDISTORTION_REFINED) {                                     // :1146-1147
    icp_refine_undistortion(s);
    if (last) s.final_refine = 1;
  }
  if (last) {                                                                           // :1151-1155
    s.last_lm = lm;
    s.lm_valid = 1;
    s.done = 1;
  }
}

}  // namespace lsb

// External-sensor constraints of the localization solve: the host-side managers that
// turn wheel-odometer and IMU measurements into the two extra residual blocks of lm_core.h
// (reference slam_lib/include/LidarSlam/SensorConstraints.h:13-118, src/SensorConstraints.cxx:8-190;
// call sites Slam.cxx:256-261, 346-352, 1123-1131, 1584-1598).  Plain host arithmetic on a handful of
// scalars per frame (like common/motion_limits.h): the residuals themselves are evaluated inside
// the LM loop (lm_add_sensor_terms).
#pragma once
#include <cmath>
#include <cstdio>
#include <vector>
#include "lm_core.h"

namespace lsb {

struct WheelOdomMeasurement { double Time = 0., Distance = 0.; };          // SensorConstraints.h:13-17
struct GravityMeasurement { double Time = 0.; double Acceleration[3] = {0., 0., 0.}; };   // :19-23

namespace sensor_detail {
// Eigen's normalized(): v / sqrt(v.v) if v.v > 0, else v
inline void Normalized(const double* v, double* o)
{
  const double z = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
  const double n = (z > 0.) ? std::sqrt(z) : 1.;
  for (int i = 0; i < 3; ++i) o[i] = v[i] / n;
}
inline double Norm(const double* v) { return std::sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]); }
}  // namespace sensor_detail

// SensorManager<T> (SensorConstraints.h:25-76)
template <typename T>
struct SensorManager {
  int PreviousIdx = -1;
  std::vector<T> Measures;
  double Weight = 0.;
  double TimeOffset = 0.;
  bool HasResidual = false;          // Residual.Cost != nullptr

  void AddMeasurement(const T& m) { Measures.push_back(m); }
  void ResetResidual() { HasResidual = false; }
  void Reset()
  {
    ResetResidual();
    Measures.clear();
    PreviousIdx = -1;
    TimeOffset = 0.;
  }
  bool CanBeUsed() const { return Weight > 1e-6 && !Measures.empty(); }

  // Index of the last measurement before `t` starting from PreviousIdx (SensorConstraints.cxx:23-30);
  // false if no measurement brackets `t`.  The reference dereferences Measures[-1 + 1] first, i.e. the
  // walk starts at element 0; when t equals the first stamp it would leave the index at -1 and read
  // Measures[-1] (undefined): the index is clamped to 0 here.
  bool Bracket(double t, int& idx)
  {
    if (t < Measures.front().Time || t > Measures.back().Time) return false;
    if (PreviousIdx >= 0 && Measures[PreviousIdx].Time > t) PreviousIdx = -1;
    idx = PreviousIdx;
    while (idx + 1 < (int)Measures.size() - 1 && Measures[idx + 1].Time < t) ++idx;
    if (idx < 0) idx = 0;
    return idx + 1 < (int)Measures.size();
  }
};

// WheelOdometryManager (SensorConstraints.h:78-97).  Slam uses the ABSOLUTE constraint (Slam.cxx:350):
// distance travelled since the first constrained frame, measured from PreviousPose, which Slam never
// sets (identity: the origin of the trajectory).
struct WheelOdometryManager : SensorManager<WheelOdomMeasurement> {
  double PreviousPos[3] = {0., 0., 0.};
  double PreviousDistance = 0.;
  double ResidualDistance = 0.;

  // returns false with *warn set when the times do not match (PRINT_WARNING, SensorConstraints.cxx:19)
  void ComputeWheelAbsoluteConstraint(double lidarTime, bool* warn)        // SensorConstraints.cxx:8-57
  {
    ResetResidual();
    if (!CanBeUsed()) return;
    lidarTime -= TimeOffset;
    int cur;
    if (Measures.size() < 2 || !Bracket(lidarTime, cur)) { if (warn) *warn = true; return; }
    const double rt = (lidarTime - Measures[cur].Time) / (Measures[cur + 1].Time - Measures[cur].Time);
    const double dist = (1 - rt) * Measures[cur].Distance + rt * Measures[cur + 1].Distance;
    if (PreviousIdx == -1) {          // no memory yet: remember the abscissa of this frame, no constraint
      PreviousIdx = cur;
      PreviousDistance = dist;
      return;
    }
    ResidualDistance = dist - PreviousDistance;
    HasResidual = true;
    PreviousIdx = cur;
  }
};

// ImuManager (SensorConstraints.h:99-115)
struct ImuManager : SensorManager<GravityMeasurement> {
  double GravityRef[3] = {0., 0., 0.};
  double GravityDir[3] = {0., 0., 0.};

  void ComputeGravityRef(double deltaAngle)                                 // SensorConstraints.cxx:152-190
  {
    const double kPi = 3.14159265358979323846;
    const int NPhi = (int)std::ceil(2 * kPi / deltaAngle), NTheta = (int)std::ceil(kPi / deltaAngle);
    std::vector<std::vector<int>> histogram((size_t)NPhi * NTheta);
    std::vector<double> dirs(Measures.size() * 3);
    for (size_t i = 0; i < Measures.size(); ++i) {
      double* d = &dirs[3 * i];
      sensor_detail::Normalized(Measures[i].Acceleration, d);
      int iphi = (int)((std::atan2(d[1], d[0]) + kPi) / deltaAngle);
      int ith = (int)(std::acos(d[2]) / deltaAngle);
      iphi = iphi < 0 ? 0 : (iphi >= NPhi ? NPhi - 1 : iphi);                // the reference indexes out of range at phi = pi
      ith = ith < 0 ? 0 : (ith >= NTheta ? NTheta - 1 : ith);
      histogram[(size_t)iphi * NTheta + ith].push_back((int)i);
    }
    size_t best = 0;
    for (size_t b = 0; b < histogram.size(); ++b)                            // phi-major, first largest bin wins
      if (histogram[b].size() > histogram[best].size()) best = b;
    double sum[3] = {0., 0., 0.};
    for (int i : histogram[best])
      for (int a = 0; a < 3; ++a) sum[a] += dirs[3 * i + a];
    sensor_detail::Normalized(sum, GravityRef);
  }

  void ComputeGravityConstraint(double lidarTime, bool* warn)              // SensorConstraints.cxx:110-150
  {
    ResetResidual();
    if (!CanBeUsed()) return;
    lidarTime -= TimeOffset;
    if (Measures.size() < 2 || lidarTime < Measures.front().Time || lidarTime > Measures.back().Time) { if (warn) *warn = true; return; }
    if (sensor_detail::Norm(GravityRef) < 1e-6) ComputeGravityRef(5.f * 3.14159265358979323846 / 180.);
    int cur;
    if (!Bracket(lidarTime, cur)) { if (warn) *warn = true; return; }
    const double rt = (lidarTime - Measures[cur].Time) / (Measures[cur + 1].Time - Measures[cur].Time);
    double a[3], b[3], g[3];
    sensor_detail::Normalized(Measures[cur].Acceleration, a);
    sensor_detail::Normalized(Measures[cur + 1].Acceleration, b);
    for (int i = 0; i < 3; ++i) g[i] = (1 - rt) * a[i] + rt * b[i];
    if (!(sensor_detail::Norm(g) > 1e-6)) return;                            // inconsistent IMU measure
    sensor_detail::Normalized(g, GravityDir);
    HasResidual = true;
    PreviousIdx = cur;
  }
};

// the residual blocks the managers hold, in the form the LM loop consumes (Slam.cxx:1123-1131)
inline LmSensorTerms MakeSensorTerms(const WheelOdometryManager& w, const ImuManager& g)
{
  LmSensorTerms c;
  lm_sensor_terms_clear(c);
  if (w.HasResidual) {
    c.use_wheel = 1;
    for (int i = 0; i < 3; ++i) c.wheel_prev[i] = w.PreviousPos[i];
    c.wheel_dist = w.ResidualDistance;
    c.wheel_weight = w.Weight;
  }
  if (g.HasResidual) {
    c.use_gravity = 1;
    for (int i = 0; i < 3; ++i) { c.g_ref[i] = g.GravityRef[i]; c.g_cur[i] = g.GravityDir[i]; }
    c.gravity_weight = g.Weight;
  }
  return c;
}

}  // namespace lsb

// Slam::CheckMotionLimits (reference slam_lib/src/Slam.cxx:1391-1484): velocity / acceleration of the
// pose over a time window of the logged trajectory against user limits.  Host-only arithmetic on a
// handful of poses; shared by the library's host driver and the CPU oracle.
#pragma once
#include <cfloat>
#include <cmath>
#include "linalg.h"

namespace lsb {

struct MotionLimitsState {
  float PreviousVelocity[2] = {0.f, 0.f};   // Slam.h:663 (m/s, deg/s)
  bool Comply = true;                       // Slam.h:660
};

// log[i].T (Iso3d) / log[i].time (s): the trajectory BEFORE the current pose is appended.
// Returns warning bits: 1 = window too short / not enough poses, 2 = does not comply.
template <class Log>
inline int check_motion_limits(const Log& log, const Iso3d& Tworld, double currentTimeStamp, float timeWindowDuration, const float velocityLimits[2],
                               const float accelerationLimits[2], unsigned nbrFrameProcessed, MotionLimitsState& st)
{
  int warn = 0;
  const int nPoses = (int)log.size();
  if (nPoses == 0) return 0;
  double deltaTime = currentTimeStamp - log.back().time;
  double nextDeltaTime = FLT_MAX;
  int startIndex = nPoses - 1;
  if (deltaTime < timeWindowDuration) {
    while (startIndex >= 0) {
      deltaTime = nextDeltaTime;
      nextDeltaTime = currentTimeStamp - log[startIndex].time;
      if (nextDeltaTime >= timeWindowDuration) break;
      --startIndex;
    }
    if (startIndex < 0) { warn |= 1; startIndex = 0; }
    else if (std::abs(deltaTime - timeWindowDuration) < std::abs(nextDeltaTime - timeWindowDuration)) ++startIndex;
    deltaTime = currentTimeStamp - log[startIndex].time;
  } else warn |= 1;
  st.Comply = true;
  const Iso3d TWindow = inverse(log[startIndex].T) * Tworld;
  float angle = (float)rotation_angle(TWindow.R);          // Eigen::AngleAxisd(R).angle()
  if (angle > M_PI) angle = (float)(2 * M_PI - angle);
  angle = (float)(angle / M_PI * 180.);                    // Utils::Rad2Deg<float>
  const float distance = (float)norm(TWindow.t);
  const float velocity[2] = {(float)(distance / deltaTime), (float)(angle / deltaTime)};
  if (nbrFrameProcessed >= 2) {
    const float dt = (float)deltaTime;                     // Array2f / double: the scalar is cast to float
    const float acc[2] = {(velocity[0] - st.PreviousVelocity[0]) / dt, (velocity[1] - st.PreviousVelocity[1]) / dt};
    const bool complyVelocity = velocity[0] < velocityLimits[0] && velocity[1] < velocityLimits[1];
    const bool complyAcceleration = std::fabs(acc[0]) < accelerationLimits[0] && std::fabs(acc[1]) < accelerationLimits[1];
    st.Comply = complyVelocity && complyAcceleration;
  }
  st.PreviousVelocity[0] = velocity[0];
  st.PreviousVelocity[1] = velocity[1];
  if (!st.Comply) warn |= 2;
  return warn;
}

}  // namespace lsb

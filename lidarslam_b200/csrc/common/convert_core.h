// Per-point arithmetic of the driver-cloud adaptors (reference
// ros_wrapping/lidar_conversions/src/Utilities.h:58-108, VelodyneToLidarNode.cxx:80-106,
// RobosenseToLidarNode.cxx:82-131), shared by the CUDA kernels (cuda/convert.cu) and the oracle.
#pragma once
#include "detmath.h"
#include "../../../include/lidarslam_b200.h"

namespace lsb {

#define LSB_CONV_PI 3.14159265358979323846

// normalised azimuth in [0, 1] (Utilities.h:86): std::atan2 on the float fields is the float overload
LSB_HD double conv_point_advancement(float x, float y)
{
  return (LSB_CONV_PI - (double)dm::atan2f(y, x)) / (2 * LSB_CONV_PI);
}

// advancement relative to the first point of the frame, wrapped to [0, 1) (Utilities.h:92-94); fmod is exact
LSB_HD double conv_wrap_advancement(double pointAdvancement, double initAdvancement)
{
  return ::fmod(1. + ::fmod(pointAdvancement - initAdvancement, 1.), 1.);
}

// VelodyneToLidarNode.cxx:103-104
LSB_HD double conv_time_from_advancement(double frameAdvancement, int timestampFirstPacket, double rpm)
{
  return (timestampFirstPacket ? frameAdvancement : frameAdvancement - 1) / rpm * 60.;
}

// RobosenseToLidarNode.cxx:129-130
LSB_HD double conv_rs_time(unsigned i, unsigned pointsPerRing, double rpm)
{
  const double frameAdvancement = static_cast<double>(i % pointsPerRing) / pointsPerRing;
  return (frameAdvancement - 1) / rpm * 60.;
}

// RobosenseToLidarNode.cxx:32 LASER_ID_MAPPING_RS16
LSB_HD unsigned conv_rs16_laser_id(unsigned id)
{
  return id < 8 ? id : 23 - id;     // 0..7, 15, 14, ..., 8
}

// custom mapping (vector<int> LaserIdMapping); an id outside the table is left unchanged (the reference reads out of range)
LSB_HD unsigned conv_map_laser_id(unsigned id, const int32_t* mapping, int n_mapping)
{
  return (mapping && (int)id < n_mapping) ? (unsigned)mapping[id] : id;
}

LSB_HD lsb_point conv_make_point(float x, float y, float z, float intensity, unsigned laser_id, int device_id, double time)
{
  lsb_point p;
  p.x = x; p.y = y; p.z = z; p.w = 1.f;
  p.time = time;
  p.intensity = intensity;
  p.laser_id = (uint16_t)laser_id;
  p.device_id = (uint8_t)device_id;
  p.label = 0;
  return p;
}

}  // namespace lsb

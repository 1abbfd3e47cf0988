// CPU ORACLE (test infrastructure) — the driver-cloud adaptors, restated loop for loop
// (reference ros_wrapping/lidar_conversions/src/VelodyneToLidarNode.cxx:50-111,
// RobosenseToLidarNode.cxx:58-137, Utilities.h:58-108): sequential, with the estimator's
// per-ring map and the "last kept point" comparison exactly as the reference carries them.
// Per-point arithmetic from common/convert_core.h (std::atan2f -> dm::atan2f).
#include <cmath>
#include <map>
#include <vector>
#include "orc_slam.h"
#include "../lidarslam_b200/csrc/common/convert_core.h"

#define ORC_API extern "C" __attribute__((visibility("default")))
using orc::Slam;

namespace {

// Utilities.h:58-108
struct SpinningFrameAdvancementEstimator {
  std::map<int, double> PreviousAdvancementPerRing;
  double InitAdvancement = 0.;
  double operator()(const lsb_point& point)
  {
    const double pointAdvancement = lsb::conv_point_advancement(point.x, point.y);
    if (PreviousAdvancementPerRing.empty()) InitAdvancement = pointAdvancement;
    double frameAdvancement = lsb::conv_wrap_advancement(pointAdvancement, InitAdvancement);
    if (frameAdvancement < PreviousAdvancementPerRing[point.laser_id]) frameAdvancement += 1.;
    PreviousAdvancementPerRing[point.laser_id] = frameAdvancement;
    return frameAdvancement;
  }
};

lsb_conversion_params Defaults(const lsb_conversion_params* p)
{
  lsb_conversion_params d;
  d.laser_id_mapping = nullptr; d.n_laser_id_mapping = 0; d.device_id = 0; d.rpm = 600.; d.timestamp_first_packet = 0;
  return p ? *p : d;
}

long Out(Slam* h, lsb_point* out, long cap)
{
  const std::vector<lsb_point>& c = h->Converted;
  if (out) for (long i = 0; i < cap && i < (long)c.size(); ++i) out[i] = c[i];
  return (long)c.size();
}

}  // namespace

ORC_API long orc_convert_velodyne(Slam* h, const lsb_velodyne_point* pts, size_t n, const lsb_conversion_params* pp, lsb_point* out, long cap)
{
  const lsb_conversion_params p = Defaults(pp);
  std::vector<lsb_point>& cloudS = h->Converted;
  cloudS.clear();
  if (n == 0) return 0;
  const bool useLaserIdMapping = p.laser_id_mapping && p.n_laser_id_mapping > 0;
  const bool isTimeValid = pts[n - 1].time - pts[0].time > 1e-8;
  SpinningFrameAdvancementEstimator frameAdvancementEstimator;
  for (size_t i = 0; i < n; ++i) {
    const lsb_velodyne_point& v = pts[i];
    const unsigned laser = useLaserIdMapping ? lsb::conv_map_laser_id(v.ring, p.laser_id_mapping, p.n_laser_id_mapping) : v.ring;
    lsb_point s = lsb::conv_make_point(v.x, v.y, v.z, v.intensity, laser, p.device_id, 0.);
    if (isTimeValid) s.time = v.time;
    else s.time = lsb::conv_time_from_advancement(frameAdvancementEstimator(s), p.timestamp_first_packet, p.rpm);
    cloudS.push_back(s);
  }
  return Out(h, out, cap);
}

ORC_API long orc_convert_robosense(Slam* h, const lsb_xyzi_point* pts, uint32_t width, uint32_t height, const lsb_conversion_params* pp,
                                   lsb_point* out, long cap)
{
  const lsb_conversion_params p = Defaults(pp);
  std::vector<lsb_point>& cloudS = h->Converted;
  cloudS.clear();
  const size_t n = (size_t)width * height;
  if (n == 0) return 0;
  const unsigned nLasers = height, pointsPerRing = (unsigned)(n / nLasers);
  const bool useLaserIdMapping = p.laser_id_mapping && p.n_laser_id_mapping > 0;
  for (unsigned i = 0; i < n; ++i) {
    const lsb_xyzi_point& r = pts[i];
    if (!(std::isfinite(r.x) && std::isfinite(r.y) && std::isfinite(r.z))) continue;
    if (!cloudS.empty() && r.x == cloudS.back().x && r.y == cloudS.back().y && r.z == cloudS.back().z) continue;
    unsigned laser = i / width;
    laser = useLaserIdMapping ? lsb::conv_map_laser_id(laser, p.laser_id_mapping, p.n_laser_id_mapping)
                              : (nLasers == 16 ? lsb::conv_rs16_laser_id(laser) : laser);
    cloudS.push_back(lsb::conv_make_point(r.x, r.y, r.z, r.intensity, laser, p.device_id, lsb::conv_rs_time(i, pointsPerRing, p.rpm)));
  }
  return Out(h, out, cap);
}

ORC_API int orc_add_converted_frame(Slam* h, uint64_t stamp_us, uint32_t seq)
{
  std::vector<orc::Frame> f;
  f.push_back(orc::Frame{h->Converted.data(), h->Converted.size(), stamp_us, seq});
  return h->AddFrames(f);
}

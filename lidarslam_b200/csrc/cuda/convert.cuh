// Input adaptors on the device: raw driver clouds -> LidarPoint frames, so that a
// driver cloud goes host -> HBM once and never comes back (reference ros_wrapping/lidar_conversions:
// VelodyneToLidarNode.cxx:50-111, RobosenseToLidarNode.cxx:58-137, Utilities.h:58-108).
#pragma once
#include "dev_util.h"
#include "../common/convert_core.h"

namespace lsb {

struct ConvertWorkspace {
  DevBuf<lsb_point> out;            // the converted frame (input of lsb_add_converted_frame)
  size_t n_out = 0;
  DevBuf<unsigned char> raw;        // the driver cloud as uploaded
  DevBuf<int32_t> mapping;
  DevBuf<uint16_t> key_a, key_b;
  DevBuf<uint32_t> idx_a, idx_b, first, flags, scan;
  DevBuf<int32_t> last_finite_in, last_finite;
  DevBuf<double> adv;
  DevBuf<unsigned char> cub_tmp;
  PinnedBuf<uint32_t> h_count;

  // both return the number of points of the converted frame (left in `out`)
  size_t velodyne(const lsb_velodyne_point* host_pts, size_t n, const lsb_conversion_params& p, cudaStream_t s, LaunchCounter& lc);
  size_t robosense(const lsb_xyzi_point* host_pts, uint32_t width, uint32_t height, const lsb_conversion_params& p, cudaStream_t s,
                   LaunchCounter& lc);
private:
  void upload_mapping(const lsb_conversion_params& p, cudaStream_t s);
};

}  // namespace lsb

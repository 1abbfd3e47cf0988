// PCD wire format of keypoint maps: what pcl::io::savePCDFile{ASCII,Binary,
// BinaryCompressed}<LidarPoint> write and pcl::io::loadPCDFile<LidarPoint> reads, restated
// without PCL (reference call sites: PointCloudStorage.h:85-115, Slam.cxx:504-543; point
// fields: LidarPoint.h:68-77).  Host-only byte work: the maps live on the device, files on
// the host, and a map is saved a few times per run.
#pragma once
#include "../../../include/lidarslam_b200.h"
#include <string>
#include <vector>

namespace lsb {
namespace pcd {

// PCD v0.7 header of a LidarPoint cloud of n points (unorganised: WIDTH n, HEIGHT 1), without the DATA line.
std::string Header(size_t n);

// LZF (liblzf wire format, the codec of "DATA binary_compressed").
// Compress returns the number of bytes written (0 if `cap` is too small).
size_t LzfCompress(const unsigned char* in, size_t n, unsigned char* out, size_t cap);
// Decompress returns the number of bytes produced, 0 on a malformed stream or if `cap` is too small.
size_t LzfDecompress(const unsigned char* in, size_t n, unsigned char* out, size_t cap);

// 0 on success; -1 cannot open / write, -3 empty cloud, -4 unknown format (codes of savePointCloudToPCD).
int Write(const std::string& path, const lsb_point* pts, size_t n, int format);
// 0 on success, -1 on any failure (as pcl::io::loadPCDFile); `err` gets the reason.
int Read(const std::string& path, std::vector<lsb_point>& out, std::string* err);

}  // namespace pcd
}  // namespace lsb

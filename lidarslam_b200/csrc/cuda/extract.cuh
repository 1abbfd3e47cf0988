// GPU SpinningSensorKeypointExtractor (kernels K1-K5 of DESIGN.md).
// Replaces reference slam_lib/src/SpinningSensorKeypointExtractor.cxx:118-637.
#pragma once
#include "dev_util.h"

namespace lsb {

constexpr int KE_MAX_RINGS = 1024;          // laser_id capacity (LSB_E_CAPACITY beyond)
constexpr int KE_MAX_WIDTH = 8;             // neighbor_width capacity
constexpr size_t KE_MAX_POINTS = (1u << 21) - 1;   // 21-bit packed scan fields

struct ExtractorGpu {
  lsb_ke_params P;
  size_t n = 0;                    // points of the last frame
  unsigned nb_rings = 0;           // max laser_id + 1 (never shrinks, SSKE.cxx:156-164)
  uint32_t counts[3] = {0, 0, 0};  // keypoints per type of the last frame
  unsigned az_estimations = 0;     // times the azimuthal resolution was estimated (once per extractor unless reset, SSKE.cxx:169)
  double time_min[3] = {0, 0, 0}, time_max[3] = {0, 0, 0};   // range of their `time` field (valid where counts[k] > 0)

  // ring-major ("scan line") order: position j <-> (ring, index in ring)
  DevBuf<uint32_t> key_a, key_b, idx_a, src;   // src[j] = input index of ring-ordered point j
  DevBuf<float4> p4;                           // (x, y, z, intensity)
  DevBuf<int> ring_start, ring_end;            // [KE_MAX_RINGS]
  DevBuf<uint8_t> valid, label, nms_state;     // bit k = Keypoint k
  DevBuf<float> angles, depth_gap, saliency, intensity_gap, nms_vals;
  DevBuf<unsigned long long> scan_in, scan_out;
  DevBuf<lsb_point> kp[3];                     // extracted keypoints, ring-major / index-minor
  DevBuf<uint32_t> d_counts;                   // 3 counts, nb_rings, [4] azimuth count, [8..19] time ranges
  DevBuf<float> az_angles, az_sorted;          // azimuthal resolution estimation
  DevBuf<float> d_az;                          // [2]: resolution, number of angles
  DevBuf<unsigned char> cub_tmp;
  PinnedBuf<uint32_t> h_counts;
  PinnedBuf<float> h_az;

  ExtractorGpu();
  // parameters and the state that outlives a frame (azimuth estimate, ring count): handed from one instance to the next when
  // several instances take turns (lsb_prefetch_frames)
  void adopt_scalars(const ExtractorGpu& o) { P = o.P; nb_rings = o.nb_rings; az_estimations = o.az_estimations; }
  // SpinningSensorKeypointExtractor::ComputeKeyPoints on a DEVICE cloud.
  // Synchronises the stream once (keypoint counts).
  void compute(const lsb_point* d_pts, size_t n, cudaStream_t s, LaunchCounter& lc);
  // GetDebugArray field (see lsb_get_ke_debug) into d_out[n], input order
  void debug_field(int field, float* d_out, cudaStream_t s, LaunchCounter& lc);
  // labels in input order (bit0 EDGE, bit1 PLANE, bit2 BLOB)
  void labels_input_order(uint8_t* d_out, cudaStream_t s, LaunchCounter& lc);
};

}  // namespace lsb

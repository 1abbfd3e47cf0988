// GPU RollingGrid: the keypoint map as a sort-based hash (kernels K7 / K12).
// Replaces reference slam_lib/src/RollingGrid.cxx:40-464
//   unordered_map<int, unordered_map<int, Voxel>>  ->  one array of leaves sorted by
//   the 64-bit key (outer 1-D voxel index << 32 | biased inner 1-D leaf index),
//   i.e. ascending (outer, inner): the canonical iteration order of the map.
#pragma once
#include "dev_util.h"
#include "knn.cuh"
#include "select.cuh"
#include "sort32.cuh"

namespace lsb {

struct RollingGridGpu {
  // parameters (RollingGrid.h:173-212)
  int GridSize = 50;
  double VoxelResolution = 10.;
  double LeafSize = 0.2;
  unsigned MinFramesPerVoxel = 0;
  int Sampling = LSB_SAMPLING_MAX_INTENSITY;
  double DecayingThreshold = -1.;
  float Position[3] = {0.f, 0.f, 0.f};   // VoxelGridPosition
  unsigned NbPoints = 0;                 // reference bookkeeping (not decremented by ClearOldPoints)

  // the map: n_entries leaves sorted by key
  size_t n_entries = 0;
  DevBuf<unsigned long long> keys, keys_b;
  DevBuf<lsb_point> pts, pts_b;
  DevBuf<uint32_t> counts, counts_b;
  DevBuf<unsigned long long> merge_keys;   // output of the merge launched ahead of the insert count (swapped in when there are inserts)
  DevBuf<lsb_point> merge_pts;
  DevBuf<uint32_t> merge_counts;

  // submap (target of the localization kNN)
  DevBuf<lsb_point> submap;
  size_t n_submap = 0;
  bool tree_valid = false;
  KnnGrid grid;
  float knn_cell_factor = 2.5f;          // kNN cell edge = factor * LeafSize

  // scratch
  DevBuf<unsigned long long> nk_a, nk_b, ins_keys;
  DevBuf<uint32_t> ni_a, ni_b, flags, scan, ins_counts;
  DevBuf<lsb_point> ins_pts;
  DevBuf<uint32_t> c_eff, c_tpos;        // SamplingMode::CENTROID: effective-point flags and their arrival positions
  DevBuf<uint32_t> d_info;               // [8]
  DevBuf<unsigned char> cub_tmp;
  PinnedBuf<uint32_t> h_info;            // [16]: [0..7] counters, [8] sequence word of the kernel that published them
  uint32_t info_seq = 0;
  bool box_by_kernel = false;            // h_bbox is being written by a fused box kernel with sequence word box_seq (else: by a copy on the stream)
  int box_seq = 0;                       // sequence word of the last box published in h_bbox[6] (fused box kernels)
  int NextBoxSeq() { box_seq = (box_seq + 1 > 0) ? box_seq + 1 : 1; return box_seq; }
  PinnedBuf<float> h_bbox;
  DevBuf<float> d_bbox;
  Sort32 sorter;                         // single-launch radix sort of the cloud's compact leaf keys (sort32.cu)
  SelectState sel_state;                 // single-pass compaction (select.cuh) of this map's launches, ordered on its stream
  DevBuf<int> box_partial;               // per-block partial boxes + ticket of the fused box kernels (slam_kernels.cu)
  DevBuf<unsigned> box_counter;

  void Reset(const float position[3], cudaStream_t s);                     // :40-48
  void Clear();                                                            // :51-56
  void SetGridSize(int size, cudaStream_t s, LaunchCounter& lc);           // :59-70
  void SetVoxelResolution(double res, cudaStream_t s, LaunchCounter& lc);  // :73-88
  // Get (:95-114): copies into d_out (device, capacity >= n_entries); returns the count
  size_t Get(bool clean, lsb_point* d_out, cudaStream_t s, LaunchCounter& lc);
  unsigned Size() const { return NbPoints; }
  void Roll(const float mn[3], const float mx[3], cudaStream_t s, LaunchCounter& lc);   // :117-157
  // Add (:160-318) of a DEVICE cloud
  void Add(const lsb_point* d_cloud, size_t n, bool fixed, double currentTime, bool roll, cudaStream_t s, LaunchCounter& lc);
  // the same in three host phases (one stream per map, read-backs overlapped)
  void AddBegin(const lsb_point* d_cloud, size_t n, bool roll, cudaStream_t s, LaunchCounter& lc);
  void AddSort(const lsb_point* d_cloud, size_t n, bool fixed, double currentTime, bool roll, cudaStream_t s, LaunchCounter& lc);
  void AddMerge(size_t n, cudaStream_t s, LaunchCounter& lc);
  void ClearOldPoints(double currentTime, cudaStream_t s, LaunchCounter& lc);           // :325-350
  void BuildSubMapKdTree(cudaStream_t s, LaunchCounter& lc);                             // :353-359
  void BuildSubMapKdTree(const float mn[3], const float mx[3], int minNbPoints, cudaStream_t s, LaunchCounter& lc);   // :362-442
  void SubMapSelect(const float mn[3], const float mx[3], int minNbPoints, cudaStream_t s, LaunchCounter& lc);
  void SubMapSelectDevice(int minNbPoints, cudaStream_t s, LaunchCounter& lc);   // bounds from d_bbox (device), no read-back first
  void SubMapFinish(cudaStream_t s, LaunchCounter& lc);
  bool IsSubMapKdTreeValid() const { return tree_valid && n_submap > 0; }               // RollingGrid.h:132-135
  bool IsTimeThreshold() const { return DecayingThreshold > 0; }
  // KdTree.Reset() (RollingGrid.cxx:55, :317): the search structure goes, the SubMap cloud stays readable (RollingGrid.h:160, :198)
  void ResetTree() { tree_valid = false; grid.clear(); }

 private:
  void build_grid(const float* lo, const float* hi, cudaStream_t s, LaunchCounter& lc);
  size_t compact_entries(cudaStream_t s, LaunchCounter& lc);   // keeps entries with flags[i] != 0
  void ensure_scratch(size_t n, cudaStream_t s);
  void submap_pass(int mode, size_t offset, cudaStream_t s, LaunchCounter& lc, const int* d_bb = nullptr);
  bool sub_from_device = false;
  int sub_lo[3] = {0, 0, 0}, sub_hi[3] = {-1, -1, -1}, sub_min_nb = -1;
  bool sub_filter = false;
};

// min/max of the xyz of a device cloud (pcl::getMinMax3D): enqueue + decode after a sync, or both with one sync
void cloud_bbox_async(const lsb_point* d_cloud, size_t n, DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox, cudaStream_t s, LaunchCounter& lc);
void bbox_decode(const float* h_bbox, float mn[3], float mx[3]);
void cloud_bbox(const lsb_point* d_cloud, size_t n, float mn[3], float mx[3], DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox,
                cudaStream_t s, LaunchCounter& lc);

}  // namespace lsb

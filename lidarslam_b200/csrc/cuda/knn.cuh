// Exact k-nearest-neighbour search on a uniform grid (kernels K7b/K8 of DESIGN.md).
// Replaces KDTreePCLAdaptor (reference slam_lib/include/LidarSlam/KDTreePCLAdaptor.h:57-105,
// nanoflann KDTreeSingleIndexAdaptor<L2_Simple_Adaptor<float>, 3>, leaf 16).
// Semantics kept: exact search, float distance d2 = ((dx^2 + dy^2) + dz^2), query
// = double world point cast to float (:96-101), results ascending, min(k, size)
// found.  Canonical tie order (north star): ascending (d2, index in the cloud).
#pragma once
#include "dev_util.h"
#include "../common/linalg.h"

namespace lsb {

constexpr int KNN_MAX_K = 32;

struct KnnGridParams {
  float ox, oy, oz;     // origin (min corner)
  float cell;           // cell edge
  int nx, ny, nz;
};

// Device-side inputs of a search inside the device-resident ICP loop (cuda/icp.cuh), all nullable
struct KnnDeviceArgs {
  const Iso3d* prior = nullptr;   // pose in device memory that moves the keypoints (overrides the host `prior`)
  const int* done = nullptr;      // when *done != 0 the search kernels return at once (outputs untouched)
  const Interpolator* refine = nullptr;   // with reuse: undistortion refinement (device memory) applied IN PLACE to the keypoints first
  bool reuse = false;             // idx_out / d2_out / nfound_out still hold the lists of the same queries against the same
                                  // target at a slightly different pose: they bound the search
};

bool knn_reuse_enabled();   // LSB_KNN_REUSE != 0 (A/B hook)

struct KnnGrid {
  const lsb_point* cloud = nullptr;   // device, not owned; indices refer to this order
  size_t m = 0;
  KnnGridParams gp{};
  DevBuf<uint32_t> key_a, idx_a;   // cell key and arrival rank inside the cell of every target point
  DevBuf<float4> sorted;          // cell-major points: (x, y, z, bits of the cloud index)
  DevBuf<uint32_t> cell_start;    // [ncells + 1] prefix offsets into sorted
  DevBuf<uint32_t> cell_count;    // [ncells + 1] points per cell (histogram; its exclusive scan is cell_start)
  DevBuf<unsigned char> cub_tmp;
  mutable DevBuf<int32_t> unres_list;    // queries handed to the brute-force kernel
  mutable DevBuf<unsigned> unres_count;  // [4]: unsettled count, cursor, brute-force count, cursor
  mutable bool counters_clean = false;   // unres_count is known to be zero (the matcher's fit kernel clears it behind a search), else a search memsets it first
  mutable DevBuf<int32_t> bf_list;       // exact search: isolated queries for the brute-force kernel
  mutable DevBuf<unsigned long long> bf_partial;   // their per-segment partial lists
  mutable DevBuf<unsigned> bf_tickets;   // per-query segment counter (self-resetting)
  DevBuf<float> d_bbox;           // [6]
  PinnedBuf<float> h_bbox;

  bool empty() const { return m == 0; }
  void clear() { cloud = nullptr; m = 0; }
  // Builds the grid over d_cloud[0..m).  lo/hi: host-known bounds containing every
  // point (nullptr => computed on the device, one stream sync).
  // d_m != nullptr (needs lo / hi): the number of points is read from device memory, `m` is an upper bound that sizes the
  // launches; the caller sets the true count with set_count() once it knows it (before any search).
  void build(const lsb_point* d_cloud, size_t m, float cell, const float* lo, const float* hi, cudaStream_t s, LaunchCounter& lc,
             const uint32_t* d_m = nullptr);
  void set_count(size_t true_m) { m = true_m; if (m == 0) cloud = nullptr; }

  // kNN of nq float queries (xyz packed).  idx_out/d2_out: [nq * k], missing
  // entries -1 / +inf.  nfound_out may be nullptr.  Layout: query-major [nq][k], or
  // entry-major [k][nq] (entry_major, coalesced for one-thread-per-query consumers).
  // cap_dist < 0: exact for every query (grid walk, then brute force for queries whose
  // neighbours are far away).  cap_dist >= 0: exact for every query whose k-th neighbour
  // lies within cap_dist; the others report nfound = min(k, m) with d2[k-1] > cap_dist^2
  // (possibly +inf), which is all the plane / blob matchers need (NEIGHBORS_TOO_FAR).
  void search(const float* d_queries_xyz, size_t nq, int k, float cap_dist, int32_t* d_idx_out, float* d_d2_out, int32_t* d_nfound_out,
              bool entry_major, cudaStream_t s, LaunchCounter& lc) const;
  // kNN of the keypoints kp[0..nq) moved by `prior` in double then cast to float
  // (KeypointsMatcher.cxx:118 + KDTreePCLAdaptor.h:96-101).
  void search_points(const lsb_point* d_kp, size_t nq, const Iso3d& prior, int k, float cap_dist, int32_t* d_idx_out, float* d_d2_out,
                     int32_t* d_nfound_out, bool entry_major, cudaStream_t s, LaunchCounter& lc, const KnnDeviceArgs* da = nullptr) const;
};

// Exact kNN of nq DEVICE queries in a DEVICE cloud without any allocation (static __device__ scratch, see knn.cu); enqueued on s.
void knn_search_noalloc(int device, const lsb_point* d_cloud, size_t m, const float* d_queries_xyz, size_t nq, int k, int32_t* d_idx_out,
                        float* d_d2_out, cudaStream_t s);

}  // namespace lsb

// Keypoint-to-model matching (K9) and the device Gauss-Newton / LM loop (K10/K11).
// Replaces reference slam_lib/src/KeypointsMatcher.cxx:33-480 (after the kNN),
// include/LidarSlam/CeresCostFunctions.h:105-152 and src/LocalOptimizer.cxx:74-102.
#pragma once
#include "dev_util.h"
#include "knn.cuh"
#include "../common/lm_core.h"
#include "icp.cuh"

namespace lsb {

// Everything one ICP iteration needs on the device.  Records / status / weight are
// index-aligned with the keypoints, the three types concatenated EDGE | PLANE | BLOB.
struct MatchWorkspace {
  DevBuf<int32_t> nn_idx[3], nn_found[3];   // per type: the three types are matched on their own streams
  DevBuf<float> nn_d2[3];
  DevBuf<double> records;           // 13 planes x rec_stride (structure of arrays)
  size_t rec_stride = 0;
  DevBuf<uint8_t> status;
  DevBuf<double> weight;
  DevBuf<int> hist;                 // [ICP iterations of the frame][3][LSB_MATCH_N_STATUS]
  static constexpr int kHistBlock = 3 * LSB_MATCH_N_STATUS;
  int hist_iters = 1;
  int icp_stage = ICP_STAGE_LOC;    // which pose of the device-resident loop moves the queries (match_type)
  DevBuf<double> part_a;            // LM chunk partial sums, double-buffered
  DevBuf<LmState> lm_state;         // [1] final state of the last solve
  DevBuf<unsigned> lm_counter;      // [2] grid barrier of the persistent LM kernel
  int lm_grid_limit = 0;            // co-resident CTAs of that kernel on this device
  int lm_grid_share = 1;            // handles in flight on this device: a solve takes lm_grid_limit / share CTAs, so that the
                                    // cooperative launches of several sequences run side by side instead of one after the other
  PinnedBuf<int> h_hist;
  PinnedBuf<LmState> h_lm;
  // result block of a solve in pinned (device-visible) host memory: k_lm_solve stores the final state and the match
  // histograms there itself and raises `seq` last; the host polls `seq` instead of copying and synchronising
  struct LmHostBlock { LmState state; int hist[3 * LSB_MATCH_N_STATUS]; IcpState icp; unsigned seq; };
  PinnedBuf<LmHostBlock> h_block;
  unsigned solve_seq = 0;
  KernelTimer t_knn, t_fit, t_lm;   // lsb_set_profiling
  size_t d2h_bytes = 0;
  size_t total = 0;                 // records of the current iteration
  size_t off[4] = {0, 0, 0, 0};     // offsets of the three types

  // buffers and zeroed match histograms for up to n_iters ICP iterations on nk keypoints per type
  void begin_frame(const size_t nk[3], int n_iters, cudaStream_t s);
  void begin_iteration(const size_t nk[3], cudaStream_t s) { begin_frame(nk, 1, s); }
  // the reuse search kernel visits every keypoint exactly once only when it runs at all (non-empty target, reuse enabled)
  static bool can_fuse_refinement(const KnnGrid& target, size_t nk) { return nk > 0 && !target.empty() && knn_reuse_enabled(); }
  // KeypointsMatcher::BuildMatchResiduals for one keypoint type (kNN + model fit) in ICP iteration icp_iter.
  // icp != nullptr: device-resident loop - the queries are moved by the pose held in *icp (`prior` is ignored) and the
  // kernels return at once when icp->done is set.  reuse: the neighbour lists of the previous iteration (same keypoints,
  // same target) are still in the workspace and bound the search.  refine (with reuse): the undistortion refinement the previous
  // iteration left in icp->Pending is applied to the keypoints inside the search kernel (else the caller applies it first).
  void match_type(int type, const lsb_point* d_kp, size_t nk, const KnnGrid& target, const Iso3d& prior, const lsb_match_params& mp,
                  cudaStream_t s, LaunchCounter& lc, int icp_iter = 0, const IcpState* icp = nullptr, bool reuse = false, bool refine = false);
  // reads the status histograms back (one stream sync); hist_out[3][LSB_MATCH_N_STATUS]
  void read_histograms(int hist_out[3][LSB_MATCH_N_STATUS], cudaStream_t s);
  // LocalOptimizer::Solve on the current records; returns the final state and the match histograms
  // with ONE launch and ONE stream sync.  The "enough matches" test of the caller (Slam.cxx:919,
  // :1098) is taken on the device too: below min_matches the kernel exits at once and the state
  // comes back with term == LSB_LM_SKIPPED (x == x0).
  LmState solve(const double x0[6], int max_iter, int two_d, double saturation, unsigned min_matches,
                int hist_out[3][LSB_MATCH_N_STATUS], cudaStream_t s, LaunchCounter& lc, const LmSensorTerms* sensor_terms = nullptr);
  // The same in two halves for the device-resident ICP loop: every iteration's solve is enqueued with the sequence number
  // taken once per loop (next_solve_seq); the kernel whose icp_control ends the loop hands LmHostBlock over and raises the
  // flag wait_result polls.  icp == nullptr: x0 is the start and the kernel hands over unconditionally.
  unsigned next_solve_seq();
  void enqueue_solve(const double x0[6], int max_iter, int two_d, double saturation, unsigned min_matches, cudaStream_t s, LaunchCounter& lc,
                     const LmSensorTerms* sensor_terms, IcpState* icp, const IcpCtl* ctl);
  const LmHostBlock& wait_result(cudaStream_t s);
};

}  // namespace lsb

// K9 model fit + K10/K11 device LM — see match.cuh.
// Per keypoint (one thread): neighbourhood selection (per-ring pick or inlier-line
// vote, KeypointsMatcher.cxx:349-480), 3x3 FP64 covariance + closed-form eigen
// (common/pca.h), model checks and the residual record (A, P, X, w)
// (KeypointsMatcher.cxx:106-346).  Plain IEEE arithmetic, no FMA contraction.
#include "match.cuh"
#include <cstring>
#include <atomic>
#include "../common/pca.h"
#include <cfloat>
#include <cmath>

namespace lsb {

namespace {

constexpr int FIT_TPB = 128;

struct FitParams {
  int single_edge_per_ring;
  double sq_max_neighbors_distance;
  unsigned edge_nb, edge_min_nb;
  double edge_max_model_error, sq_edge_max_model_error;
  float sq_edge_inlier;          // (float)(EdgeMaxModelError^2)  (KeypointsMatcher.cxx:423)
  unsigned plane_nb;
  double planarity_threshold, plane_max_model_error, sq_plane_max_model_error;
  unsigned blob_nb;
};

__device__ __forceinline__ Vec3f cloud_xyz(const lsb_point* __restrict__ cloud, int i)
{
  const float4 p = __ldg(reinterpret_cast<const float4*>(cloud) + 2 * (size_t)i);
  return Vec3f(p.x, p.y, p.z);
}
__device__ __forceinline__ int cloud_ring(const lsb_point* __restrict__ cloud, int i)
{
  const uint4 h = __ldg(reinterpret_cast<const uint4*>(cloud) + 2 * (size_t)i + 1);
  return (int)(h.w & 0xffffu);
}

__device__ __forceinline__ bool dev_isfinite(double v) { return (v - v) == 0.0; }

// Residual records live in a structure-of-arrays layout (13 planes of doubles, plane f of
// record i at rec[f * stride + i]: M 0..5, P 6..8, X 9..11, w 12) so that the per-thread accesses
// of the fit and LM kernels coalesce.
constexpr int REC_PLANES = 13;
constexpr int REC_W = 12;   // plane of the weight (w < 0: no residual)
__device__ __forceinline__ void write_record(double* __restrict__ rec, size_t stride, const Mat3d& A, const Vec3d& mean, const Vec3d& X, double w)
{
  LmRecord r;
  const double P[3] = {mean.x, mean.y, mean.z}, Xa[3] = {X.x, X.y, X.z};
  lm_record_set(r, A.m, P, Xa, w);
#pragma unroll
  for (int t = 0; t < 6; ++t) rec[t * stride] = r.M[t];
#pragma unroll
  for (int t = 0; t < 3; ++t) { rec[(6 + t) * stride] = r.P[t]; rec[(9 + t) * stride] = r.X[t]; }
  rec[REC_W * stride] = w;
}
__device__ __forceinline__ LmRecord load_record(const double* __restrict__ rec, size_t stride)
{
  LmRecord r;
  r.w = rec[REC_W * stride];
  if (r.w >= 0.0) {
#pragma unroll
    for (int t = 0; t < 6; ++t) r.M[t] = rec[t * stride];
#pragma unroll
    for (int t = 0; t < 3; ++t) { r.P[t] = rec[(6 + t) * stride]; r.X[t] = rec[(9 + t) * stride]; }
  }
  return r;
}

// One thread per keypoint.  sel[] holds positions in the kNN list, in list order.
template <int TYPE>
__global__ void __launch_bounds__(FIT_TPB)
k_fit(const lsb_point* __restrict__ kp, int nq, const lsb_point* __restrict__ cloud, const int32_t* __restrict__ nn_idx,
      const float* __restrict__ nn_d2, const int32_t* __restrict__ nn_found, int k, FitParams fp, double* __restrict__ records, size_t rec_stride,
      uint8_t* __restrict__ status, double* __restrict__ weight, int* __restrict__ hist, const int* __restrict__ done,
      unsigned* __restrict__ knn_counters)
{
  if (done && *done) return;   // device-resident ICP loop: the loop ended in an earlier iteration (cuda/icp.cuh)
  // the search that fed this kernel is over: its work counters (KnnGrid::unres_count) go back to zero here, which saves the next
  // search of the grid a memset at the head of its chain
  if (knn_counters && blockIdx.x == 0 && threadIdx.x < 4) knn_counters[threadIdx.x] = 0u;
  __shared__ int s_hist[LSB_MATCH_N_STATUS];
  if (threadIdx.x < LSB_MATCH_N_STATUS) s_hist[threadIdx.x] = 0;
  __syncthreads();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nq) {
    int st = LSB_MATCH_SUCCESS;
    double w = 0.0;
    // neighbour lists are stored entry-major (entry t of query i at [t * nq + i]): coalesced
    int32_t idx[KNN_MAX_K];
    float d2[KNN_MAX_K];
    const int n = nn_found[i];
    for (int t = 0; t < k; ++t) { idx[t] = nn_idx[(size_t)t * nq + i]; d2[t] = nn_d2[(size_t)t * nq + i]; }
    int sel[KNN_MAX_K];
    int nsel = 0;
    if (TYPE == LSB_EDGE) {
      if (fp.edge_nb < 2 || fp.edge_min_nb < 2) st = LSB_MATCH_BAD_MODEL_PARAMETRIZATION;
      else if (fp.single_edge_per_ring) {
        // GetPerRingLineNeighbors (:349-405): the nearest's ring and rings farther than 4
        // from it are excluded; then the first hit of each remaining ring, in list order.
        if (n > 0) {
          const int closest = cloud_ring(cloud, idx[0]);
          int rings[KNN_MAX_K];
          for (int t = 0; t < n; ++t) rings[t] = cloud_ring(cloud, idx[t]);
          for (int t = 0; t < n; ++t) {
            const int r = rings[t];
            if (r == closest || abs(closest - r) > 4) continue;
            bool seen = false;
            for (int u = 0; u < nsel; ++u) seen |= (rings[sel[u]] == r);
            if (!seen) sel[nsel++] = t;
          }
        }
      } else {
        // GetRansacLineNeighbors (:408-480)
        if (n >= 2) {
          Vec3f nb[KNN_MAX_K];
          for (int t = 0; t < n; ++t) nb[t] = cloud_xyz(cloud, idx[t]);
          const Vec3f P1 = nb[0];
          unsigned bestMask = 0;
          int bestCount = 0;
          for (int pt = 1; pt < n; ++pt) {
            const Vec3f dir = normalized(nb[pt] - P1);
            unsigned mask = 0;
            for (int cd = 1; cd < n; ++cd) {
              if (cd == pt) mask |= 1u << cd;
              else if (sqnorm(cross(nb[cd] - P1, dir)) < fp.sq_edge_inlier) mask |= 1u << cd;
            }
            const int cnt = __popc(mask);
            if (cnt > bestCount) { bestCount = cnt; bestMask = mask; }
          }
          sel[nsel++] = 0;
          for (int cd = 1; cd < n; ++cd)
            if (bestMask & (1u << cd)) sel[nsel++] = cd;
        }
      }
      if (st == LSB_MATCH_SUCCESS && (unsigned)nsel < fp.edge_min_nb) st = LSB_MATCH_NOT_ENOUGH_NEIGHBORS;
    } else {
      const unsigned need = (TYPE == LSB_PLANE) ? fp.plane_nb : fp.blob_nb;
      if (need < ((TYPE == LSB_PLANE) ? 3u : 4u)) st = LSB_MATCH_BAD_MODEL_PARAMETRIZATION;
      else if ((unsigned)n < need) st = LSB_MATCH_NOT_ENOUGH_NEIGHBORS;
      else { nsel = n; for (int t = 0; t < n; ++t) sel[t] = t; }
    }
    if (st == LSB_MATCH_SUCCESS && (double)d2[sel[nsel - 1]] > fp.sq_max_neighbors_distance) st = LSB_MATCH_NEIGHBORS_TOO_FAR;
    if (st == LSB_MATCH_SUCCESS) {
      // Utils::ComputeMeanAndPCA in double over the selected neighbours, in order
      CovAccum<double> acc;
      for (int t = 0; t < nsel; ++t) { const Vec3f p = cloud_xyz(cloud, idx[sel[t]]); acc.add(p.x, p.y, p.z); }
      Vec3d mean;
      Mat3d cov, V;
      double ev[3];
      acc.finish(mean, cov);
      eigen33(cov, V, ev);
      const float4 q = reinterpret_cast<const float4*>(kp)[2 * (size_t)i];
      const Vec3d base((double)q.x, (double)q.y, (double)q.z);
      Mat3d A;
      if (TYPE == LSB_EDGE) {
        const Vec3d nn = V.col(2);
        for (int r = 0; r < 3; ++r)
          for (int c = 0; c < 3; ++c) A(r, c) = ((r == c) ? 1.0 : 0.0) - nn[r] * nn[c];
        const double mse = ev[0] + ev[1];
        if (!dev_isfinite(A(0, 0))) st = LSB_MATCH_INVALID_NUMERICAL;
        else if (mse >= fp.sq_edge_max_model_error) st = LSB_MATCH_MSE_TOO_LARGE;
        else w = (mse <= 1e-6) ? 1. : 1. - ::sqrt(mse) / fp.edge_max_model_error;
      } else if (TYPE == LSB_PLANE) {
        if (ev[1] / ev[2] < fp.planarity_threshold) st = LSB_MATCH_BAD_PCA_STRUCTURE;
        else {
          const Vec3d nn = V.col(0);
          for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c) A(r, c) = nn[r] * nn[c];
          const double mse = ev[0];
          if (!dev_isfinite(A(0, 0))) st = LSB_MATCH_INVALID_NUMERICAL;
          else if (mse >= fp.sq_plane_max_model_error) st = LSB_MATCH_MSE_TOO_LARGE;
          else w = (mse <= 1e-6) ? 1. : 1. - ::sqrt(mse) / fp.plane_max_model_error;
        }
      } else {
        if (ev[0] <= 0. || ev[1] <= 0.) st = LSB_MATCH_BAD_PCA_STRUCTURE;
        else {
          const double s0 = 1.0 / ::sqrt(ev[0]), s1 = 1.0 / ::sqrt(ev[1]), s2 = 1.0 / ::sqrt(ev[2]);
          for (int r = 0; r < 3; ++r)
            for (int c = 0; c < 3; ++c)
              A(r, c) = sum3((V(r, 0) * s0) * V(c, 0), (V(r, 1) * s1) * V(c, 1), (V(r, 2) * s2) * V(c, 2));
          if (!dev_isfinite(A(0, 0)) || !dev_isfinite(s0 * (s1 * s2))) st = LSB_MATCH_INVALID_NUMERICAL;
          else w = 1.0;
        }
      }
      if (st == LSB_MATCH_SUCCESS) write_record(records + i, rec_stride, A, mean, base, w);
    }
    if (st != LSB_MATCH_SUCCESS) { w = 0.0; records[REC_W * rec_stride + i] = -1.0; }
    status[i] = (uint8_t)st;
    weight[i] = w;
    atomicAdd(&s_hist[st], 1);
  }
  __syncthreads();
  if (threadIdx.x < LSB_MATCH_N_STATUS && s_hist[threadIdx.x]) atomicAdd(hist + threadIdx.x, s_hist[threadIdx.x]);
}

// target empty: MatchingResults::Reset values (KeypointsMatcher.h:108-118)
__global__ void k_fit_empty(int nq, double* __restrict__ records, size_t rec_stride, uint8_t* __restrict__ status, double* __restrict__ weight,
                            const int* __restrict__ done)
{
  if (done && *done) return;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq) return;
  records[REC_W * rec_stride + i] = -1.0;
  status[i] = LSB_MATCH_UNKNOWN;
  weight[i] = 0.0;
}

// ---------------------------------------------------------------- K10/K11 ----
// lm_propose (common/lm_core.h) spread over the lanes of one warp: lane r < 6 owns row r of the
// scaled normal matrix, of its Cholesky factor and of the right-hand side; pivots, columns and
// the solution travel by shuffles.  Every floating-point operation of the sequential function is
// performed once, on the same operands and in the same order (see the canonical order stated at
// lm_chol_solve6), so the CPU oracle and this kernel take bit-identical steps; what changes is
// that independent rows advance together and nothing spills (the sequential form needs three
// 6x6 FP64 matrices per thread).  `s` lives in shared memory; all 32 lanes must call.
__device__ __forceinline__ void lm_propose_warp(LmState& s)
{
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int r = lane < 6 ? lane : 5;            // lanes >= 6 shadow row 5 (their results are never read)
  int iter = s.iter, num_invalid = s.num_invalid;
  const int max_iter = s.max_iter;
  const bool two_d = s.two_d != 0;
  double radius = s.radius, dec = s.decrease_factor;
  const bool free_r = !(two_d && (r == 2 || r == 3 || r == 4));
  const double Sr = s.S[r];
  double Hs[6];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const double h = (r <= j) ? s.H[lm_hidx(r, j)] : s.H[lm_hidx(j, r)];
    const bool free_j = !(two_d && (j == 2 || j == 3 || j == 4));
    Hs[j] = (free_r && free_j) ? h * Sr * s.S[j] : 0.0;
  }
  const double gs = free_r ? s.g[r] * Sr : 0.0;
  double hdiag = Hs[0];
#pragma unroll
  for (int c = 1; c < 6; ++c) hdiag = (r == c) ? Hs[c] : hdiag;
  const double xr = s.x[r];
  int done = 0, term = LSB_LM_RUNNING;
  double mcc = 0.0, step_norm = 0.0, xcr = xr;
  bool stepped = false;
  while (true) {
    if (iter >= max_iter) { done = 1; term = LSB_LM_MAX_ITER; break; }
    if (radius < 1e-32) { done = 1; term = LSB_LM_CONV_RADIUS; break; }
    iter += 1;
    double mdiag = 1.0;                        // held parameter: decoupled dummy row -> delta = 0
    if (free_r) {
      double d = hdiag;
      d = (d < 1e-6) ? 1e-6 : d;
      d = (d > 1e32) ? 1e32 : d;
      mdiag = hdiag + d / radius;
    }
    double Mr[6], Lr[6], inv[6];
#pragma unroll
    for (int c = 0; c < 6; ++c) Mr[c] = (c == r) ? mdiag : ((c < r) ? Hs[c] : 0.0);
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double d = __shfl_sync(FULL, Mr[k], k);
      if (!(d > 0.0)) ok = false;
      const double ljj = ::sqrt(d);
      inv[k] = 1.0 / ljj;
      const double Lik = (r == k) ? ljj : ((r > k) ? Mr[k] * inv[k] : 0.0);
      Lr[k] = Lik;
#pragma unroll
      for (int j = k + 1; j < 6; ++j) {
        const double Ljk = __shfl_sync(FULL, Lik, j);
        if (j <= r) Mr[j] = Mr[j] - Lik * Ljk;
      }
    }
    double db[6], db_r = 0.0;
    if (ok) {
      double sv = -gs, z_r = 0.0;
#pragma unroll
      for (int k = 0; k < 6; ++k) {                       // forward substitution
        const double zk = __shfl_sync(FULL, sv * inv[k], k);
        if (r == k) z_r = zk;
        if (r > k) sv = sv - Lr[k] * zk;
      }
      double Lc[6];                                        // column r of L: Lc[k] = L[k][r], k > r
#pragma unroll
      for (int k = 1; k < 6; ++k) {
        double v = 0.0;
#pragma unroll
        for (int c = 0; c < k; ++c) {
          const double t = __shfl_sync(FULL, Lr[c], k);
          if (r == c) v = t;
        }
        Lc[k] = v;
      }
      sv = z_r;
#pragma unroll
      for (int k = 5; k >= 0; --k) {                       // backward substitution, k descending
        const double yk = __shfl_sync(FULL, sv * inv[k], k);
        db[k] = yk;
        if (r == k) db_r = yk;
        if (r < k) sv = sv - Lc[k] * yk;
      }
#pragma unroll
      for (int i = 0; i < 6; ++i)
        if (!(db[i] - db[i] == 0.0)) ok = false;           // NaN / inf
    }
    mcc = 0.0;
    if (ok) {
      // model_cost_change = -delta^T (g + 0.5 H delta)   (scaled quantities)
      double hd = 0.0;
#pragma unroll
      for (int j = 0; j < 6; ++j) hd += Hs[j] * db[j];
      const double t = db_r * (gs + 0.5 * hd);
#pragma unroll
      for (int i = 0; i < 6; ++i) mcc -= __shfl_sync(FULL, t, i);
      ok = (mcc > 0.0);
    }
    if (!ok) {
      num_invalid += 1;
      if (num_invalid >= 5) { done = 1; term = LSB_LM_FAIL_INVALID; break; }
      radius = radius * 0.5;              // StepIsInvalid (see lm_propose)
      continue;
    }
    num_invalid = 0;
    const double dstep = db_r * Sr;
    xcr = xr + dstep;
    const double e = xr - xcr;
    const double ee = e * e;
    double sn = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i) sn += __shfl_sync(FULL, ee, i);
    step_norm = ::sqrt(sn);
    stepped = true;
    break;
  }
  __syncwarp();
  if (stepped && lane < 6) s.xc[lane] = xcr;
  if (lane == 0) {
    s.iter = iter;
    s.num_invalid = num_invalid;
    s.radius = radius;
    s.decrease_factor = dec;
    if (stepped) { s.model_cost_change = mcc; s.step_norm = step_norm; }
    if (done) { s.done = 1; s.term = term; }
  }
  __syncwarp();
}

// rotation at the evaluation point x (LmPoint): the six sin/cos on six lanes, assembled by lane 0
__device__ __noinline__ void lm_point_warp(const double* x, unsigned long long* s_pt_raw)
{
  const int lane = threadIdx.x & 31;
  double xl[6];
#pragma unroll
  for (int q = 0; q < 6; ++q) xl[q] = x[q];
  const double tv = (lane < 6) ? lm_point_trig(xl, lane) : 0.0;
  double trig[6];
#pragma unroll
  for (int q = 0; q < 6; ++q) trig[q] = __shfl_sync(0xffffffffu, tv, q);
  if (lane == 0) lm_make_point_from_trig(xl, trig, *reinterpret_cast<LmPoint*>(s_pt_raw));
}

// tools/probes/lm_eval_probe.cu defines LSB_LM_PROBE to time the phases of one evaluation (CTA 0)
#ifdef LSB_LM_PROBE
__device__ unsigned long long g_lm_probe[64 * 8];
#define LM_PROBE_STAMP(e, k) do { if (blockIdx.x == 0 && threadIdx.x == 0 && (e) < 64) { unsigned long long t_; \
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_lm_probe[(e) * 8 + (k)] = t_; } } while (0)
#else
#define LM_PROBE_STAMP(e, k) do { } while (0)
#endif

// The trust-region step, run by warp 0 of every CTA on the CTA's own copy of the state (shared
// memory) and kept out of line so that its registers do not weigh on the evaluating threads.
// Every CTA holds the same state and the same 28 sums, and every operation below is
// deterministic, so the copies stay bit-identical without any exchange.  Lane 0 takes the accept /
// reject decision (scalar), the warp proposes the next step, and the six sin/cos of the next
// evaluation point are computed by six lanes in parallel.
__device__ __noinline__ void lm_control(unsigned long long* s_state_raw, unsigned long long* s_pt_raw, const double* sE, int first, int e = 0)
{
  constexpr unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  LmState& s = *reinterpret_cast<LmState*>(s_state_raw);
  int propose = 0;
  if (lane == 0) propose = (first ? lm_first_update(s, sE) : lm_next_update(s, sE)) ? 1 : 0;
  __syncwarp();
  propose = __shfl_sync(FULL, propose, 0);
  LM_PROBE_STAMP(e, 5);
  if (propose) lm_propose_warp(s);
  __syncwarp();
  LM_PROBE_STAMP(e, 6);
  if (s.done) return;
  lm_point_warp(s.xc, s_pt_raw);
}

// Device-resident ICP loop (cuda/icp.cuh): the steps between two ICP iterations, out of line so that their FP64 pose
// arithmetic (quaternions, slerp, trig) does not weigh on the registers of the evaluating threads (inlined they cost
// k_lm_solve 2.7 KB of spills).  One thread; the state lives in shared memory, staged by the whole CTA.
__device__ __noinline__ void icp_control_dev(unsigned long long* s_icp_raw, IcpCtl ctl, const unsigned long long* s_state_raw, unsigned matched)
{
  icp_control(*reinterpret_cast<IcpState*>(s_icp_raw), ctl, *reinterpret_cast<const LmState*>(s_state_raw), matched);
}
// LocalOptimizer::SetPosePrior (:44-48) on the pose the previous iteration left in device memory
__device__ __noinline__ void icp_prior_dev(const IcpState* icp, int stage, double* x0)
{
  iso_to_xyzrpy(stage == ICP_STAGE_LOC ? icp->Tworld : icp->Trelative, x0);
}

constexpr int LM_THREADS = 512;   // 16 warps; each warp evaluates 2 groups of 32 records of the CTA's chunk
constexpr int LM_WARPS = LM_THREADS / 32;
constexpr int LM_GROUPS = LSB_LM_CHUNK / 32;
static_assert(LM_GROUPS == 32, "a chunk is a fold32 of 32 groups");

// fold32 (lm_core.h) of 28 per-lane values at once: instead of 28 butterflies (140 shuffles and
// adds) the lanes split the quantities as they fold (14 + 7 + 4 + 2 + 1 = 28 shuffles and adds).
// After step `off` the lanes whose bit `off` is clear hold the first half of the quantities, the
// others the second half; each partial sum is v[i] + v[i + off] exactly as in the butterfly.
// On return lane L holds the complete fold32 of quantity q(L) (q < 0: padding).
__device__ __forceinline__ double warp_fold28(const double (&v)[LSB_LM_NSUM], int lane, int& q)
{
  constexpr unsigned FULL = 0xffffffffu;
  double a1[14], a2[7], a3[4], a4[2];
  const bool b16 = lane & 16, b8 = lane & 8, b4 = lane & 4, b2 = lane & 2, b1 = lane & 1;
#pragma unroll
  for (int i = 0; i < 14; ++i) {
    const double keep = b16 ? v[i + 14] : v[i], send = b16 ? v[i] : v[i + 14];
    a1[i] = keep + __shfl_xor_sync(FULL, send, 16);
  }
#pragma unroll
  for (int i = 0; i < 7; ++i) {
    const double keep = b8 ? a1[i + 7] : a1[i], send = b8 ? a1[i] : a1[i + 7];
    a2[i] = keep + __shfl_xor_sync(FULL, send, 8);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const double hi = (i + 4 < 7) ? a2[(i + 4 < 7) ? i + 4 : 0] : 0.0;
    const double keep = b4 ? hi : a2[i], send = b4 ? a2[i] : hi;
    a3[i] = keep + __shfl_xor_sync(FULL, send, 4);
  }
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const double keep = b2 ? a3[i + 2] : a3[i], send = b2 ? a3[i] : a3[i + 2];
    a4[i] = keep + __shfl_xor_sync(FULL, send, 2);
  }
  const double keep = b1 ? a4[1] : a4[0], send = b1 ? a4[0] : a4[1];
  const double r = keep + __shfl_xor_sync(FULL, send, 1);
  const int l7 = lane & 7;
  q = (l7 < 7) ? (b16 ? 14 : 0) + (b8 ? 7 : 0) + l7 : -1;
  return r;
}

__device__ __forceinline__ double warp_fold32(double v)
{
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
  return v;   // complete in lane 0
}

// Arrive-and-wait over the whole (cooperatively launched, hence co-resident) grid.  `bar` only
// grows during a solve: the e-th barrier completes when it reaches (e + 1) * gridDim.x.
__device__ __forceinline__ void lm_grid_barrier(unsigned* bar, unsigned target)
{
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    unsigned v;
    do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(bar) : "memory"); } while (v < target);
  }
  __syncthreads();
}

struct LmStart { double x[6]; };

// LocalOptimizer::Solve as ONE persistent kernel: the whole trust-region loop (common/lm_core.h)
// with a grid barrier per cost/Jacobian evaluation instead of a launch.  CTA b owns the chunks
// b, b + gridDim.x, ... of 1024 records; the first one is staged in shared memory once (13 planes,
// 104 KB) and re-read from there at every evaluation.  One evaluation, in the canonical radix-32
// reduction order: a warp folds the 28 sums of a group of 32 records, the CTA folds the 32 groups
// of a chunk and publishes the chunk sums; after the barrier EVERY CTA folds the chunk sums in
// blocks of 32 (one warp per quantity), adds the block sums in order and advances its own copy of
// the LM state (lm_control), so no second barrier and no state round trip through memory is needed.
// part: [2][LSB_LM_NSUM][pstride] chunk sums (double-buffered by evaluation parity: a CTA may start
// the next evaluation while a slower one still reads the previous sums), pstride = 32 * blocks.
// bar: [2] = arrival counter, exit counter; both are zero again when the grid has exited.
__global__ void __launch_bounds__(LM_THREADS, 1)
k_lm_solve(const double* __restrict__ records, size_t rec_stride, int n, int nchunks, int pstride, double sat, LmStart start, int max_iter,
           int two_d, const int* __restrict__ hist, unsigned min_matches, LmState* __restrict__ state_out, double* __restrict__ part,
           unsigned* __restrict__ bar, const LmSensorTerms terms, MatchWorkspace::LmHostBlock* host_out, unsigned host_seq,
           IcpState* __restrict__ icp, const IcpCtl ctl)
{
  constexpr int NW = (int)(sizeof(LmState) / 8), NP = (int)(sizeof(LmPoint) / 8);
  static_assert(sizeof(LmState) % 8 == 0 && sizeof(LmPoint) % 8 == 0, "copied as 64-bit words");
  extern __shared__ __align__(16) double s_rec[];                 // [REC_PLANES][LSB_LM_CHUNK]: the CTA's first chunk
  __shared__ __align__(16) unsigned long long s_pt_raw[NP];
  __shared__ __align__(16) unsigned long long s_state_raw[NW];
  __shared__ double s_group[LSB_LM_NSUM][LM_GROUPS];
  __shared__ double s_E[LSB_LM_NSUM];
  LmState& s_state = *reinterpret_cast<LmState*>(s_state_raw);
  const LmPoint& s_pt = *reinterpret_cast<const LmPoint*>(s_pt_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ int s_off;
  __shared__ unsigned s_matched;

  // Device-resident ICP loop: when an earlier iteration ended the loop this launch does nothing.  The flag is read by one
  // thread per CTA (CTA 0 raises it at its tail; a CTA still in its prologue then can only be one whose solve never
  // starts, see below), and a CTA that stands aside still takes part in the exit count that re-arms the barrier.
  if (threadIdx.x == 0) s_off = (icp && icp->done) ? 1 : 0;
  __syncthreads();
  if (s_off) {
    if (threadIdx.x == 0) {
      const unsigned t = atomicAdd(bar + 1, 1u);
      if (t == gridDim.x - 1) { bar[0] = 0u; bar[1] = 0u; __threadfence(); }
    }
    return;
  }

  // prologue: initial state (lm_begin + the "enough matches" gate of the caller) and first chunk
  if (warp == 0) {
    if (lane == 0) {
      double x0[6];
      if (icp) icp_prior_dev(icp, ctl.stage, x0);
      else for (int q = 0; q < 6; ++q) x0[q] = start.x[q];
      lm_begin(s_state, x0, max_iter, two_d);
      unsigned matched = 0;
      for (int t = 0; t < 3; ++t) matched += (unsigned)hist[t * LSB_MATCH_N_STATUS + LSB_MATCH_SUCCESS];
      s_matched = matched;
      if (matched < min_matches) { s_state.done = 1; s_state.term = LSB_LM_SKIPPED; }
    }
    __syncwarp();
    lm_point_warp(s_state.x, s_pt_raw);
  }
  {
    const int base = blockIdx.x * LSB_LM_CHUNK;
    for (int j = threadIdx.x; j < REC_PLANES * LSB_LM_CHUNK; j += LM_THREADS) {
      const int q = j / LSB_LM_CHUNK, r = j % LSB_LM_CHUNK;
      const int i = base + r;
      s_rec[j] = (i < n) ? records[(size_t)q * rec_stride + i] : ((q == REC_W) ? -1.0 : 0.0);
    }
  }
  __syncthreads();

  const int nblocks = pstride >> 5;
  for (int e = 0; !s_state.done; ++e) {           // uniform over the grid: every CTA holds the same state
    double* my_part = part + (size_t)(e & 1) * LSB_LM_NSUM * pstride;
    LM_PROBE_STAMP(e, 0);
    for (int c = blockIdx.x; c < nchunks; c += gridDim.x) {
      // one code path for both sources (generic loads): the chunk staged in shared memory, or global
      const bool cached = (c == (int)blockIdx.x);
      const double* src = cached ? s_rec : records + (size_t)c * LSB_LM_CHUNK;
      const size_t stride = cached ? (size_t)LSB_LM_CHUNK : rec_stride;
      const int left = cached ? LSB_LM_CHUNK : n - c * LSB_LM_CHUNK;     // records of this chunk that exist
#pragma unroll 1
      for (int g = warp; g < LM_GROUPS; g += LM_WARPS) {
        double out[LSB_LM_NSUM];
        const int j = g * 32 + lane;
        if (j < left) lm_eval_record(load_record(src + j, stride), s_pt, sat, out);
        else {
#pragma unroll
          for (int q = 0; q < LSB_LM_NSUM; ++q) out[q] = 0.0;
        }
        int q;
        const double r = warp_fold28(out, lane, q);
        if (q >= 0) s_group[q][g] = r;
      }
      __syncthreads();
      {
        const int q1 = warp + LM_WARPS;
        double v0 = s_group[warp][lane], v1 = (q1 < LSB_LM_NSUM) ? s_group[q1 < LSB_LM_NSUM ? q1 : 0][lane] : 0.0;
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
          v0 += __shfl_down_sync(0xffffffffu, v0, off);
          v1 += __shfl_down_sync(0xffffffffu, v1, off);
        }
        if (lane == 0) {
          my_part[(size_t)warp * pstride + c] = v0;
          if (q1 < LSB_LM_NSUM) my_part[(size_t)q1 * pstride + c] = v1;
        }
      }
      __syncthreads();
    }
    LM_PROBE_STAMP(e, 1);
    lm_grid_barrier(bar, (unsigned)(e + 1) * gridDim.x);
    LM_PROBE_STAMP(e, 2);
    {
      // warp w folds the quantities w and w + 16: the (up to) eight block folds of a pass are independent chains, loaded
      // together and folded in lock step (same pairwise tree and the same in-order sum of the block sums as before)
      constexpr unsigned FULL = 0xffffffffu;
      double acc[2] = {0.0, 0.0};
      for (int b0 = 0; b0 < nblocks; b0 += 4) {
        double v[2][4];
#pragma unroll
        for (int qi = 0; qi < 2; ++qi) {
          const int q = warp + LM_WARPS * qi;
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int c = (b0 + u) * 32 + lane;
            v[qi][u] = (q < LSB_LM_NSUM && b0 + u < nblocks && c < nchunks) ? __ldcg(my_part + (size_t)q * pstride + c) : 0.0;
          }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
#pragma unroll
          for (int qi = 0; qi < 2; ++qi) {
#pragma unroll
            for (int u = 0; u < 4; ++u) v[qi][u] += __shfl_down_sync(FULL, v[qi][u], off);
          }
        }
#pragma unroll
        for (int qi = 0; qi < 2; ++qi) {
#pragma unroll
          for (int u = 0; u < 4; ++u)
            if (b0 + u < nblocks) acc[qi] = (b0 + u == 0) ? v[qi][u] : acc[qi] + v[qi][u];
        }
      }
      if (lane == 0) {
        s_E[warp] = acc[0];
        if (warp + LM_WARPS < LSB_LM_NSUM) s_E[warp + LM_WARPS] = acc[1];
      }
    }
    __syncthreads();
    LM_PROBE_STAMP(e, 3);
    if (warp == 0) {
      // wheel-odometer / IMU gravity blocks (Slam.cxx:1123-1131): added to the reduced sums at the evaluation point
      if ((terms.use_wheel | terms.use_gravity) && lane == 0) lm_add_sensor_terms(terms, s_pt, s_E);
      __syncwarp();
      lm_control(s_state_raw, s_pt_raw, s_E, e == 0, e);
    }
    __syncthreads();
    LM_PROBE_STAMP(e, 4);
  }
  if (blockIdx.x == 0) {
    if (threadIdx.x < NW) reinterpret_cast<unsigned long long*>(state_out)[threadIdx.x] = s_state_raw[threadIdx.x];
    // between two ICP iterations: pose update, undistortion refinement, the break tests (cuda/icp.cuh), by one thread on a
    // copy of the state in shared memory (s_group is free by now: every evaluation is over)
    constexpr int NI = (int)(sizeof(IcpState) / 8);
    static_assert(sizeof(IcpState) % 8 == 0, "copied as 64-bit words");
    static_assert(NI <= LSB_LM_NSUM * LM_GROUPS, "the ICP state is staged in s_group");
    unsigned long long* s_icp_raw = reinterpret_cast<unsigned long long*>(&s_group[0][0]);
    __shared__ int s_hand_over;
    if (icp) for (int j = threadIdx.x; j < NI; j += LM_THREADS) s_icp_raw[j] = reinterpret_cast<const unsigned long long*>(icp)[j];
    __syncthreads();
    if (threadIdx.x == 0) {
      s_hand_over = 1;
      if (icp) {
        icp_control_dev(s_icp_raw, ctl, s_state_raw, s_matched);
        s_hand_over = reinterpret_cast<const IcpState*>(s_icp_raw)->done;
      }
    }
    __syncthreads();
    if (icp) for (int j = threadIdx.x; j < NI; j += LM_THREADS) reinterpret_cast<unsigned long long*>(icp)[j] = s_icp_raw[j];
    if (host_out && s_hand_over) {     // the loop is over: CTA 0 hands the results to the host, then raises the flag
      if (threadIdx.x < NW) reinterpret_cast<volatile unsigned long long*>(&host_out->state)[threadIdx.x] = s_state_raw[threadIdx.x];
      if (threadIdx.x < 3 * LSB_MATCH_N_STATUS) reinterpret_cast<volatile int*>(host_out->hist)[threadIdx.x] = hist[threadIdx.x];
      if (icp) for (int j = threadIdx.x; j < NI; j += LM_THREADS) reinterpret_cast<volatile unsigned long long*>(&host_out->icp)[j] = s_icp_raw[j];
      __threadfence_system();
      __syncthreads();
      if (threadIdx.x == 0) { *reinterpret_cast<volatile unsigned*>(&host_out->seq) = host_seq; __threadfence_system(); }
    }
  }
  // the last CTA to leave re-arms the barrier for the next solve (every CTA is past its last wait)
  if (threadIdx.x == 0) {
    const unsigned t = atomicAdd(bar + 1, 1u);
    if (t == gridDim.x - 1) { bar[0] = 0u; bar[1] = 0u; __threadfence(); }
  }
}

}  // namespace

// ============================================================== host side ====
void MatchWorkspace::begin_frame(const size_t nk[3], int n_iters, cudaStream_t s)
{
  off[0] = 0;
  for (int k = 0; k < 3; ++k) off[k + 1] = off[k] + nk[k];
  total = off[3];
  for (int k = 0; k < 3; ++k) {
    const size_t q = std::max<size_t>(nk[k], 1);
    nn_idx[k].reserve(q * KNN_MAX_K, s); nn_d2[k].reserve(q * KNN_MAX_K, s); nn_found[k].reserve(q, s);
  }
  rec_stride = (total + 31) / 32 * 32;
  records.reserve(std::max<size_t>(REC_PLANES * rec_stride, REC_PLANES), s); status.reserve(total, s); weight.reserve(total, s);
  // one histogram block per ICP iteration: an iteration that comes after the loop's break leaves everything untouched
  hist_iters = std::max(n_iters, 1);
  hist.reserve((size_t)hist_iters * kHistBlock, s);
  h_hist.reserve(kHistBlock);
  LSB_CUDA(cudaMemsetAsync(hist.p, 0, (size_t)hist_iters * kHistBlock * sizeof(int), s));
}

void MatchWorkspace::match_type(int type, const lsb_point* d_kp, size_t nk, const KnnGrid& target, const Iso3d& prior,
                                const lsb_match_params& mp, cudaStream_t s, LaunchCounter& lc, int icp_iter, const IcpState* icp, bool reuse, bool refine)
{
  if (nk == 0) return;
  double* rec = records.p + off[type];
  uint8_t* st = status.p + off[type];
  double* wt = weight.p + off[type];
  int* h = hist.p + (size_t)icp_iter * kHistBlock + type * LSB_MATCH_N_STATUS;
  const int* done = icp ? &icp->done : nullptr;
  const unsigned G = div_up(nk, FIT_TPB);
  if (target.empty()) {
    k_fit_empty<<<G, FIT_TPB, 0, s>>>((int)nk, rec, rec_stride, st, wt, done); LSB_LAUNCH_CHECK(); lc.add();
    return;
  }
  FitParams fp;
  fp.single_edge_per_ring = mp.single_edge_per_ring;
  fp.sq_max_neighbors_distance = mp.max_neighbors_distance * mp.max_neighbors_distance;
  fp.edge_nb = mp.edge_nb_neighbors; fp.edge_min_nb = mp.edge_min_nb_neighbors;
  fp.edge_max_model_error = mp.edge_max_model_error;
  fp.sq_edge_max_model_error = mp.edge_max_model_error * mp.edge_max_model_error;
  fp.sq_edge_inlier = (float)(mp.edge_max_model_error * mp.edge_max_model_error);
  fp.plane_nb = mp.plane_nb_neighbors; fp.planarity_threshold = mp.planarity_threshold;
  fp.plane_max_model_error = mp.plane_max_model_error;
  fp.sq_plane_max_model_error = mp.plane_max_model_error * mp.plane_max_model_error;
  fp.blob_nb = mp.blob_nb_neighbors;
  const unsigned kk = (type == LSB_EDGE) ? mp.edge_nb_neighbors : (type == LSB_PLANE ? mp.plane_nb_neighbors : mp.blob_nb_neighbors);
  if (kk > (unsigned)KNN_MAX_K) throw CudaError(LSB_E_CAPACITY, "more than 32 neighbours requested");
  const int k = std::max(1, (int)kk);
  t_knn.begin(s);
  // planes / blobs only need neighbours within MaxNeighborsDistance (else NEIGHBORS_TOO_FAR);
  // the line neighbourhood selection needs the exact k nearest whatever their distance
  const float cap = (type == LSB_EDGE) ? -1.f : (float)mp.max_neighbors_distance * 1.0001f + 1e-3f;
  KnnDeviceArgs da;
  // the pose the queries are moved by: of the device-resident loop (Tworld for the localization, Trelative for the ego-motion), or `prior`
  da.prior = icp ? (icp_stage == ICP_STAGE_LOC ? &icp->Tworld : &icp->Trelative) : nullptr;
  da.done = done;
  da.reuse = reuse;
  da.refine = (reuse && icp && refine) ? &icp->Pending : nullptr;
  target.search_points(d_kp, nk, prior, k, cap, nn_idx[type].p, nn_d2[type].p, nn_found[type].p, /*entry_major=*/true, s, lc, &da);
  t_knn.end(s);
  const int NQ = (int)nk;
  t_fit.begin(s);
  if (type == LSB_EDGE) k_fit<LSB_EDGE><<<G, FIT_TPB, 0, s>>>(d_kp, NQ, target.cloud, nn_idx[type].p, nn_d2[type].p, nn_found[type].p, k, fp, rec, rec_stride, st, wt, h, done, target.unres_count.p);
  else if (type == LSB_PLANE) k_fit<LSB_PLANE><<<G, FIT_TPB, 0, s>>>(d_kp, NQ, target.cloud, nn_idx[type].p, nn_d2[type].p, nn_found[type].p, k, fp, rec, rec_stride, st, wt, h, done, target.unres_count.p);
  else k_fit<LSB_BLOB><<<G, FIT_TPB, 0, s>>>(d_kp, NQ, target.cloud, nn_idx[type].p, nn_d2[type].p, nn_found[type].p, k, fp, rec, rec_stride, st, wt, h, done, target.unres_count.p);
  LSB_LAUNCH_CHECK(); lc.add();
  target.counters_clean = target.unres_count.p != nullptr;
  t_fit.end(s);
}

void MatchWorkspace::read_histograms(int hist_out[3][LSB_MATCH_N_STATUS], cudaStream_t s)
{
  d2h_bytes += 3 * LSB_MATCH_N_STATUS * sizeof(int);
  LSB_CUDA(cudaMemcpyAsync(h_hist.p, hist.p, 3 * LSB_MATCH_N_STATUS * sizeof(int), cudaMemcpyDeviceToHost, s));   // block of iteration 0
  LSB_CUDA(cudaStreamSynchronize(s));
  for (int t = 0; t < 3; ++t)
    for (int q = 0; q < LSB_MATCH_N_STATUS; ++q) hist_out[t][q] = h_hist.p[t * LSB_MATCH_N_STATUS + q];
}

void MatchWorkspace::enqueue_solve(const double x0[6], int max_iter, int two_d, double saturation, unsigned min_matches, cudaStream_t s,
                                   LaunchCounter& lc, const LmSensorTerms* sensor_terms, IcpState* icp, const IcpCtl* ctl)
{
  const int n = (int)total;
  LmSensorTerms terms;
  if (sensor_terms) terms = *sensor_terms; else lm_sensor_terms_clear(terms);
  const int nchunks = std::max(1, (int)div_up(total, LSB_LM_CHUNK));
  const int pstride = (nchunks + 31) / 32 * 32;
  constexpr size_t kSmem = (size_t)REC_PLANES * LSB_LM_CHUNK * sizeof(double);
  if (lm_grid_limit == 0) {      // co-resident CTAs of the persistent kernel on this device
    int dev = 0, sms = 0, per_sm = 0;
    LSB_CUDA(cudaGetDevice(&dev));
    LSB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    LSB_CUDA(cudaFuncSetAttribute(k_lm_solve, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    LSB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_lm_solve, LM_THREADS, kSmem));
    if (per_sm < 1) throw CudaError(LSB_E_CUDA, "k_lm_solve does not fit on a multiprocessor");
    lm_grid_limit = sms * per_sm;
  }
  part_a.reserve((size_t)2 * pstride * LSB_LM_NSUM, s);
  lm_state.reserve(1, s); h_lm.reserve(1);
  if (!h_block.p) { h_block.reserve(1); h_block.p->seq = 0; }
  if (!lm_counter.p) {           // [arrivals, exits] of the grid barrier: zero between solves
    lm_counter.reserve(2, s);
    LSB_CUDA(cudaMemsetAsync(lm_counter.p, 0, 2 * sizeof(unsigned), s));
  }
  LmStart start;
  for (int i = 0; i < 6; ++i) start.x[i] = x0 ? x0[i] : 0.0;
  IcpCtl c{};
  if (ctl) c = *ctl;
  const int iter = ctl ? ctl->icp_iter : 0;
  t_lm.begin(s);
  // one cooperative launch = the whole LM loop (co-residency of the grid is what the barrier needs)
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeCooperative;
  attr.val.cooperative = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)std::min(nchunks, std::max(lm_grid_limit / std::max(lm_grid_share, 1), 8)));
  cfg.blockDim = dim3(LM_THREADS);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = s;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  LSB_CUDA(cudaLaunchKernelEx(&cfg, k_lm_solve, (const double*)records.p, rec_stride, n, nchunks, pstride, saturation, start, max_iter, two_d,
                              (const int*)(hist.p + (size_t)iter * kHistBlock), min_matches, lm_state.p, part_a.p, lm_counter.p, terms, h_block.p,
                              solve_seq, icp, c));
  lc.add();
  t_lm.end(s);
}

unsigned MatchWorkspace::next_solve_seq()
{
  if (!h_block.p) { h_block.reserve(1); h_block.p->seq = 0; }
  return solve_seq = (++solve_seq ? solve_seq : ++solve_seq);   // never 0 (the initial value of the flag)
}

const MatchWorkspace::LmHostBlock& MatchWorkspace::wait_result(cudaStream_t s)
{
  d2h_bytes += sizeof(LmHostBlock);
  // wait for the kernel's own hand-off (no copy, no stream synchronisation); the stream is queried now and then so that a
  // failed launch surfaces as an error instead of an endless wait
  const unsigned seq = solve_seq;
  volatile unsigned* flag = &h_block.p->seq;
  for (unsigned spins = 1; *flag != seq; ++spins) {
    if ((spins & 4095u) == 0) {
      const cudaError_t q = cudaStreamQuery(s);
      if (q == cudaSuccess) { if (*flag != seq) throw CudaError(LSB_E_CUDA, "k_lm_solve finished without handing its result over"); break; }
      if (q != cudaErrorNotReady) cuda_check(q, "k_lm_solve", __FILE__, __LINE__);
    }
  }
  std::atomic_thread_fence(std::memory_order_acquire);
  return *h_block.p;
}

LmState MatchWorkspace::solve(const double x0[6], int max_iter, int two_d, double saturation, unsigned min_matches,
                              int hist_out[3][LSB_MATCH_N_STATUS], cudaStream_t s, LaunchCounter& lc, const LmSensorTerms* sensor_terms)
{
  next_solve_seq();
  enqueue_solve(x0, max_iter, two_d, saturation, min_matches, s, lc, sensor_terms, nullptr, nullptr);
  const LmHostBlock& blk = wait_result(s);
  for (int t = 0; t < 3; ++t)
    for (int q = 0; q < LSB_MATCH_N_STATUS; ++q) hist_out[t][q] = blk.hist[t * LSB_MATCH_N_STATUS + q];
  LmState st;
  std::memcpy(&st, (const void*)&blk.state, sizeof(LmState));
  return st;
}

}  // namespace lsb

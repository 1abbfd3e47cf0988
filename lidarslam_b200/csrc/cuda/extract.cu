// GPU SpinningSensorKeypointExtractor — kernels K1-K5.
// Reference behaviour: slam_lib/src/SpinningSensorKeypointExtractor.cxx
//   K1 ring bucket        <- ConvertAndSortScanLines      :139-171
//   K2 invalidate         <- InvalidateNotUsablePoints    :207-308
//   K3 curvature scores   <- ComputeCurvature             :311-471 (+ LineFitting :33-115)
//   K4 label              <- SetKeyPointsLabels           :474-573
//   K5 compaction         <- SetKeyPointsLabels           :575-589
//   azimuth estimation    <- EstimateAzimuthalResolution  :593-637
// Design: points are stably bucketed by laser_id with a radix sort (arrival order
// kept inside a ring), all per-point work runs one thread per point over the
// whole scan, and the order-dependent greedy labelling of the reference (sort by
// score, walk, suppress neighbours) is evaluated WITHOUT sorting as a parallel
// non-maximum suppression over the total order (score, index): a candidate is
// selected iff no selected higher-priority candidate lies within its half width,
// which is exactly what the sequential walk computes.
// All arithmetic is plain IEEE fp32 in the reference's operation order
// (-fmad=false); trigonometry comes from common/detmath.h.
#include "extract.cuh"
#include <cstring>
#include "../common/pca.h"
#include <cub/cub.cuh>
#include <cfloat>

namespace lsb {

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ Vec3f xyz(const float4& p) { return Vec3f(p.x, p.y, p.z); }

// ---------------------------------------------------------------- K1 ---------
__global__ void k_ring_keys(const lsb_point* __restrict__ pts, int n, uint32_t* __restrict__ key, uint32_t* __restrict__ idx)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  // laser_id lives in the second 16-byte half of the 32-byte point
  const uint4 hi = reinterpret_cast<const uint4*>(pts)[2 * i + 1];
  key[i] = hi.w & 0xffffu;
  idx[i] = (uint32_t)i;
}

__global__ void k_ring_gather(const lsb_point* __restrict__ pts, int n, const uint32_t* __restrict__ key,
                              const uint32_t* __restrict__ src, float4* __restrict__ p4, int* __restrict__ ring_start,
                              int* __restrict__ ring_end, uint32_t* __restrict__ d_counts)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t s = key[j];
  if (s < KE_MAX_RINGS) {
    if (j == 0 || key[j - 1] != s) ring_start[s] = j;
    if (j == n - 1 || key[j + 1] != s) ring_end[s] = j + 1;
  }
  if (j == n - 1) d_counts[3] = s + 1;
  const uint32_t i = src[j];
  const float4 lo = reinterpret_cast<const float4*>(pts)[2 * i];
  const float4 hi = reinterpret_cast<const float4*>(pts)[2 * i + 1];
  p4[j] = make_float4(lo.x, lo.y, lo.z, hi.z);   // hi = {time lo, time hi, intensity, ids}
}

// ---------------------------------------------------------------- K2 ---------
__global__ void k_invalidate(const float4* __restrict__ p4, const uint32_t* __restrict__ key, int n,
                             const int* __restrict__ ring_start, const int* __restrict__ ring_end,
                             uint8_t* __restrict__ valid, int W, float minDist, float maxPosDiffCoeff)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  const uint32_t s = key[j];
  if (s >= KE_MAX_RINGS) { valid[j] = 0; return; }
  const int start = ring_start[s];
  const int Npts = ring_end[s] - start;
  const int i = j - start;
  if (Npts < 2 * W + 1 || i < W || i >= Npts - W) { valid[j] = 0; return; }
  const Vec3f cur = xyz(p4[j]);
  const float L = norm(cur);
  if (L < minDist) valid[j] = 0;
  const float a = L * maxPosDiffCoeff;
  const float maxPosDiff = (a < 0.02f) ? 0.02f : a;
  const float sq = maxPosDiff * maxPosDiff;
  const Vec3f nxt = xyz(p4[j + 1]);
  if (sqnorm(nxt - cur) > sq) {
    if (L < norm(nxt)) {
      valid[j + 1] = 0;
      for (int q = j + 1; q < j + W; ++q) {
        if (sqnorm(xyz(p4[q + 1]) - xyz(p4[q])) > sq) break;
        valid[q + 1] = 0;
      }
    } else {
      valid[j] = 0;
      for (int q = j - 1; q > j - W; --q) {
        if (sqnorm(xyz(p4[q + 1]) - xyz(p4[q])) > sq) break;
        valid[q] = 0;
      }
    }
  }
}

// ---------------------------------------------------------------- K3 ---------
struct Line { Vec3f dir, pos; };

__device__ __forceinline__ float line_sqdist(const Line& l, const Vec3f& p) { return sqnorm(cross(p - l.pos, l.dir)); }

// LineFitting::FitPCA (:59-84) over q[0..n).  NT > 0: the count is known at compile time (n == NT), the loops unroll and the
// caller's neighbour arrays stay in registers (with a run-time count they live in local memory); same operations, same order.
template <int NT>
__device__ __forceinline__ bool fit_pca(const Vec3f* q, int n_rt, Line& l)
{
  const int n = NT > 0 ? NT : n_rt;
  CovAccum<float> acc;
#pragma unroll
  for (int t = 0; t < (NT > 0 ? NT : 2 * KE_MAX_WIDTH); ++t) if (t < n) acc.add(q[t].x, q[t].y, q[t].z);
  Mat3f cov, evecs;
  float evals[3];
  acc.finish(l.pos, cov);
  eigen33(cov, evecs, evals);
  l.dir = evecs.col(2);
  const float maxd = 0.02f;
  const float sqMax = maxd * maxd;
  bool ok = true;
#pragma unroll
  for (int t = 0; t < (NT > 0 ? NT : 2 * KE_MAX_WIDTH); ++t)
    if (t < n && line_sqdist(l, q[t]) > sqMax) ok = false;
  return ok;
}

// LineFitting::FitPCAAndCheckConsistency (:87-108)
template <int NT>
__device__ __forceinline__ bool fit_pca_consistent(const Vec3f* q, int n_rt, Line& l, float maxSinAngle)
{
  const int n = NT > 0 ? NT : n_rt;
  const Vec3f U = normalized(q[n - 1] - q[0]);
  bool ok = true;
#pragma unroll
  for (int t = 0; t < (NT > 0 ? NT - 1 : KE_MAX_WIDTH - 1); ++t) {
    if (t < n - 1) {
      const Vec3f V = normalized(q[t + 1] - q[t]);
      const float sinAngle = norm(cross(U, V));
      if (sinAngle > maxSinAngle) ok = false;
    }
  }
  if (!ok) return false;
  return fit_pca<NT>(q, n, l);
}

// WT > 0: neighbour half-width known at compile time (W == WT; the reference's default is 4), WT == 0: any width up to KE_MAX_WIDTH
template <int WT>
__global__ void __launch_bounds__(128)
k_curvature(const float4* __restrict__ p4, const uint32_t* __restrict__ key, int n, const int* __restrict__ ring_start,
            const int* __restrict__ ring_end, const uint8_t* __restrict__ valid, float* __restrict__ angles,
            float* __restrict__ depth_gap, float* __restrict__ saliency, float* __restrict__ intensity_gap, int W_rt,
            float sqDistToLineThreshold, float maxSinAngle)
{
  const int W = WT > 0 ? WT : W_rt;
  constexpr int WMAX = WT > 0 ? WT : KE_MAX_WIDTH;
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  float oAngle = 0.f, oDepth = 0.f, oSal = 0.f, oInt = 0.f;
  const uint32_t s = key[j];
  bool active = s < KE_MAX_RINGS && valid[j] != 0;
  int start = 0, Npts = 0;
  if (active) {
    start = ring_start[s];
    Npts = ring_end[s] - start;
    const int i = j - start;
    active = !(Npts < 2 * W + 1) && i >= W && i + W < Npts;
  }
  if (active) {
    Vec3f left[WMAX], right[WMAX];
    const float4 c4 = p4[j];
    const Vec3f central = xyz(c4);
#pragma unroll
    for (int t = 0; t < WMAX; ++t) {      // from the centre outward (:363-366)
      if (t < W) {
        left[t] = xyz(p4[j - 1 - t]);
        right[t] = xyz(p4[j + 1 + t]);
      }
    }
    const float ig = p4[j + 1].w - p4[j - 1].w;
    oInt = (ig < 0.f) ? -ig : ig;
    Line leftLine, rightLine;
    const bool leftFlat = fit_pca_consistent<WT>(left, W, leftLine, maxSinAngle);
    const bool rightFlat = fit_pca_consistent<WT>(right, W, rightLine, maxSinAngle);
    float distLeft = 0.f, distRight = 0.f;
    if (leftFlat && rightFlat) {
      distLeft = line_sqdist(leftLine, central);
      distRight = line_sqdist(rightLine, central);
      if (distLeft < sqDistToLineThreshold && distRight < sqDistToLineThreshold)
        oAngle = norm(cross(leftLine.dir, rightLine.dir));
    } else if (!leftFlat && rightFlat) {
      distLeft = FLT_MAX;
#pragma unroll
      for (int t = 0; t < WMAX; ++t) if (t < W) { const float d = line_sqdist(rightLine, left[t]); distLeft = (d < distLeft) ? d : distLeft; }
      distLeft *= 0.25f;
    } else if (leftFlat && !rightFlat) {
      distRight = FLT_MAX;
#pragma unroll
      for (int t = 0; t < WMAX; ++t) if (t < W) { const float d = line_sqdist(leftLine, right[t]); distRight = (d < distRight) ? d : distRight; }
      distRight *= 0.25f;
    } else {
      const float sqCurrDepth = sqnorm(central);
      const float minDepthGapDist = 1.5f;
      Vec3f far[2 * WMAX];
      bool gapL = false, gapR = false, stopL = false, stopR = false;
      int nfar = 0;
#pragma unroll
      for (int t = 0; t < WMAX; ++t) {
        if (t < W && !stopL) {
          const float d = sqnorm(left[t]) - sqCurrDepth;
          if (((d < 0.f) ? -d : d) > minDepthGapDist) { gapL = true; far[nfar++] = left[t]; }
          else if (gapL) stopL = true;
        }
      }
#pragma unroll
      for (int t = 0; t < WMAX; ++t) {
        if (t < W && !stopR) {
          const float d = sqnorm(right[t]) - sqCurrDepth;
          if (((d < 0.f) ? -d : d) > minDepthGapDist) { gapR = true; far[nfar++] = right[t]; }
          else if (gapR) stopR = true;
        }
      }
      if (nfar > W) {
        Line farLine;
        fit_pca<0>(far, nfar, farLine);
        oSal = line_sqdist(farLine, central);
      }
    }
    oDepth = (distLeft < distRight) ? distRight : distLeft;
  }
  angles[j] = oAngle;
  depth_gap[j] = oDepth;
  saliency[j] = oSal;
  intensity_gap[j] = oInt;
}

// ---------------------------------------------------------------- K4 ---------
// One CTA per ring.  Parallel non-maximum suppression over the total order the
// reference's sort+walk defines (:499-563).  state: 0 = not selectable (not a
// candidate, invalid, or suppressed), 1 = undecided, 2 = selected.
constexpr int LABEL_THREADS = 512;
constexpr int LABEL_SMEM_LEN = 4096;   // longer rings use global scratch

template <bool PLANE>
__device__ __forceinline__ bool nms_higher(float vq, int q, float vi, int i)
{
  // EDGE criteria: (value desc, index asc).  PLANE: the reference walks the
  // descending Angles order backwards => (value asc, index desc).
  if (PLANE) return (vq < vi) || (vq == vi && q > i);
  return (vq > vi) || (vq == vi && q < i);
}

template <bool PLANE>
__device__ void nms_phase(const float* __restrict__ gvals, float* vals, uint8_t* v, uint8_t* lab, uint8_t* st, int L, float thr,
                          int hw, uint8_t bit)
{
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    const float val = gvals[i];
    vals[i] = val;
    bool cand;
    if (PLANE) cand = !(val > thr) && !((double)val < 1e-6);
    else cand = !(val < thr);
    st[i] = (cand && (v[i] & bit)) ? 1 : 0;
  }
  __syncthreads();
  int changed;
  do {
    changed = 0;
    for (int i = threadIdx.x; i < L; i += blockDim.x) {
      if (st[i] != 1) continue;
      const float vi = vals[i];
      const int b = max(0, i - hw), e = min(L - 1, i + hw);
      bool rej = false, wait = false;
      for (int q = b; q <= e; ++q) {
        if (q == i) continue;
        const uint8_t sq = st[q];
        if (sq == 2) rej = true;
        else if (sq == 1 && nms_higher<PLANE>(vals[q], q, vi, i)) wait = true;
      }
      if (rej) { st[i] = 0; changed = 1; }
      else if (!wait) { st[i] = 2; changed = 1; }
    }
    changed = __syncthreads_or(changed);
  } while (changed);
  // selected points take the label and clear the validity bit of idx +- hw (:513-523)
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    const int b = max(0, i - hw), e = min(L - 1, i + hw);
    bool near = false;
    for (int q = b; q <= e; ++q) near |= (st[q] == 2);
    if (st[i] == 2) lab[i] |= bit;
    if (near) v[i] &= (uint8_t)~bit;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(LABEL_THREADS)
k_label(const int* __restrict__ ring_start, const int* __restrict__ ring_end, const float* __restrict__ angles,
        const float* __restrict__ depth_gap, const float* __restrict__ saliency, const float* __restrict__ intensity_gap,
        uint8_t* __restrict__ valid, uint8_t* __restrict__ label, float* __restrict__ g_vals, uint8_t* __restrict__ g_state,
        int W, float sqEdgeDepthGapThr, float edgeSinAngleThr, float sqEdgeSaliencyThr, float edgeIntensityGapThr,
        float planeSinAngleThr)
{
  __shared__ float s_vals[LABEL_SMEM_LEN];
  __shared__ uint8_t s_v[LABEL_SMEM_LEN], s_lab[LABEL_SMEM_LEN], s_st[LABEL_SMEM_LEN];
  const int s = blockIdx.x;
  const int start = ring_start[s];
  const int L = ring_end[s] - start;
  if (L < 2 * W + 1) return;   // IsScanLineAlmostEmpty: labels stay 0, validity already 0
  const bool in_smem = L <= LABEL_SMEM_LEN;
  float* vals = in_smem ? s_vals : g_vals + start;
  uint8_t* st = in_smem ? s_st : g_state + start;
  uint8_t* v = in_smem ? s_v : valid + start;
  uint8_t* lab = in_smem ? s_lab : label + start;
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    if (in_smem) v[i] = valid[start + i];
    lab[i] = 0;
  }
  __syncthreads();
  // Edges: depth gap, angle, saliency, intensity gap, in this order (:527-533)
  nms_phase<false>(depth_gap + start, vals, v, lab, st, L, sqEdgeDepthGapThr, W - 1, 1);
  nms_phase<false>(angles + start, vals, v, lab, st, L, edgeSinAngleThr, W, 1);
  nms_phase<false>(saliency + start, vals, v, lab, st, L, sqEdgeSaliencyThr, W - 1, 1);
  nms_phase<false>(intensity_gap + start, vals, v, lab, st, L, edgeIntensityGapThr, 1, 1);
  // Planes (:536-563): ascending sin angle, suppress +-4
  nms_phase<true>(angles + start, vals, v, lab, st, L, planeSinAngleThr, 4, 2);
  // Blobs (:568-572) and validity re-set of labelled points (:575-589)
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    uint8_t l = lab[i];
    if (i % 3 == 0 && (v[i] & 4)) l |= 4;
    label[start + i] = l;
    valid[start + i] = v[i] | l;
  }
}

// ---------------------------------------------------------------- K5 ---------
__global__ void k_pack_flags(const uint8_t* __restrict__ label, int n, unsigned long long* __restrict__ out, uint32_t* __restrict__ d_counts)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < 6) reinterpret_cast<unsigned long long*>(d_counts + 8)[j] = (j & 1) ? 0ull : ~0ull;   // time (min, max) per type
  if (j >= n) return;
  const unsigned long long l = label[j];
  out[j] = (l & 1ull) | (((l >> 1) & 1ull) << 21) | (((l >> 2) & 1ull) << 42);
}

// order-preserving map double -> uint64 (atomicMin / atomicMax on the point times)
__device__ __forceinline__ unsigned long long time_ord(double d)
{
  const unsigned long long u = (unsigned long long)__double_as_longlong(d);
  return (u & 0x8000000000000000ull) ? ~u : (u | 0x8000000000000000ull);
}

// Scatter + per-type time range of the keypoints (Slam::InitUndistortion needs the min / max of the
// `time` field over the keypoints, Slam.cxx:1293-1300: taken here so that it comes back with the counts).
// d_counts: [0..2] counts, [3] nb rings, [8..19] = 3 x (min, max) as ordered uint64 (initialised by k_pack_flags).
__global__ void k_scatter_keypoints(const lsb_point* __restrict__ pts, const uint32_t* __restrict__ src,
                                    const uint8_t* __restrict__ label, const unsigned long long* __restrict__ pos, int n,
                                    lsb_point* __restrict__ out0, lsb_point* __restrict__ out1, lsb_point* __restrict__ out2,
                                    uint32_t* __restrict__ d_counts)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  uint8_t l = 0;
  unsigned long long t_ord = 0ull;
  if (j < n) {
    l = label[j];
    const unsigned long long p = pos[j];
    const uint32_t p0 = (uint32_t)(p & 0x1fffffull), p1 = (uint32_t)((p >> 21) & 0x1fffffull), p2 = (uint32_t)((p >> 42) & 0x1fffffull);
    if (l) {
      const uint4* sp = reinterpret_cast<const uint4*>(pts) + 2 * (size_t)src[j];
      const uint4 a = sp[0], b = sp[1];
      if (l & 1) { uint4* o = reinterpret_cast<uint4*>(out0) + 2 * (size_t)p0; o[0] = a; o[1] = b; }
      if (l & 2) { uint4* o = reinterpret_cast<uint4*>(out1) + 2 * (size_t)p1; o[0] = a; o[1] = b; }
      if (l & 4) { uint4* o = reinterpret_cast<uint4*>(out2) + 2 * (size_t)p2; o[0] = a; o[1] = b; }
      t_ord = time_ord(__longlong_as_double((long long)(((unsigned long long)b.y << 32) | b.x)));
    }
    if (j == n - 1) {
      d_counts[0] = p0 + (l & 1);
      d_counts[1] = p1 + ((l >> 1) & 1);
      d_counts[2] = p2 + ((l >> 2) & 1);
    }
  }
  if (__ballot_sync(0xffffffffu, l != 0) == 0u) return;
  unsigned long long* tmm = reinterpret_cast<unsigned long long*>(d_counts + 8);
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const bool in = (l >> k) & 1;
    unsigned long long mn = in ? t_ord : ~0ull, mx = in ? t_ord : 0ull;
#pragma unroll
    for (int off = 16; off; off >>= 1) {
      mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, off));
      mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    }
    // the bounds only tighten: a (possibly stale) read that already beats this warp's value makes the atomic pointless,
    // and a stale read can only be looser than the truth, so no needed update is ever skipped
    if ((threadIdx.x & 31) == 0 && mn != ~0ull) {
      if (mn < __ldcg(tmm + 2 * k)) atomicMin(tmm + 2 * k, mn);
      if (mx > __ldcg(tmm + 2 * k + 1)) atomicMax(tmm + 2 * k + 1, mx);
    }
  }
}

// ------------------------------------------------- azimuthal resolution ------
__global__ void k_az_angles(const float4* __restrict__ p4, const uint32_t* __restrict__ key, int n, float* __restrict__ out)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  float angle = -1.f;   // sentinel: dropped by the select
  if (j > 0 && key[j - 1] == key[j]) {
    const float x1 = p4[j - 1].x, y1 = p4[j - 1].y, x2 = p4[j].x, y2 = p4[j].y;
    const float d = x1 * x2 + y1 * y2;
    const float n1 = sqrtf(x1 * x1 + y1 * y1), n2 = sqrtf(x2 * x2 + y2 * y2);
    float a = dm::acosf(d / (n1 * n2));
    if (a < 0.f) a = -a;
    if ((double)a > 1e-4) angle = a;
  }
  out[j] = angle;
}

struct AzKeep { __device__ bool operator()(float a) const { return a >= 0.f; } };

// the iterative median of :622-634 on the sorted angles, one thread
__global__ void k_az_median(const float* __restrict__ sorted, const int* __restrict__ n_sel, float* __restrict__ d_az)
{
  const int n = *n_sel;
  d_az[1] = (float)n;
  if (n < 100) { d_az[0] = -1.f; return; }
  int maxInliersIdx = n;
  float maxAngle = (float)(5. / 180. * 3.14159265358979323846);
  float medianAngle = 0.f;
  while ((double)maxAngle > 1.8 * (double)medianAngle) {
    int lo = 0, hi = maxInliersIdx;   // std::upper_bound: first element > maxAngle
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (!(maxAngle < sorted[mid])) lo = mid + 1; else hi = mid;
    }
    maxInliersIdx = lo;
    medianAngle = sorted[maxInliersIdx / 2];
    const double a = (double)medianAngle * 2., b = (double)maxAngle / 1.8;
    maxAngle = (float)((a < b) ? a : b);
  }
  d_az[0] = medianAngle;
}

// ------------------------------------------------------------ debug ----------
__global__ void k_debug_field(const float* __restrict__ f, const uint8_t* __restrict__ bits, int shift, const uint32_t* __restrict__ src,
                              int n, float* __restrict__ out)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  out[src[j]] = f ? f[j] : (float)((bits[j] >> shift) & 1);
}
__global__ void k_labels_input_order(const uint8_t* __restrict__ label, const uint32_t* __restrict__ src, int n, uint8_t* __restrict__ out)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= n) return;
  out[src[j]] = label[j];
}

}  // namespace

// ============================================================== host side ====
ExtractorGpu::ExtractorGpu()
{
  lsb_default_ke_params(&P);
}

void ExtractorGpu::compute(const lsb_point* d_pts, size_t n_, cudaStream_t s, LaunchCounter& lc)
{
  if (n_ > KE_MAX_POINTS) throw CudaError(LSB_E_CAPACITY, "frame has more than 2^21-1 points");
  if (P.neighbor_width < 1 || P.neighbor_width > KE_MAX_WIDTH) throw CudaError(LSB_E_INVALID, "neighbor_width must be in [1, 8]");
  n = n_;
  counts[0] = counts[1] = counts[2] = 0;
  if (n == 0) return;
  const int N = (int)n;
  const int W = P.neighbor_width;
  key_a.reserve(n, s); key_b.reserve(n, s); idx_a.reserve(n, s); src.reserve(n, s);
  p4.reserve(n + 1, s); valid.reserve(n, s); label.reserve(n, s); nms_state.reserve(n, s);
  angles.reserve(n, s); depth_gap.reserve(n, s); saliency.reserve(n, s); intensity_gap.reserve(n, s); nms_vals.reserve(n, s);
  scan_in.reserve(n, s); scan_out.reserve(n, s);
  for (int k = 0; k < 3; ++k) kp[k].reserve(n, s);
  ring_start.reserve(KE_MAX_RINGS, s); ring_end.reserve(KE_MAX_RINGS, s);
  d_counts.reserve(24, s); d_az.reserve(4, s);
  h_counts.reserve(24); h_az.reserve(4);

  const unsigned G = div_up(n, TPB);
  // K1: stable bucket by laser_id
  k_ring_keys<<<G, TPB, 0, s>>>(d_pts, N, key_a.p, idx_a.p); LSB_LAUNCH_CHECK(); lc.add();
  size_t tmp = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, tmp, key_a.p, key_b.p, idx_a.p, src.p, N, 0, 16, s);
  size_t tmp2 = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp2, scan_in.p, scan_out.p, N, s);
  size_t tmp3 = 0;
  cub::DeviceSelect::If(nullptr, tmp3, angles.p, depth_gap.p, (int*)d_counts.p, N, AzKeep(), s);
  size_t tmp4 = 0;
  cub::DeviceRadixSort::SortKeys(nullptr, tmp4, angles.p, depth_gap.p, N, 0, 32, s);
  cub_tmp.reserve(std::max(std::max(tmp, tmp2), std::max(tmp3, tmp4)) + 256, s);
  size_t tb = cub_tmp.cap;
  LSB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tb, key_a.p, key_b.p, idx_a.p, src.p, N, 0, 16, s)); lc.add(3);
  LSB_CUDA(cudaMemsetAsync(ring_start.p, 0, KE_MAX_RINGS * sizeof(int), s));
  LSB_CUDA(cudaMemsetAsync(ring_end.p, 0, KE_MAX_RINGS * sizeof(int), s));
  k_ring_gather<<<G, TPB, 0, s>>>(d_pts, N, key_b.p, src.p, p4.p, ring_start.p, ring_end.p, d_counts.p); LSB_LAUNCH_CHECK(); lc.add();

  // azimuthal resolution (:169): estimated when outside [1e-6, pi/4]
  float az = P.azimuthal_resolution;
  if (az < 1e-6 || M_PI / 4. < az) {
    ++az_estimations;
    az_angles.reserve(n, s); az_sorted.reserve(n, s);
    k_az_angles<<<G, TPB, 0, s>>>(p4.p, key_b.p, N, az_angles.p); LSB_LAUNCH_CHECK(); lc.add();
    tb = cub_tmp.cap;
    LSB_CUDA(cub::DeviceSelect::If(cub_tmp.p, tb, az_angles.p, az_sorted.p, (int*)(d_counts.p + 4), N, AzKeep(), s)); lc.add(2);
    // runs on the first frame only: one extra sync to size the sort
    LSB_CUDA(cudaMemcpyAsync(h_counts.p + 4, d_counts.p + 4, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    LSB_CUDA(cudaStreamSynchronize(s));
    const int nsel = (int)h_counts.p[4];
    if (nsel > 0) {
      tb = cub_tmp.cap;
      LSB_CUDA(cub::DeviceRadixSort::SortKeys(cub_tmp.p, tb, az_sorted.p, az_angles.p, nsel, 0, 32, s)); lc.add(4);
    }
    k_az_median<<<1, 1, 0, s>>>(az_angles.p, (const int*)(d_counts.p + 4), d_az.p); LSB_LAUNCH_CHECK(); lc.add();
    LSB_CUDA(cudaMemcpyAsync(h_az.p, d_az.p, 2 * sizeof(float), cudaMemcpyDeviceToHost, s));
    LSB_CUDA(cudaStreamSynchronize(s));
    if (h_az.p[0] >= 0.f) az = h_az.p[0];   // fewer than 100 angles: keep the previous value (:616-620)
    P.azimuthal_resolution = az;
  }

  // K2
  const float angleBeamNormal = (float)((double)(90 - P.min_beam_surface_angle) / 180. * M_PI);   // :210
  float azr = az;
  if (azr < 1e-6 || M_PI / 4 < azr) azr = (float)(0.2 / 180. * M_PI);                              // :215-219
  const float maxPosDiffCoeff = dm::sinf(azr) / dm::cosf(azr + angleBeamNormal);                    // :222
  LSB_CUDA(cudaMemsetAsync(valid.p, 7, n, s));
  k_invalidate<<<G, TPB, 0, s>>>(p4.p, key_b.p, N, ring_start.p, ring_end.p, valid.p, W, P.min_distance_to_sensor, maxPosDiffCoeff);
  LSB_LAUNCH_CHECK(); lc.add();
  // K3
  const float maxSinAngle = dm::sinf((float)(40. * 0.017453293));   // LineFitting::MaxAngle (:55)
  if (W == 4)
    k_curvature<4><<<div_up(n, 128), 128, 0, s>>>(p4.p, key_b.p, N, ring_start.p, ring_end.p, valid.p, angles.p, depth_gap.p, saliency.p,
                                                  intensity_gap.p, W, P.dist_to_line_threshold * P.dist_to_line_threshold, maxSinAngle);
  else
    k_curvature<0><<<div_up(n, 128), 128, 0, s>>>(p4.p, key_b.p, N, ring_start.p, ring_end.p, valid.p, angles.p, depth_gap.p, saliency.p,
                                                  intensity_gap.p, W, P.dist_to_line_threshold * P.dist_to_line_threshold, maxSinAngle);
  LSB_LAUNCH_CHECK(); lc.add();
  // K4
  LSB_CUDA(cudaMemsetAsync(label.p, 0, n, s));
  k_label<<<KE_MAX_RINGS, LABEL_THREADS, 0, s>>>(ring_start.p, ring_end.p, angles.p, depth_gap.p, saliency.p, intensity_gap.p, valid.p,
                                                 label.p, nms_vals.p, nms_state.p, W,
                                                 P.edge_depth_gap_threshold * P.edge_depth_gap_threshold, P.edge_sin_angle_threshold,
                                                 P.edge_saliency_threshold * P.edge_saliency_threshold,
                                                 P.edge_intensity_gap_threshold, P.plane_sin_angle_threshold);
  LSB_LAUNCH_CHECK(); lc.add();
  // K5
  k_pack_flags<<<G, TPB, 0, s>>>(label.p, N, scan_in.p, d_counts.p); LSB_LAUNCH_CHECK(); lc.add();
  tb = cub_tmp.cap;
  LSB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, scan_in.p, scan_out.p, N, s)); lc.add(2);
  k_scatter_keypoints<<<G, TPB, 0, s>>>(d_pts, src.p, label.p, scan_out.p, N, kp[0].p, kp[1].p, kp[2].p, d_counts.p);
  LSB_LAUNCH_CHECK(); lc.add();
  LSB_CUDA(cudaMemcpyAsync(h_counts.p, d_counts.p, 20 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  LSB_CUDA(cudaStreamSynchronize(s));
  for (int k = 0; k < 3; ++k) {
    counts[k] = h_counts.p[k];
    for (int e = 0; e < 2; ++e) {
      unsigned long long u;
      std::memcpy(&u, h_counts.p + 8 + 2 * (2 * k + e), 8);
      u = (u & 0x8000000000000000ull) ? (u & 0x7fffffffffffffffull) : ~u;
      std::memcpy(e ? &time_max[k] : &time_min[k], &u, 8);
    }
  }
  if (h_counts.p[3] > (uint32_t)KE_MAX_RINGS) throw CudaError(LSB_E_CAPACITY, "laser_id >= 1024 is not supported");
  if (h_counts.p[3] > nb_rings) nb_rings = h_counts.p[3];
}

void ExtractorGpu::debug_field(int field, float* d_out, cudaStream_t s, LaunchCounter& lc)
{
  if (n == 0) return;
  LSB_CUDA(cudaMemsetAsync(d_out, 0, n * sizeof(float), s));
  const float* f = nullptr;
  const uint8_t* bits = nullptr;
  int shift = 0;
  switch (field) {
    case 0: f = angles.p; break;
    case 1: f = saliency.p; break;
    case 2: f = depth_gap.p; break;
    case 3: f = intensity_gap.p; break;
    case 4: case 5: case 6: bits = label.p; shift = field - 4; break;
    case 7: case 8: case 9: bits = valid.p; shift = field - 7; break;
    default: throw CudaError(LSB_E_INVALID, "unknown extractor debug field");
  }
  k_debug_field<<<div_up(n, TPB), TPB, 0, s>>>(f, bits, shift, src.p, (int)n, d_out); LSB_LAUNCH_CHECK(); lc.add();
}

void ExtractorGpu::labels_input_order(uint8_t* d_out, cudaStream_t s, LaunchCounter& lc)
{
  if (n == 0) return;
  k_labels_input_order<<<div_up(n, TPB), TPB, 0, s>>>(label.p, src.p, (int)n, d_out); LSB_LAUNCH_CHECK(); lc.add();
}

}  // namespace lsb

// Exact kNN on a uniform grid — see knn.cuh and DESIGN.md §3 K8.  Hand-written search kernels;
// CUB is used only for the exclusive scan of the cell histogram.
//
// Cells of one (z, y) row are contiguous in the cell-major point array, so a row of cells is ONE
// index range (two cell_start loads), which makes empty space cheap.  After the cube of Chebyshev
// radius r around the query's cell is complete every unvisited point is farther than
//   bound(r) = r * cell + min over axes of the query's distance to its cell faces,
// so a query is settled as soon as its k-th best distance is strictly below bound(r) (minus a
// rounding margin) or the cube covers the whole grid: exact, with ties at equal d2 resolved by
// the smaller cloud index.  Cube pass (r = 1): k_knn (four lanes per query) or k_knn_stream (one
// thread per query, warp-synchronous insertions); unsettled queries: k_knn_warp (box of the
// sphere of the tentative k-th distance).
#include "knn.cuh"
#include <cub/cub.cuh>
#include <cfloat>
#include <climits>
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <atomic>

namespace lsb {

namespace {

constexpr int TPB = 256;
constexpr long long KNN_MAX_CELLS = 4ll << 20;

__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return (i >= 0) ? i : (i ^ 0x7fffffff); }
__host__ __device__ __forceinline__ float ord2f(int i) { const int j = (i >= 0) ? i : (i ^ 0x7fffffff);
#ifdef __CUDA_ARCH__
  return __int_as_float(j);
#else
  float f; memcpy(&f, &j, 4); return f;
#endif
}

__global__ void k_bbox_init(int* bb) { if (threadIdx.x < 3) bb[threadIdx.x] = INT_MAX; else if (threadIdx.x < 6) bb[threadIdx.x] = INT_MIN; }

__global__ void k_bbox(const lsb_point* __restrict__ pts, int n, int* __restrict__ bb)
{
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = reinterpret_cast<const float4*>(pts)[2 * i];
    const int v[3] = {f2ord(p.x), f2ord(p.y), f2ord(p.z)};
#pragma unroll
    for (int a = 0; a < 3; ++a) { mn[a] = min(mn[a], v[a]); mx[a] = max(mx[a], v[a]); }
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    for (int off = 16; off; off >>= 1) {
      mn[a] = min(mn[a], __shfl_xor_sync(0xffffffffu, mn[a], off));
      mx[a] = max(mx[a], __shfl_xor_sync(0xffffffffu, mx[a], off));
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(bb + a, mn[a]); atomicMax(bb + 3 + a, mx[a]); }
  }
}

__device__ __forceinline__ int cell_coord(float v, float o, float cell, int n)
{
  float f = floorf((v - o) / cell);
  f = fminf(fmaxf(f, 0.f), (float)(n - 1));
  return (int)f;
}

// Cell ordering as a counting sort (no radix sort): the histogram of the cell keys IS the cell_start table once scanned.
// Pass 1: cell key of every point, its arrival rank inside the cell (atomic counter per cell).  Pass 2 (after an exclusive
// scan of the counters): the point goes to cell_start[key] + rank.  The order of the points INSIDE a cell follows the
// atomics and is not reproducible; no result depends on it (every search orders its candidates by (d2, cloud index)).
// d_m != nullptr: the point count comes from device memory (launch sized for an upper bound before the host knew it)
__global__ void k_cell_count(const lsb_point* __restrict__ pts, int m, KnnGridParams gp, uint32_t* __restrict__ key, uint32_t* __restrict__ rank,
                             uint32_t* __restrict__ count, const uint32_t* __restrict__ d_m)
{
  if (d_m) m = (int)*d_m;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float4 p = reinterpret_cast<const float4*>(pts)[2 * i];
  const int cx = cell_coord(p.x, gp.ox, gp.cell, gp.nx), cy = cell_coord(p.y, gp.oy, gp.cell, gp.ny), cz = cell_coord(p.z, gp.oz, gp.cell, gp.nz);
  const uint32_t k = (uint32_t)((cz * gp.ny + cy) * gp.nx + cx);
  key[i] = k;
  rank[i] = atomicAdd(count + k, 1u);
}

// The histogram is left all zero again (its scan has been taken): the next build starts from a clean table without a
// memset of the whole grid (10 MB for the 2.5 M cells of a blob sub-map) on the frame's critical chain.
__global__ void k_cell_scatter(const lsb_point* __restrict__ pts, int m, const uint32_t* __restrict__ key, const uint32_t* __restrict__ rank,
                               const uint32_t* __restrict__ cell_start, float4* __restrict__ sorted, uint32_t* __restrict__ count,
                               const uint32_t* __restrict__ d_m)
{
  if (d_m) m = (int)*d_m;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float4 p = reinterpret_cast<const float4*>(pts)[2 * (size_t)i];
  const uint32_t k = key[i];
  sorted[__ldg(cell_start + k) + rank[i]] = make_float4(p.x, p.y, p.z, __int_as_float(i));
  count[k] = 0u;
}

// ------------------------------------------------------------------ search ---
// A candidate is one 64-bit key: (bits of d2) << 32 | cloud index.  d2 >= +0, so the unsigned
// order of the keys is the canonical order of the results: ascending d2, ties by smaller index.
typedef unsigned long long Key;
constexpr Key KEY_NONE = ((Key)0x7f800000u << 32) | 0x7fffffffu;     // (+inf, INT_MAX)
__device__ __forceinline__ Key make_key(float d2, int id) { return ((Key)__float_as_uint(d2) << 32) | (unsigned)id; }
__device__ __forceinline__ float key_d2(Key k) { return __uint_as_float((unsigned)(k >> 32)); }
__device__ __forceinline__ int key_id(Key k) { return (int)(unsigned)k; }

template <int KMAX>
struct TopK {
  Key key[KMAX];     // ascending
  __device__ __forceinline__ void init()
  {
#pragma unroll
    for (int t = 0; t < KMAX; ++t) key[t] = KEY_NONE;
  }
  __device__ __forceinline__ float kth() const { return key_d2(key[KMAX - 1]); }
  __device__ __forceinline__ void offer(Key x)
  {
    if (x < key[KMAX - 1]) {
#pragma unroll
      for (int t = 0; t < KMAX; ++t) {          // bubble: x carries the displaced larger key upward
        const bool lt = key[t] < x;
        const Key lo = lt ? key[t] : x;
        x = lt ? x : key[t];
        key[t] = lo;
      }
    }
  }
  __device__ __forceinline__ void pop()
  {
#pragma unroll
    for (int u = 0; u + 1 < KMAX; ++u) key[u] = key[u + 1];
    key[KMAX - 1] = KEY_NONE;
  }
};

// scans sorted[b, e): two candidates per trip, both loads issued before the first insertion
template <int KMAX>
__device__ __forceinline__ void scan_range(TopK<KMAX>& tk, const float4* __restrict__ sorted, uint32_t b, uint32_t e, float qx, float qy, float qz)
{
  uint32_t p = b;
  for (; p + 1 < e; p += 2) {
    const float4 s0 = __ldg(sorted + p), s1 = __ldg(sorted + p + 1);
    const float ax = qx - s0.x, ay = qy - s0.y, az = qz - s0.z;
    const float bx = qx - s1.x, by = qy - s1.y, bz = qz - s1.z;
    tk.offer(make_key((ax * ax + ay * ay) + az * az, __float_as_int(s0.w)));
    tk.offer(make_key((bx * bx + by * by) + bz * bz, __float_as_int(s1.w)));
  }
  if (p < e) {
    const float4 s0 = __ldg(sorted + p);
    const float ax = qx - s0.x, ay = qy - s0.y, az = qz - s0.z;
    tk.offer(make_key((ax * ax + ay * ay) + az * az, __float_as_int(s0.w)));
  }
}

struct QueryCell {
  float qx, qy, qz;
  int cx, cy, cz;
  float minoff;
  float ox, oy, oz;      // offset of the query inside its cell, in [0, cell]
  int rcover, rfirst;
  bool finite;
};

// The pose that moves the keypoints to the target's frame: a kernel argument, or (device-resident ICP loop, cuda/icp.cuh)
// the pose the previous solve left in device memory.  `done` set = the loop is over, the kernel returns at once.
struct KnnPose {
  Iso3d value;
  const Iso3d* dev;
  const int* done;
  const Interpolator* refine;   // k_knn_reuse only: undistortion refinement applied to the keypoint before it is searched
};
__device__ __forceinline__ bool knn_off(const KnnPose& kp) { return kp.done && *kp.done; }

template <bool FROM_POINTS>
__device__ __forceinline__ QueryCell make_query(const float* __restrict__ q_xyz, const lsb_point* __restrict__ kp, const KnnPose& pose, int i,
                                                const KnnGridParams& gp)
{
  QueryCell q;
  if (FROM_POINTS) {
    const float4 p = reinterpret_cast<const float4*>(kp)[2 * (size_t)i];
    const Iso3d prior = pose.dev ? *pose.dev : pose.value;
    const Vec3d w = apply(prior, Vec3d((double)p.x, (double)p.y, (double)p.z));
    q.qx = (float)w.x; q.qy = (float)w.y; q.qz = (float)w.z;
  } else {
    q.qx = q_xyz[3 * (size_t)i]; q.qy = q_xyz[3 * (size_t)i + 1]; q.qz = q_xyz[3 * (size_t)i + 2];
  }
  const float cell = gp.cell;
  // unclamped cell coordinates of the query (bounded so that the int conversion is defined)
  const float fx = fminf(fmaxf(floorf((q.qx - gp.ox) / cell), -1.0e6f), 1.0e6f);
  const float fy = fminf(fmaxf(floorf((q.qy - gp.oy) / cell), -1.0e6f), 1.0e6f);
  const float fz = fminf(fmaxf(floorf((q.qz - gp.oz) / cell), -1.0e6f), 1.0e6f);
  q.cx = (int)fx; q.cy = (int)fy; q.cz = (int)fz;
  const float lx = gp.ox + fx * cell, ly = gp.oy + fy * cell, lz = gp.oz + fz * cell;
  q.minoff = fminf(fminf(fminf(q.qx - lx, lx + cell - q.qx), fminf(q.qy - ly, ly + cell - q.qy)), fminf(q.qz - lz, lz + cell - q.qz));
  q.minoff = fmaxf(q.minoff, 0.f);
  q.ox = q.qx - lx; q.oy = q.qy - ly; q.oz = q.qz - lz;
  // radius at which the cube covers the whole grid / first touches it
  q.rcover = max(max(max(q.cx, gp.nx - 1 - q.cx), max(q.cy, gp.ny - 1 - q.cy)), max(q.cz, gp.nz - 1 - q.cz));
  q.rfirst = max(max(max(-q.cx, q.cx - (gp.nx - 1)), max(-q.cy, q.cy - (gp.ny - 1))), max(max(-q.cz, q.cz - (gp.nz - 1)), 0));
  q.finite = (q.qx == q.qx) && (q.qy == q.qy) && (q.qz == q.qz);
  return q;
}

// every point outside the completed cube of radius r is farther than this (minus a rounding margin)
__device__ __forceinline__ bool resolved_at(float kth_d2, int r, float cell, float minoff)
{
  if (!(kth_d2 < INFINITY)) return false;
  const float reach = (float)r * cell + minoff;
  const float bound = reach - (1.0e-3f + 1.0e-4f * (reach + cell));
  return bound > 0.f && kth_d2 < bound * bound;
}

// lower bound (squared, with a rounding margin) of the distance from the query to any point of
// a cell that lies d cells away along one axis (d = 0: the query's own cell)
__device__ __forceinline__ float axis_gap(int d, float off, float cell)
{
  if (d == 0) return 0.f;
  const float g = (d > 0) ? ((float)(d - 1) * cell + (cell - off)) : ((float)(-d - 1) * cell + off);
  return fmaxf(g - (1.0e-3f + 1.0e-4f * cell), 0.f);
}

// visiting order of the 27 cells of the 3x3x3 cube: the query's cell, the 6 face
// neighbours, the 12 edge neighbours, the 8 corners (packed (dz+1)*9 + (dy+1)*3 + (dx+1))
__constant__ unsigned char c_cube_order[28] = {13, 12, 14, 10, 16, 4, 22, 9, 11, 15, 17, 3, 5, 21, 23, 1, 7, 19, 25, 0, 2, 6, 8, 18, 20, 24, 26, 255};

// Phase 1: FOUR lanes per query (a thread per query leaves the GPU three quarters empty at
// 10^5 queries and serialises ~30 dependent cell visits per thread).  The 27 cells of the
// 3x3x3 cube around the query's cell are dealt nearest-first to the four lanes (lane s takes
// positions s, s+4, ... of c_cube_order); a cell whose lower distance bound is not below the
// smallest k-th distance held by the four lanes is skipped.  The four sorted lists are then
// merged (k rounds of a 2-step min).  The merged list is always written: for a query that the
// cube does not settle it is a tentative answer whose k-th distance bounds the true one, which
// is what the warp kernel below starts from.
template <int KMAX, bool FROM_POINTS>
__global__ void __launch_bounds__(128)
k_knn(const float* __restrict__ q_xyz, const lsb_point* __restrict__ kp, KnnPose prior, int nq, KnnGridParams gp,
      const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_start, int k, size_t qs, size_t ts, int32_t* __restrict__ idx_out,
      float* __restrict__ d2_out, int32_t* __restrict__ nfound_out, int32_t* __restrict__ unres_list, unsigned* __restrict__ unres_count)
{
  constexpr unsigned FULL = 0xffffffffu;
  if (knn_off(prior)) return;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int sub = t & 3;
  const bool live = (t >> 2) < nq;
  const int i = live ? (t >> 2) : nq - 1;      // idle quads of the last warp shadow the last query (shuffles need all lanes)
  const QueryCell q = make_query<FROM_POINTS>(q_xyz, kp, prior, i, gp);
  TopK<KMAX> tk;
  tk.init();
  const bool scan = q.finite && q.rfirst <= 1;   // NaN query: nothing found, as an exact search would
  const float gx[3] = {axis_gap(-1, q.ox, gp.cell), 0.f, axis_gap(1, q.ox, gp.cell)};
  const float gy[3] = {axis_gap(-1, q.oy, gp.cell), 0.f, axis_gap(1, q.oy, gp.cell)};
  const float gz[3] = {axis_gap(-1, q.oz, gp.cell), 0.f, axis_gap(1, q.oz, gp.cell)};
  // the lane's 7 cells: index ranges first (independent loads), then the scans
  uint32_t cb[7], ce[7];
  float lb[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    const int c = c_cube_order[sub + 4 * j];
    const int jz = c / 9, jy = (c / 3) % 3, jx = c % 3;
    const int z = q.cz + jz - 1, y = q.cy + jy - 1, x = q.cx + jx - 1;
    const bool in = scan && c != 255 && z >= 0 && z < gp.nz && y >= 0 && y < gp.ny && x >= 0 && x < gp.nx;
    const int cell_id = in ? (z * gp.ny + y) * gp.nx + x : 0;
    cb[j] = in ? __ldg(cell_start + cell_id) : 0u;
    ce[j] = in ? __ldg(cell_start + cell_id + 1) : 0u;
    lb[j] = in ? (gx[jx] * gx[jx] + gy[jy] * gy[jy]) + gz[jz] * gz[jz] : INFINITY;
  }
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    float thr = tk.kth();
    thr = fminf(thr, __shfl_xor_sync(FULL, thr, 1));
    thr = fminf(thr, __shfl_xor_sync(FULL, thr, 2));
    if (lb[j] < thr) scan_range(tk, sorted, cb[j], ce[j], q.qx, q.qy, q.qz);
    __syncwarp();
  }
  // merge: round r yields the r-th smallest key of the four lists
  int nf = 0;
  float kd = INFINITY;
#pragma unroll 1
  for (int r = 0; r < KMAX; ++r) {
    if (r >= k) break;
    Key m = tk.key[0];
    { const Key o = __shfl_xor_sync(FULL, m, 1); m = (o < m) ? o : m; }
    { const Key o = __shfl_xor_sync(FULL, m, 2); m = (o < m) ? o : m; }
    if (tk.key[0] == m && m != KEY_NONE) tk.pop();
    const bool ok = key_d2(m) < INFINITY;
    if (live && sub == 0) {
      idx_out[(size_t)i * qs + r * ts] = ok ? key_id(m) : -1;
      d2_out[(size_t)i * qs + r * ts] = ok ? key_d2(m) : INFINITY;
    }
    nf += ok ? 1 : 0;
    kd = key_d2(m);
  }
  if (!live || sub != 0) return;
  if (nfound_out) nfound_out[i] = nf;
  const bool done = !q.finite || (scan && ((q.rcover <= 1) || resolved_at(kd, 1, gp.cell, q.minoff)));
  if (!done) unres_list[atomicAdd(unres_count, 1u)] = i;
}

// Phase 1 for large query sets: ONE thread per query, every trip taken by the 32 lanes together.
// With a scan loop per cell a warp pays, cell after cell, for the longest cell among its lanes,
// and for a ~60-instruction sorted insertion on nearly every trip (some lane always inserts).
// Here each lane runs ONE candidate stream over the non-empty rows of the 3x3x3 cube (a row =
// three x-adjacent cells = one contiguous index range; own row first, a row skipped when its
// lower bound is not below the lane's k-th distance); a trip is branch-free (advance-row step,
// load, distance, compare) and STREAM_TRIPS candidates are loaded together; each is then offered
// in turn, the sorted insertion being executed only when some lane of the warp needs it (the
// other lanes bubble KEY_NONE through, unchanged).  Row ranges and bounds live in shared memory
// (per-lane dynamic index).
constexpr int STREAM_TRIPS = 4;
__constant__ unsigned char c_row_order[9] = {4, 1, 7, 3, 5, 0, 2, 6, 8};   // (dz+1)*3 + (dy+1), nearest first

template <int KMAX, bool FROM_POINTS>
__global__ void __launch_bounds__(128)
k_knn_stream(const float* __restrict__ q_xyz, const lsb_point* __restrict__ kp, KnnPose prior, int nq, KnnGridParams gp,
             const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_start, int k, size_t qs, size_t ts,
             int32_t* __restrict__ idx_out, float* __restrict__ d2_out, int32_t* __restrict__ nfound_out, int32_t* __restrict__ unres_list,
             unsigned* __restrict__ unres_count)
{
  constexpr unsigned FULL = 0xffffffffu;
  __shared__ uint32_t s_rb[9][128], s_re[9][128];
  __shared__ float s_rl[9][128];
  if (knn_off(prior)) return;
  const int tid = threadIdx.x;
  const int i0 = blockIdx.x * blockDim.x + tid;
  const bool live = i0 < nq;
  const int i = live ? i0 : nq - 1;
  const QueryCell q = make_query<FROM_POINTS>(q_xyz, kp, prior, i, gp);
  const bool scan = live && q.finite && q.rfirst <= 1;
  const float gy[3] = {axis_gap(-1, q.oy, gp.cell), 0.f, axis_gap(1, q.oy, gp.cell)};
  const float gz[3] = {axis_gap(-1, q.oz, gp.cell), 0.f, axis_gap(1, q.oz, gp.cell)};
  // the lane's non-empty rows, in visiting order
  uint32_t rb[9], re[9];
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const int r = c_row_order[t];
    const int z = q.cz + r / 3 - 1, y = q.cy + r % 3 - 1;
    const bool in = scan && z >= 0 && z < gp.nz && y >= 0 && y < gp.ny;
    const int base = (z * gp.ny + y) * gp.nx;
    const int xa = min(max(q.cx - 1, 0), gp.nx), xb = min(max(q.cx + 2, 0), gp.nx);   // clamped: out-of-grid cells are empty
    rb[t] = in ? __ldg(cell_start + base + xa) : 0u;
    re[t] = in ? __ldg(cell_start + base + xb) : 0u;
  }
  int nrow = 0;
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    const int r = c_row_order[t];
    if (rb[t] < re[t]) {
      s_rb[nrow][tid] = rb[t];
      s_re[nrow][tid] = re[t];
      s_rl[nrow][tid] = gy[r % 3] * gy[r % 3] + gz[r / 3] * gz[r / 3];
      ++nrow;
    }
  }
  TopK<KMAX> tk;
  tk.init();
  int o = 0;
  uint32_t p = 0, e = 0;
  while (true) {
    // one advance-row step (skips a row whose lower bound cannot beat the lane's k-th)
    const bool adv = p >= e && o < nrow;
    const int oo = adv ? o : 0;
    const bool take = adv && s_rl[oo][tid] < tk.kth();
    p = take ? s_rb[oo][tid] : p;
    e = take ? s_re[oo][tid] : e;
    o += adv ? 1 : 0;
    // STREAM_TRIPS candidates loaded together (one memory round trip); each is offered in turn, and
    // a sorted insertion is executed only if some lane of the warp needs it
    float4 sv[STREAM_TRIPS];
#pragma unroll
    for (int j = 0; j < STREAM_TRIPS; ++j) sv[j] = __ldg(sorted + ((p + j < e) ? p + j : 0u));
    Key kj[STREAM_TRIPS];
#pragma unroll
    for (int j = 0; j < STREAM_TRIPS; ++j) {
      const float ax = q.qx - sv[j].x, ay = q.qy - sv[j].y, az = q.qz - sv[j].z;
      kj[j] = (p + j < e) ? make_key((ax * ax + ay * ay) + az * az, __float_as_int(sv[j].w)) : KEY_NONE;
    }
    p = min(p + (uint32_t)STREAM_TRIPS, max(p, e));
#pragma unroll
    for (int j = 0; j < STREAM_TRIPS; ++j) {
      Key x = (kj[j] < tk.key[KMAX - 1]) ? kj[j] : KEY_NONE;
      if (__any_sync(FULL, x != KEY_NONE)) {
#pragma unroll
        for (int t = 0; t < KMAX; ++t) {
          const bool lt = tk.key[t] < x;
          const Key lo = lt ? tk.key[t] : x;
          x = lt ? x : tk.key[t];
          tk.key[t] = lo;
        }
      }
    }
    if (!__any_sync(FULL, p < e || o < nrow)) break;
  }
  if (!live) return;
  int nf = 0;
#pragma unroll
  for (int t = 0; t < KMAX; ++t) {
    if (t < k) {
      const bool ok = key_d2(tk.key[t]) < INFINITY;
      idx_out[(size_t)i * qs + t * ts] = ok ? key_id(tk.key[t]) : -1;
      d2_out[(size_t)i * qs + t * ts] = ok ? key_d2(tk.key[t]) : INFINITY;
      nf += ok ? 1 : 0;
    }
  }
  if (nfound_out) nfound_out[i] = nf;
  const bool done = !q.finite || (scan && ((q.rcover <= 1) || resolved_at(tk.kth(), 1, gp.cell, q.minoff)));
  if (!done) unres_list[atomicAdd(unres_count, 1u)] = i;
}

__device__ __forceinline__ void axis_cells(float lo, float hi, float o, float cell, int n, int& c0, int& c1);

// Cube pass of the LATER ICP iterations of a frame (KnnDeviceArgs::reuse): the target grid has not changed and the query
// has moved by millimetres, so the k neighbours found in the previous iteration, re-evaluated at the new query position,
// give an exact upper bound T of the new k-th distance before anything is searched.  With T known up front
//   * only the rows of the box of the sphere of radius sqrt(T) are visited (typically 2 x 2 rows of 2 cells instead of the
//     9 rows of 3 cells of the cube), rows whose lower bound exceeds T are dropped;
//   * no top-k list is kept while scanning: a candidate with d2 <= T is appended to a per-thread buffer in shared memory
//     (about k of them: T is the k-th distance up to the motion of the query), a predicated store instead of a sorted
//     insertion, and the k smallest are then placed by rank;
//   * the answer is exact by construction (every point within T was seen), so no query is left to the warp pass, except
//     the ones without a usable previous list (fewer than k found, "too far" results), a box of more than 16 rows or a
//     buffer overflow (many points at equal distance): those go to k_knn_warp with T as their threshold.
// One thread per query, trips taken by the 32 lanes together as in k_knn_stream.
template <int KMAX, bool FROM_POINTS>
__global__ void __launch_bounds__(128)
k_knn_reuse(const float* __restrict__ q_xyz, const lsb_point* __restrict__ kp, KnnPose prior, int nq, KnnGridParams gp,
            const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_start, const lsb_point* __restrict__ cloud, int k, size_t qs,
            size_t ts, int32_t* __restrict__ idx_out, float* __restrict__ d2_out, int32_t* __restrict__ nfound_out,
            int32_t* __restrict__ unres_list, unsigned* __restrict__ unres_count)
{
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int CAP = (KMAX >= 32) ? KMAX + 6 : KMAX + 8;   // 48 KB of static shared memory at most
  __shared__ Key s_buf[CAP][128];
  constexpr int MAXROWS = (KMAX >= 32) ? 9 : 16;          // rows of the sphere's box a thread takes on itself (shared memory budget)
  __shared__ uint32_t s_rb[MAXROWS][128], s_re[MAXROWS][128];
  if (knn_off(prior)) return;
  const int tid = threadIdx.x;
  const int i0 = blockIdx.x * blockDim.x + tid;
  const bool live = i0 < nq;
  const int i = live ? i0 : nq - 1;
  if (FROM_POINTS && prior.refine && live) {
    // Slam::RefineUndistortion of the previous ICP iteration (Slam.cxx:1345-1350) on this keypoint, in place: the same
    // arithmetic as k_undistort_state (one thread per keypoint there too), fused here to save a launch per type
    lsb_point* pt = const_cast<lsb_point*>(kp) + i;
    float4 v = reinterpret_cast<float4*>(pt)[0];
    const Iso3d H = prior.refine->at(pt->time);
    transform_point_f(H, v.x, v.y, v.z);
    reinterpret_cast<float4*>(pt)[0] = v;
  }
  const QueryCell q = make_query<FROM_POINTS>(q_xyz, kp, prior, i, gp);
  // the bound: largest distance from the moved query to its previous k neighbours
  bool ok = live && q.finite && (!nfound_out || nfound_out[i] >= k);
  float T = 0.f;
#pragma unroll
  for (int t = 0; t < KMAX; ++t) {
    if (t < k) {
      const int id = idx_out[(size_t)i * qs + t * ts];
      ok = ok && id >= 0;
      const float4 p = __ldg(reinterpret_cast<const float4*>(cloud) + 2 * (size_t)max(id, 0));
      const float ax = q.qx - p.x, ay = q.qy - p.y, az = q.qz - p.z;
      T = fmaxf(T, (ax * ax + ay * ay) + az * az);
    }
  }
  ok = ok && T < INFINITY;
  int x0 = 0, x1 = 0, y0 = 0, y1 = -1, z0 = 0, z1 = -1;
  if (ok) {
    const float r = sqrtf(T);
    const float rr = r + (1.0e-3f + 1.0e-4f * (r + gp.cell));
    axis_cells(q.qx - rr, q.qx + rr, gp.ox, gp.cell, gp.nx, x0, x1);
    axis_cells(q.qy - rr, q.qy + rr, gp.oy, gp.cell, gp.ny, y0, y1);
    axis_cells(q.qz - rr, q.qz + rr, gp.oz, gp.cell, gp.nz, z0, z1);
  }
  const int nyr = y1 - y0 + 1, rows = (z1 - z0 + 1) * nyr;
  const bool fits = ok && rows <= MAXROWS;
  int nrow = 0;
#pragma unroll
  for (int t = 0; t < MAXROWS; ++t) {
    if (fits && t < rows) {
      const int z = z0 + t / nyr, y = y0 + t % nyr;
      const float gy = axis_gap(y - q.cy, q.oy, gp.cell), gz = axis_gap(z - q.cz, q.oz, gp.cell);
      if (gy * gy + gz * gz <= T) {
        const int base = (z * gp.ny + y) * gp.nx;
        const uint32_t b = __ldg(cell_start + base + x0), e = __ldg(cell_start + base + x1 + 1);
        if (b < e) { s_rb[nrow][tid] = b; s_re[nrow][tid] = e; ++nrow; }
      }
    }
  }
  int cnt = 0, o = 0;
  uint32_t p = 0, e = 0;
  while (true) {
    const bool adv = p >= e && o < nrow;
    const int oo = adv ? o : 0;
    p = adv ? s_rb[oo][tid] : p;
    e = adv ? s_re[oo][tid] : e;
    o += adv ? 1 : 0;
    float4 sv[STREAM_TRIPS];
#pragma unroll
    for (int j = 0; j < STREAM_TRIPS; ++j) sv[j] = __ldg(sorted + ((p + j < e) ? p + j : 0u));
#pragma unroll
    for (int j = 0; j < STREAM_TRIPS; ++j) {
      const float ax = q.qx - sv[j].x, ay = q.qy - sv[j].y, az = q.qz - sv[j].z;
      const float d2 = (ax * ax + ay * ay) + az * az;
      if (p + j < e && d2 <= T) {
        if (cnt < CAP) s_buf[cnt][tid] = make_key(d2, __float_as_int(sv[j].w));
        ++cnt;
      }
    }
    p = min(p + (uint32_t)STREAM_TRIPS, max(p, e));
    if (!__any_sync(FULL, p < e || o < nrow)) break;
  }
  if (!live) return;
  if (!fits || cnt > CAP || cnt < k) {
    // handed to the warp pass: it reads its threshold from the k-th distance slot (an upper bound, or +inf for "none")
    if (q.finite) {
      // a large box (isolated query: the k-th neighbour is many cells away) is not worth one warp's scan: without a threshold
      // the warp pass takes its usual route for such queries (cube of the cap, then the shared brute-force kernel)
      d2_out[(size_t)i * qs + (size_t)(k - 1) * ts] = (ok && rows <= 64) ? T : INFINITY;
      unres_list[atomicAdd(unres_count, 1u)] = i;
    } else {
      for (int t = 0; t < k; ++t) { idx_out[(size_t)i * qs + t * ts] = -1; d2_out[(size_t)i * qs + t * ts] = INFINITY; }
      if (nfound_out) nfound_out[i] = 0;
    }
    return;
  }
  for (int j = 0; j < cnt; ++j) {
    const Key mine = s_buf[j][tid];
    int rank = 0;
    for (int u = 0; u < cnt; ++u) rank += (s_buf[u][tid] < mine) ? 1 : 0;
    if (rank < k) { idx_out[(size_t)i * qs + rank * ts] = key_id(mine); d2_out[(size_t)i * qs + rank * ts] = key_d2(mine); }
  }
  if (nfound_out) nfound_out[i] = k;
}

// merges the 32 per-lane sorted lists: round t yields the t-th smallest key of the warp
// (two warp-wide integer min reductions per round).  Writes the first k; returns the number
// found and the k-th distance.
template <int KMAX>
__device__ __forceinline__ int warp_merge_write(TopK<KMAX>& tk, int k, int lane, int32_t* __restrict__ idx_row, float* __restrict__ d2_row,
                                                size_t ts, float& kth_d2)
{
  constexpr unsigned FULL = 0xffffffffu;
  int nf = 0;
  kth_d2 = INFINITY;
#pragma unroll 1
  for (int t = 0; t < k; ++t) {
    const unsigned hi = (unsigned)(tk.key[0] >> 32), lo = (unsigned)tk.key[0];
    const unsigned mhi = __reduce_min_sync(FULL, hi);
    const unsigned mlo = __reduce_min_sync(FULL, (hi == mhi) ? lo : 0xffffffffu);
    const Key m = ((Key)mhi << 32) | mlo;
    if (tk.key[0] == m && m != KEY_NONE) tk.pop();     // the winning lane pops its head
    const bool ok = key_d2(m) < INFINITY;
    if (lane == 0) { idx_row[t * ts] = ok ? key_id(m) : -1; d2_row[t * ts] = ok ? key_d2(m) : INFINITY; }
    nf += ok ? 1 : 0;
    kth_d2 = key_d2(m);
  }
  return nf;
}

// cells of [lo, hi] along one axis, clamped to the grid (monotone in the same floorf((v - o) / cell)
// that binned the points, so every point with a coordinate in [lo, hi] lies in the returned cells)
__device__ __forceinline__ void axis_cells(float lo, float hi, float o, float cell, int n, int& c0, int& c1)
{
  c0 = (int)fminf(fmaxf(floorf((lo - o) / cell), 0.f), (float)(n - 1));
  c1 = (int)fminf(fmaxf(floorf((hi - o) / cell), 0.f), (float)(n - 1));
}

// The lanes of a warp share the rows of a box of cells (each row one contiguous index range) and
// keep per-lane top-K lists.  Rows that cannot hold anything nearer than the lane's k-th candidate,
// or than T, are skipped.
template <int KMAX>
__device__ __forceinline__ void warp_scan_box(TopK<KMAX>& tk, const QueryCell& q, const KnnGridParams& gp, const float4* __restrict__ sorted,
                                              const uint32_t* __restrict__ cell_start, int x0, int x1, int y0, int y1, int z0, int z1, float T,
                                              int lane)
{
  tk.init();
  const int nyr = y1 - y0 + 1, nrows = (z1 - z0 + 1) * nyr;
  if (!q.finite || x0 > x1 || nyr <= 0) return;
  for (int t = lane; t < nrows; t += 32) {
    const int z = z0 + t / nyr, y = y0 + t % nyr;
    const float gy = axis_gap(y - q.cy, q.oy, gp.cell), gz = axis_gap(z - q.cz, q.oz, gp.cell);
    const float lb = gy * gy + gz * gz;
    if (!(lb < tk.kth()) || lb > T) continue;
    const int base = (z * gp.ny + y) * gp.nx;
    scan_range(tk, sorted, __ldg(cell_start + base + x0), __ldg(cell_start + base + x1 + 1), q.qx, q.qy, q.qz);
  }
}

// Phase 2: one warp per query that the 3x3x3 cube did not settle.
// (a) With a threshold T such that every point of the answer has d2 <= T (the k-th distance of
//     phase 1's tentative list; capped mode: at most the cap) no top-k list is needed while
//     scanning: the lanes share the rows of the box of the sphere of radius sqrt(T) (a row = one
//     contiguous index range), every point with d2 <= T is appended to a per-warp buffer in
//     shared memory (a few dozen keys: about k of them when T is the tentative k-th distance), and
//     the k smallest are then placed by rank (each lane counts the keys smaller than its own):
//     sorted output without any sorted insertion.  Exact by construction; in capped mode fewer
//     than k keys within the cap means "k-th neighbour beyond the cap", all the caller needs.
//     A buffer overflow (dense box under a loose cap) falls back to per-lane lists.
// (b) Exact mode without a threshold (fewer than k points in the cube): cube of radius rcap with
//     per-lane lists; a query that even this does not settle goes to the brute-force kernel.
constexpr int WARP_BUF = 128;

template <int KMAX, bool FROM_POINTS>
__global__ void __launch_bounds__(128)
k_knn_warp(const float* __restrict__ q_xyz, const lsb_point* __restrict__ kp, KnnPose prior, KnnGridParams gp,
           const float4* __restrict__ sorted, const uint32_t* __restrict__ cell_start, int m, int k, int rcap, int capped, float cap2,
           size_t qs, size_t ts, int32_t* __restrict__ idx_out, float* __restrict__ d2_out, int32_t* __restrict__ nfound_out,
           const int32_t* __restrict__ unres_list, unsigned* __restrict__ unres_count, int32_t* __restrict__ bf_list, unsigned bf_cap)
{
  __shared__ Key s_keys[4][WARP_BUF];
  __shared__ int s_cnt[4];
  if (knn_off(prior)) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const unsigned count = unres_count[0];
  while (true) {
    unsigned slot = 0;
    if (lane == 0) slot = atomicAdd(unres_count + 1, 1u);
    slot = __shfl_sync(0xffffffffu, slot, 0);
    if (slot >= count) break;
    const int i = unres_list[slot];
    const QueryCell q = make_query<FROM_POINTS>(q_xyz, kp, prior, i, gp);
    int32_t* idx_row = idx_out + (size_t)i * qs;
    float* d2_row = d2_out + (size_t)i * qs;
    const float tentative = d2_row[(size_t)(k - 1) * ts];
    float T = tentative;
    if (capped) T = fminf(T, cap2);
    int x0, x1, y0, y1, z0, z1, nf = 0;
    TopK<KMAX> tk;
    float kd;
    if (T < INFINITY) {
      // without a tentative k-th distance T is the cap, a loose bound: try the sphere of two cells first
      const float T2 = 4.f * gp.cell * gp.cell;
      const bool loose = !(tentative < INFINITY) && T2 < T;
      bool placed = false;
      for (int stage = loose ? 0 : 1; stage < 2 && !placed; ++stage) {
        const float Ts = (stage == 0) ? T2 : T;
        const float r = sqrtf(Ts);
        const float rr = r + (1.0e-3f + 1.0e-4f * (r + gp.cell));
        axis_cells(q.qx - rr, q.qx + rr, gp.ox, gp.cell, gp.nx, x0, x1);
        axis_cells(q.qy - rr, q.qy + rr, gp.oy, gp.cell, gp.ny, y0, y1);
        axis_cells(q.qz - rr, q.qz + rr, gp.oz, gp.cell, gp.nz, z0, z1);
        const int nyr = y1 - y0 + 1, nrows = (z1 - z0 + 1) * nyr;
        if (lane == 0) s_cnt[w] = 0;
        __syncwarp();
        for (int t = lane; t < nrows; t += 32) {
          const int z = z0 + t / nyr, y = y0 + t % nyr;
          const float gy = axis_gap(y - q.cy, q.oy, gp.cell), gz = axis_gap(z - q.cz, q.oz, gp.cell);
          if (gy * gy + gz * gz > Ts) continue;       // nothing of this row can be within Ts
          const int base = (z * gp.ny + y) * gp.nx;
          const uint32_t b = __ldg(cell_start + base + x0), e = __ldg(cell_start + base + x1 + 1);
          for (uint32_t p = b; p < e; p += 2) {
            const bool two = p + 1 < e;
            const float4 s0 = __ldg(sorted + p), s1 = __ldg(sorted + (two ? p + 1 : p));
            const float ax = q.qx - s0.x, ay = q.qy - s0.y, az = q.qz - s0.z;
            const float bx = q.qx - s1.x, by = q.qy - s1.y, bz = q.qz - s1.z;
            const float da = (ax * ax + ay * ay) + az * az, db = (bx * bx + by * by) + bz * bz;
            if (da <= Ts) { const int o = atomicAdd(&s_cnt[w], 1); if (o < WARP_BUF) s_keys[w][o] = make_key(da, __float_as_int(s0.w)); }
            if (two && db <= Ts) { const int o = atomicAdd(&s_cnt[w], 1); if (o < WARP_BUF) s_keys[w][o] = make_key(db, __float_as_int(s1.w)); }
          }
        }
        __syncwarp();
        const int C = s_cnt[w];
        if (stage == 0 && C < k) continue;            // fewer than k points within two cells: the cap decides
        if (C > WARP_BUF) {
          if (stage == 0) continue;
          break;                                      // dense box under a loose cap: per-lane lists below
        }
        for (int j = lane; j < C; j += 32) {
          const Key mine = s_keys[w][j];
          int rank = 0;
          for (int u = 0; u < C; ++u) rank += (s_keys[w][u] < mine) ? 1 : 0;
          if (rank < k) { idx_row[(size_t)rank * ts] = key_id(mine); d2_row[(size_t)rank * ts] = key_d2(mine); }
        }
        for (int r2 = C + lane; r2 < k; r2 += 32) { idx_row[(size_t)r2 * ts] = -1; d2_row[(size_t)r2 * ts] = INFINITY; }
        nf = min(C, k);
        // fewer than k points within the cap: the k-th neighbour is farther than the cap
        if (C < k && m >= k) nf = k;
        placed = true;
        __syncwarp();
      }
      if (!placed) {
        warp_scan_box<KMAX>(tk, q, gp, sorted, cell_start, x0, x1, y0, y1, z0, z1, T, lane);
        nf = warp_merge_write<KMAX>(tk, k, lane, idx_row, d2_row, ts, kd);
        if (!(T == tentative || kd <= cap2) && m >= k) nf = k;
      }
    } else {
      const int R = max(rcap, 1);
      z0 = max(0, q.cz - R); z1 = min(gp.nz - 1, q.cz + R);
      y0 = max(0, q.cy - R); y1 = min(gp.ny - 1, q.cy + R);
      x0 = max(0, q.cx - R); x1 = min(gp.nx - 1, q.cx + R);
      warp_scan_box<KMAX>(tk, q, gp, sorted, cell_start, x0, x1, y0, y1, z0, z1, INFINITY, lane);
      nf = warp_merge_write<KMAX>(tk, k, lane, idx_row, d2_row, ts, kd);
      const bool settled = (q.rcover <= R) || resolved_at(kd, R, gp.cell, q.minoff);
      if (!settled) {            // isolated query: brute force, shared by many warps (k_knn_brute)
        unsigned o = 0;
        if (lane == 0) o = atomicAdd(unres_count + 2, 1u);
        o = __shfl_sync(0xffffffffu, o, 0);
        if (o < bf_cap) {
          if (lane == 0) bf_list[o] = i;
          continue;
        }
        tk.init();               // list full: this warp scans the whole cloud itself
        for (int p = lane; p < m; p += 32) {
          const float4 sp = __ldg(sorted + p);
          const float dx = q.qx - sp.x, dy = q.qy - sp.y, dz = q.qz - sp.z;
          tk.offer(make_key((dx * dx + dy * dy) + dz * dz, __float_as_int(sp.w)));
        }
        nf = warp_merge_write<KMAX>(tk, k, lane, idx_row, d2_row, ts, kd);
      }
    }
    if (lane == 0 && nfound_out) nfound_out[i] = nf;
  }
}

// Brute force for the isolated queries of an exact search: a work item is (query, segment of the
// cloud); the warp keeps per-lane lists over its segment, merges them, and writes a sorted partial
// list; the warp that completes a query's last segment merges the BF_SEGS partial lists.
constexpr int BF_SEGS = 16;

template <int KMAX, bool FROM_POINTS>
__global__ void __launch_bounds__(128)
k_knn_brute(const float* __restrict__ q_xyz, const lsb_point* __restrict__ kp, KnnPose prior, KnnGridParams gp,
            const float4* __restrict__ sorted, int m, int k, size_t qs, size_t ts, int32_t* __restrict__ idx_out, float* __restrict__ d2_out,
            int32_t* __restrict__ nfound_out, const int32_t* __restrict__ bf_list, unsigned* __restrict__ unres_count,
            Key* __restrict__ partial, unsigned* __restrict__ tickets, unsigned bf_cap)
{
  constexpr unsigned FULL = 0xffffffffu;
  if (knn_off(prior)) return;
  const int lane = threadIdx.x & 31;
  const unsigned nbf = min(unres_count[2], bf_cap);
  const unsigned nitems = nbf * BF_SEGS;
  const int seg_len = (m + BF_SEGS - 1) / BF_SEGS;
  while (true) {
    unsigned item = 0;
    if (lane == 0) item = atomicAdd(unres_count + 3, 1u);
    item = __shfl_sync(FULL, item, 0);
    if (item >= nitems) break;
    const unsigned qslot = item / BF_SEGS, seg = item % BF_SEGS;
    const int i = bf_list[qslot];
    const QueryCell q = make_query<FROM_POINTS>(q_xyz, kp, prior, i, gp);
    TopK<KMAX> tk;
    tk.init();
    const int p0 = (int)seg * seg_len, p1 = min(m, p0 + seg_len);
    for (int p = p0 + lane; p < p1; p += 32) {
      const float4 s = __ldg(sorted + p);
      const float dx = q.qx - s.x, dy = q.qy - s.y, dz = q.qz - s.z;
      tk.offer(make_key((dx * dx + dy * dy) + dz * dz, __float_as_int(s.w)));
    }
    Key* out = partial + ((size_t)qslot * BF_SEGS + seg) * k;
    for (int t = 0; t < k; ++t) {          // warp merge of the per-lane lists -> sorted partial list
      const unsigned hi = (unsigned)(tk.key[0] >> 32), lo = (unsigned)tk.key[0];
      const unsigned mhi = __reduce_min_sync(FULL, hi);
      const unsigned mlo = __reduce_min_sync(FULL, (hi == mhi) ? lo : 0xffffffffu);
      const Key mk = ((Key)mhi << 32) | mlo;
      if (tk.key[0] == mk && mk != KEY_NONE) tk.pop();
      if (lane == 0) out[t] = mk;
    }
    __threadfence();
    unsigned done = 0;
    if (lane == 0) done = atomicAdd(tickets + qslot, 1u);
    done = __shfl_sync(FULL, done, 0);
    if (done != BF_SEGS - 1) continue;
    __threadfence();
    // last segment of this query: merge the BF_SEGS sorted lists (lane s < BF_SEGS walks list s)
    const Key* lists = partial + (size_t)qslot * BF_SEGS * k;
    int pos = 0, nf = 0;
    for (int t = 0; t < k; ++t) {
      const Key head = (lane < BF_SEGS && pos < k) ? __ldcg(lists + (size_t)lane * k + pos) : KEY_NONE;
      const unsigned hi = (unsigned)(head >> 32), lo = (unsigned)head;
      const unsigned mhi = __reduce_min_sync(FULL, hi);
      const unsigned mlo = __reduce_min_sync(FULL, (hi == mhi) ? lo : 0xffffffffu);
      const Key mk = ((Key)mhi << 32) | mlo;
      if (head == mk && mk != KEY_NONE) ++pos;
      const bool ok = key_d2(mk) < INFINITY;
      if (lane == 0) { idx_out[(size_t)i * qs + t * ts] = ok ? key_id(mk) : -1; d2_out[(size_t)i * qs + t * ts] = ok ? key_d2(mk) : INFINITY; }
      nf += ok ? 1 : 0;
    }
    if (lane == 0) { if (nfound_out) nfound_out[i] = nf; tickets[qslot] = 0u; }
  }
}

template <bool FROM_POINTS>
void launch_knn(const KnnGrid& g, const float* q, const lsb_point* kp, const Iso3d& prior_host, size_t nq, int k, float cap_dist, int32_t* idx,
                float* d2, int32_t* nf, bool entry_major, cudaStream_t s, LaunchCounter& lc, const KnnDeviceArgs* da)
{
  KnnPose prior;
  prior.value = prior_host;
  prior.dev = da ? da->prior : nullptr;
  prior.done = da ? da->done : nullptr;
  prior.refine = (da && da->reuse) ? da->refine : nullptr;
  const size_t qs = entry_major ? 1 : (size_t)k, ts = entry_major ? nq : 1;
  const unsigned G = div_up(nq * 4, 128);   // four lanes per query
  const int NQ = (int)nq, M = (int)g.m;
  const int capped = cap_dist >= 0.f ? 1 : 0;
  // radius of the warp kernel's cube: capped => everything within cap_dist is inside it
  // (r*cell minus the rounding margin exceeds cap_dist); exact => 5 m, then brute force
  const float reach = capped ? cap_dist : 5.0f;
  const int rcap = std::max(2, (int)std::ceil(reach / g.gp.cell) + 1);
  const float cap2 = capped ? cap_dist * cap_dist : INFINITY;   // rows farther than the cap cannot matter
  // unres_count: [0] unsettled queries, [1] cursor, [2] brute-force queries, [3] cursor
  g.unres_list.reserve(nq, s);
  if (!g.unres_count.p) { g.unres_count.reserve(4, s); g.counters_clean = false; }
  if (!g.counters_clean) LSB_CUDA(cudaMemsetAsync(g.unres_count.p, 0, 4 * sizeof(unsigned), s));
  g.counters_clean = false;
  const unsigned bf_cap = (unsigned)std::min<size_t>(nq, 16384);   // more isolated queries than that scan the cloud on their own
  if (!capped) {
    g.bf_list.reserve(bf_cap, s);
    g.bf_partial.reserve((size_t)bf_cap * BF_SEGS * (size_t)k, s);
    if (g.bf_tickets.cap < bf_cap) { g.bf_tickets.reserve(bf_cap, s); LSB_CUDA(cudaMemsetAsync(g.bf_tickets.p, 0, g.bf_tickets.cap * sizeof(unsigned), s)); }
  }
  // few queries: four lanes per query (fills the SMs, short per-lane chains); many queries: one
  // thread per query with warp-synchronous insertions (fewest instructions)
  static const long quad_max = [] { const char* e = std::getenv("LSB_KNN_QUAD_MAX"); return e ? std::atol(e) : 148L * 64; }();   // tuning hook
  const bool quad = (long)nq <= quad_max;
  const unsigned G1 = div_up(nq, 128);
  const bool reuse = da && da->reuse && knn_reuse_enabled() && nf != nullptr;
#define LSB_KNN_CASE(KM)                                                                                                                   \
  do {                                                                                                                                     \
    if (reuse)                                                                                                                             \
      k_knn_reuse<KM, FROM_POINTS><<<G1, 128, 0, s>>>(q, kp, prior, NQ, g.gp, g.sorted.p, g.cell_start.p, g.cloud, k, qs, ts, idx, d2, nf,   \
                                                      g.unres_list.p, g.unres_count.p);                                                    \
    else if (quad)                                                                                                                         \
      k_knn<KM, FROM_POINTS><<<G, 128, 0, s>>>(q, kp, prior, NQ, g.gp, g.sorted.p, g.cell_start.p, k, qs, ts, idx, d2, nf, g.unres_list.p,  \
                                               g.unres_count.p);                                                                           \
    else                                                                                                                                   \
      k_knn_stream<KM, FROM_POINTS><<<G1, 128, 0, s>>>(q, kp, prior, NQ, g.gp, g.sorted.p, g.cell_start.p, k, qs, ts, idx, d2, nf,          \
                                                       g.unres_list.p, g.unres_count.p);                                                   \
    LSB_LAUNCH_CHECK(); lc.add();                                                                                                          \
    k_knn_warp<KM, FROM_POINTS><<<148 * 8, 128, 0, s>>>(q, kp, prior, g.gp, g.sorted.p, g.cell_start.p, M, k, rcap, capped, cap2, qs, ts, idx, d2, nf, \
                                                        g.unres_list.p, g.unres_count.p, g.bf_list.p, bf_cap);                             \
    LSB_LAUNCH_CHECK(); lc.add();                                                                                                          \
    if (!capped) {                                                                                                                         \
      k_knn_brute<KM, FROM_POINTS><<<148 * 8, 128, 0, s>>>(q, kp, prior, g.gp, g.sorted.p, M, k, qs, ts, idx, d2, nf, g.bf_list.p,          \
                                                           g.unres_count.p, reinterpret_cast<Key*>(g.bf_partial.p), g.bf_tickets.p, bf_cap); \
      LSB_LAUNCH_CHECK(); lc.add();                                                                                                        \
    }                                                                                                                                      \
  } while (0)
  static const bool dbg = std::getenv("LSB_KNN_DEBUG") != nullptr;
  if (k <= 1) LSB_KNN_CASE(1);
  else if (k <= 5) LSB_KNN_CASE(5);
  else if (k <= 8) LSB_KNN_CASE(8);
  else if (k <= 10) LSB_KNN_CASE(10);
  else if (k <= 16) LSB_KNN_CASE(16);
  else LSB_KNN_CASE(32);
#undef LSB_KNN_CASE
  if (dbg) {      // diagnostic only (synchronises): how many queries the first kernel left to the warp pass / brute force
    unsigned h[4];
    LSB_CUDA(cudaMemcpyAsync(h, g.unres_count.p, sizeof(h), cudaMemcpyDeviceToHost, s));
    LSB_CUDA(cudaStreamSynchronize(s));
    std::fprintf(stderr, "[lsb knn] nq %zu k %d m %zu cell %.2f %s unresolved %u brute %u\n", nq, k, g.m, g.gp.cell, reuse ? "reuse" : (quad ? "quad" : "stream"), h[0], h[2]);
  }
}


__global__ void k_fill_empty(int32_t* idx, float* d2, int32_t* nf, size_t total, size_t nq)
{
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) { idx[i] = -1; d2[i] = INFINITY; }
  if (nf && i < nq) nf[i] = 0;
}

}  // namespace

bool knn_reuse_enabled()
{
  static const bool on = [] { const char* e = std::getenv("LSB_KNN_REUSE"); return !e || e[0] != '0'; }();   // tuning / A-B hook
  return on;
}

void KnnGrid::build(const lsb_point* d_cloud, size_t m_, float cell, const float* lo, const float* hi, cudaStream_t s, LaunchCounter& lc,
                    const uint32_t* d_m)
{
  if (d_m && !(lo && hi)) throw CudaError(LSB_E_INVALID, "KnnGrid::build: a device-side count needs host-known bounds");
  cloud = d_cloud;
  m = m_;
  if (m == 0) return;
  const int M = (int)m;
  float bmin[3], bmax[3];
  if (lo && hi) {
    for (int a = 0; a < 3; ++a) { bmin[a] = lo[a]; bmax[a] = hi[a]; }
  } else {
    d_bbox.reserve(8, s); h_bbox.reserve(8);
    k_bbox_init<<<1, 32, 0, s>>>((int*)d_bbox.p); LSB_LAUNCH_CHECK(); lc.add();
    k_bbox<<<std::min(div_up(m, TPB), 592u), TPB, 0, s>>>(d_cloud, M, (int*)d_bbox.p); LSB_LAUNCH_CHECK(); lc.add();
    LSB_CUDA(cudaMemcpyAsync(h_bbox.p, d_bbox.p, 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
    LSB_CUDA(cudaStreamSynchronize(s));
    for (int a = 0; a < 3; ++a) {
      int v; memcpy(&v, h_bbox.p + a, 4); bmin[a] = ord2f(v);
      memcpy(&v, h_bbox.p + 3 + a, 4); bmax[a] = ord2f(v);
    }
  }
  for (int a = 0; a < 3; ++a) {
    if (!(bmin[a] <= bmax[a]) || !std::isfinite(bmin[a]) || !std::isfinite(bmax[a])) throw CudaError(LSB_E_INVALID, "kNN target cloud has non-finite coordinates");
  }
  // one cell of margin on each side; enlarge the cell until the grid fits the budget
  float c = cell;
  long long nx, ny, nz;
  while (true) {
    nx = (long long)std::floor((bmax[0] - bmin[0]) / c) + 3;
    ny = (long long)std::floor((bmax[1] - bmin[1]) / c) + 3;
    nz = (long long)std::floor((bmax[2] - bmin[2]) / c) + 3;
    if (nx * ny * nz <= KNN_MAX_CELLS) break;
    c *= 1.25f;
  }
  gp.cell = c;
  gp.ox = bmin[0] - c; gp.oy = bmin[1] - c; gp.oz = bmin[2] - c;
  gp.nx = (int)nx; gp.ny = (int)ny; gp.nz = (int)nz;
  const int ncells = gp.nx * gp.ny * gp.nz;
  key_a.reserve(m, s); idx_a.reserve(m, s);
  const int nc1 = ncells + 1;
  sorted.reserve(m, s); cell_start.reserve((size_t)nc1, s);
  if ((size_t)nc1 > cell_count.cap) {      // invariant: the whole table is zero between builds (k_cell_scatter cleans up behind itself)
    cell_count.reserve((size_t)nc1, s);
    LSB_CUDA(cudaMemsetAsync(cell_count.p, 0, cell_count.cap * sizeof(uint32_t), s));
  }
  const unsigned G = div_up(m, TPB);
  k_cell_count<<<G, TPB, 0, s>>>(d_cloud, M, gp, key_a.p, idx_a.p, cell_count.p, d_m); LSB_LAUNCH_CHECK(); lc.add();
  // (a single-pass scan of our own - select.cuh - was measured here: one launch less, but 26 instead of 15 us for the 2.5 M cells
  //  of a blob sub-map, and 1 - 3 % fewer frames per second)
  size_t tb = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tb, cell_count.p, cell_start.p, nc1, s);
  cub_tmp.reserve(tb + 256, s);
  tb = cub_tmp.cap;
  LSB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, cell_count.p, cell_start.p, nc1, s)); lc.add(2);   // cell_start[ncells] = m
  k_cell_scatter<<<G, TPB, 0, s>>>(d_cloud, M, key_a.p, idx_a.p, cell_start.p, sorted.p, cell_count.p, d_m); LSB_LAUNCH_CHECK(); lc.add();
}

void KnnGrid::search(const float* d_q, size_t nq, int k, float cap_dist, int32_t* d_idx, float* d_d2, int32_t* d_nf, bool entry_major, cudaStream_t s, LaunchCounter& lc) const
{
  if (nq == 0) return;
  if (k < 1 || k > KNN_MAX_K) throw CudaError(LSB_E_INVALID, "k must be in [1, 32]");
  if (m == 0) {
    k_fill_empty<<<div_up(nq * k, TPB), TPB, 0, s>>>(d_idx, d_d2, d_nf, nq * k, nq); LSB_LAUNCH_CHECK(); lc.add();
    return;
  }
  launch_knn<false>(*this, d_q, nullptr, Iso3d::identity(), nq, k, cap_dist, d_idx, d_d2, d_nf, entry_major, s, lc, nullptr);
}

void KnnGrid::search_points(const lsb_point* d_kp, size_t nq, const Iso3d& prior, int k, float cap_dist, int32_t* d_idx, float* d_d2,
                            int32_t* d_nf, bool entry_major, cudaStream_t s, LaunchCounter& lc, const KnnDeviceArgs* da) const
{
  if (nq == 0) return;
  if (k < 1 || k > KNN_MAX_K) throw CudaError(LSB_E_INVALID, "k must be in [1, 32]");
  if (m == 0) {
    k_fill_empty<<<div_up(nq * k, TPB), TPB, 0, s>>>(d_idx, d_d2, d_nf, nq * k, nq); LSB_LAUNCH_CHECK(); lc.add();
    return;
  }
  launch_knn<true>(*this, nullptr, d_kp, prior, nq, k, cap_dist, d_idx, d_d2, d_nf, entry_major, s, lc, da);
}

}  // namespace lsb

// ====================================================================================================
// Exact kNN that allocates nothing (lsb_knn_search_noalloc, include/lidarslam_b200.h).  No counterpart in the
// reference: added for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  It is the brute-force kernel of the exact search above
// (k_knn_brute: a work item = (query, segment of the cloud), per-lane lists, warp merge, partial sorted lists, the
// warp that completes a query's last segment merges them) reading the caller's cloud directly, with its scratch —
// the partial lists and the per-query tickets — in ONE file-scope __device__ array instead of DevBufs:
//   * every GPU that loads the library holds NOALLOC_SCRATCH_BYTES (4 MiB + 8 KiB) until the process exits;
//   * the limits that size it: nq <= LSB_NOALLOC_MAX_QUERIES (2048), k <= LSB_NOALLOC_MAX_K (16); larger -> LSB_E_CAPACITY;
//   * every call on a device shares the array and finds the previous call's contents in it: the tickets are cleared by
//     a memset node at the head of each call (the partial lists are fully rewritten before they are read);
//   * two calls that overlap on the GPU (different streams, no dependency) race on it.
// Host code must not pass the array's name to a kernel (on the host it is a stand-in, not a device address): the
// address is fetched once per device with cudaGetSymbolAddress, with that device current.
namespace lsb {

namespace {

constexpr int NOALLOC_SEGS = BF_SEGS;
constexpr size_t NOALLOC_PARTIAL_KEYS = (size_t)LSB_NOALLOC_MAX_QUERIES * NOALLOC_SEGS * LSB_NOALLOC_MAX_K;
constexpr size_t NOALLOC_SCRATCH_BYTES = NOALLOC_PARTIAL_KEYS * sizeof(Key) + (size_t)LSB_NOALLOC_MAX_QUERIES * sizeof(unsigned);

static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

__global__ void __launch_bounds__(128)
k_knn_brute_static(const lsb_point* __restrict__ cloud, int m, const float* __restrict__ q_xyz, int nq, int k, int32_t* __restrict__ idx_out,
                   float* __restrict__ d2_out, Key* __restrict__ partial, unsigned* __restrict__ tickets)
{
  constexpr unsigned FULL = 0xffffffffu;
  constexpr int KMAX = LSB_NOALLOC_MAX_K;
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const int seg_len = (m + NOALLOC_SEGS - 1) / NOALLOC_SEGS;
  const int nitems = nq * NOALLOC_SEGS;
  for (int item = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; item < nitems; item += warps) {
    const int i = item / NOALLOC_SEGS, seg = item % NOALLOC_SEGS;
    const float qx = q_xyz[3 * (size_t)i], qy = q_xyz[3 * (size_t)i + 1], qz = q_xyz[3 * (size_t)i + 2];
    TopK<KMAX> tk;
    tk.init();
    const int p0 = seg * seg_len, p1 = min(m, p0 + seg_len);
    for (int p = p0 + lane; p < p1; p += 32) {
      const float4 s = __ldg(reinterpret_cast<const float4*>(cloud) + 2 * (size_t)p);
      const float dx = qx - s.x, dy = qy - s.y, dz = qz - s.z;
      tk.offer(make_key((dx * dx + dy * dy) + dz * dz, p));      // same arithmetic and tie order as the grid search
    }
    Key* out = partial + ((size_t)i * NOALLOC_SEGS + seg) * k;
    for (int t = 0; t < k; ++t) {          // warp merge of the per-lane lists -> sorted partial list
      const unsigned hi = (unsigned)(tk.key[0] >> 32), lo = (unsigned)tk.key[0];
      const unsigned mhi = __reduce_min_sync(FULL, hi);
      const unsigned mlo = __reduce_min_sync(FULL, (hi == mhi) ? lo : 0xffffffffu);
      const Key mk = ((Key)mhi << 32) | mlo;
      if (tk.key[0] == mk && mk != KEY_NONE) tk.pop();
      if (lane == 0) out[t] = mk;
    }
    __threadfence();
    unsigned done = 0;
    if (lane == 0) done = atomicAdd(tickets + i, 1u);
    done = __shfl_sync(FULL, done, 0);
    if (done != NOALLOC_SEGS - 1) continue;
    __threadfence();
    // last segment of this query: merge the sorted lists (lane s < NOALLOC_SEGS walks list s)
    const Key* lists = partial + (size_t)i * NOALLOC_SEGS * k;
    int pos = 0;
    for (int t = 0; t < k; ++t) {
      const Key head = (lane < NOALLOC_SEGS && pos < k) ? __ldcg(lists + (size_t)lane * k + pos) : KEY_NONE;
      const unsigned hi = (unsigned)(head >> 32), lo = (unsigned)head;
      const unsigned mhi = __reduce_min_sync(FULL, hi);
      const unsigned mlo = __reduce_min_sync(FULL, (hi == mhi) ? lo : 0xffffffffu);
      const Key mk = ((Key)mhi << 32) | mlo;
      if (head == mk && mk != KEY_NONE) ++pos;
      const bool ok = key_d2(mk) < INFINITY;
      if (lane == 0) { idx_out[(size_t)i * k + t] = ok ? key_id(mk) : -1; d2_out[(size_t)i * k + t] = ok ? key_d2(mk) : INFINITY; }
    }
  }
}

// device address of g_noalloc_scratch, fetched once per device (index = CUDA device ordinal)
constexpr int NOALLOC_MAX_DEVICES = 64;
std::atomic<unsigned char*> g_noalloc_ptr[NOALLOC_MAX_DEVICES];

}  // namespace

void knn_search_noalloc(int device, const lsb_point* d_cloud, size_t m, const float* d_queries_xyz, size_t nq, int k, int32_t* d_idx_out,
                        float* d_d2_out, cudaStream_t s)
{
  if (device < 0 || device >= NOALLOC_MAX_DEVICES) throw CudaError(LSB_E_INVALID, "lsb_knn_search_noalloc: invalid CUDA device index");
  if (k < 1 || k > LSB_NOALLOC_MAX_K) throw CudaError(LSB_E_CAPACITY, "lsb_knn_search_noalloc: k must be in [1, 16] (LSB_NOALLOC_MAX_K)");
  if (nq > (size_t)LSB_NOALLOC_MAX_QUERIES)
    throw CudaError(LSB_E_CAPACITY, "lsb_knn_search_noalloc: more than 2048 queries (LSB_NOALLOC_MAX_QUERIES): the static scratch array is sized for that many");
  if (m > (size_t)INT_MAX) throw CudaError(LSB_E_CAPACITY, "lsb_knn_search_noalloc: cloud larger than 2^31 - 1 points");
  if (nq == 0) return;
  if (!d_queries_xyz || !d_idx_out || !d_d2_out || (m && !d_cloud)) throw CudaError(LSB_E_INVALID, "lsb_knn_search_noalloc: null device pointer");
  LSB_CUDA(cudaSetDevice(device));
  unsigned char* base = g_noalloc_ptr[device].load(std::memory_order_acquire);
  if (!base) {      // set-up, once per device, with that device current
    void* sym = nullptr;
    LSB_CUDA(cudaGetSymbolAddress(&sym, g_noalloc_scratch));
    if (!sym) throw CudaError(LSB_E_CUDA, "lsb_knn_search_noalloc: cudaGetSymbolAddress returned no address");
    base = static_cast<unsigned char*>(sym);
    g_noalloc_ptr[device].store(base, std::memory_order_release);
  }
  Key* partial = reinterpret_cast<Key*>(base);
  unsigned* tickets = reinterpret_cast<unsigned*>(base + NOALLOC_PARTIAL_KEYS * sizeof(Key));
  // the array keeps the previous call's contents: the tickets must start at zero
  LSB_CUDA(cudaMemsetAsync(tickets, 0, nq * sizeof(unsigned), s));
  const unsigned warps = (unsigned)(nq * NOALLOC_SEGS);
  const unsigned G = std::min<unsigned>(div_up((size_t)warps * 32, 128), 148u * 16u);
  k_knn_brute_static<<<G, 128, 0, s>>>(d_cloud, (int)m, d_queries_xyz, (int)nq, k, d_idx_out, d_d2_out, partial, tickets);
  LSB_LAUNCH_CHECK();
}

}  // namespace lsb

// GPU RollingGrid — see map.cuh.  The per-point serial hash inserts of the
// reference (RollingGrid.cxx:188-313) become: key computation -> stable radix sort
// by leaf key -> one thread per touched leaf replays the reference's sequential
// rule over its (short, arrival-ordered) segment -> binary-search join with the
// sorted map: existing leaves are updated in place, new leaves are merged by rank.
#include "map.cuh"
#include "select.cuh"
#include "sort32.cuh"
#include "../common/linalg.h"
#include <cub/cub.cuh>
#include <climits>
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <algorithm>

namespace lsb {

namespace {

constexpr int TPB = 256;
constexpr unsigned long long INVALID_KEY = ~0ull;

struct GridGeom {
  int G;                 // GridSize
  float ox, oy, oz;      // voxelGridOrigin (RollingGrid.cxx:177)
  float res;             // (float)VoxelResolution
  float leaf;            // (float)LeafSize
};

__host__ __device__ __forceinline__ int to1d(int x, int y, int z, int G) { return z * G * G + y * G + x; }   // :449-452
__host__ __device__ __forceinline__ void to3d(int id, int G, int& x, int& y, int& z)                           // :455-463
{
  z = id / (G * G);
  id -= z * G * G;
  y = id / G;
  x = id - y * G;
}
// Utils::PositionToVoxel (RollingGrid.h:38-42): float arithmetic, round half away from zero
__host__ __device__ __forceinline__ int pos2vox(float p, float o, float r) { return (int)roundf((p - o) / r); }

__device__ __forceinline__ unsigned long long make_key(int idxOut, int idxIn)
{
  return ((unsigned long long)(unsigned)idxOut << 32) | (unsigned long long)((unsigned)idxIn ^ 0x80000000u);
}

__device__ __forceinline__ int f2ord(float f) { const int i = __float_as_int(f); return (i >= 0) ? i : (i ^ 0x7fffffff); }
__device__ __forceinline__ float ord2f_dev(int i) { return __int_as_float((i >= 0) ? i : (i ^ 0x7fffffff)); }

__global__ void k_bbox_init(int* bb) { if (threadIdx.x < 3) bb[threadIdx.x] = INT_MAX; else if (threadIdx.x < 6) bb[threadIdx.x] = INT_MIN; }
__global__ void k_bbox(const lsb_point* __restrict__ pts, int n, int* __restrict__ bb)
{
  int mn[3] = {INT_MAX, INT_MAX, INT_MAX}, mx[3] = {INT_MIN, INT_MIN, INT_MIN};
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const float4 p = reinterpret_cast<const float4*>(pts)[2 * (size_t)i];
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

// leaf key of every incoming point (RollingGrid.cxx:190-202)
__global__ void k_point_keys(const lsb_point* __restrict__ pts, int n, GridGeom g, unsigned long long* __restrict__ key, uint32_t* __restrict__ idx)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = reinterpret_cast<const float4*>(pts)[2 * (size_t)i];
  const int vx = pos2vox(p.x, g.ox, g.res), vy = pos2vox(p.y, g.oy, g.res), vz = pos2vox(p.z, g.oz, g.res);
  unsigned long long k = INVALID_KEY;
  if (vx >= 0 && vx < g.G && vy >= 0 && vy < g.G && vz >= 0 && vz < g.G) {
    const float cx = (float)vx * g.res + g.ox, cy = (float)vy * g.res + g.oy, cz = (float)vz * g.res + g.oz;
    const int ix = pos2vox(p.x, cx, g.leaf), iy = pos2vox(p.y, cy, g.leaf), iz = pos2vox(p.z, cz, g.leaf);
    k = make_key(to1d(vx, vy, vz, g.G), to1d(ix, iy, iz, g.G));
  }
  key[i] = k;
  idx[i] = (uint32_t)i;
}

// Compact 32-bit form of the leaf key for the sort of one incoming cloud: the outer voxel index RELATIVE to the voxel box of
// the cloud (same z-major / y / x-minor nesting as to1d, so the order is that of the canonical key) above the inner leaf index
// biased to be non-negative.  Fits 31 bits for the usual geometries (OS-128: 14 + 17), which takes the radix sort from seven
// 8-bit passes over 64-bit keys to four over 32-bit ones; the canonical 64-bit keys are rebuilt after the sort.
struct CompactKey {
  int x0, y0, z0, dx, dy, dz;   // voxel box of the cloud
  int bits_in;                  // bits of the biased inner index
  int bias;                     // added to the (signed) inner 1-D index
};
constexpr uint32_t INVALID_KEY32 = 0xffffffffu;

__global__ void k_point_keys32(const lsb_point* __restrict__ pts, int n, GridGeom g, CompactKey ck, uint32_t* __restrict__ key, uint32_t* __restrict__ idx,
                               uint32_t* __restrict__ info)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 p = reinterpret_cast<const float4*>(pts)[2 * (size_t)i];
  const int vx = pos2vox(p.x, g.ox, g.res), vy = pos2vox(p.y, g.oy, g.res), vz = pos2vox(p.z, g.oz, g.res);
  uint32_t k = INVALID_KEY32;
  if (vx >= 0 && vx < g.G && vy >= 0 && vy < g.G && vz >= 0 && vz < g.G) {
    const float cx = (float)vx * g.res + g.ox, cy = (float)vy * g.res + g.oy, cz = (float)vz * g.res + g.oz;
    const int ix = pos2vox(p.x, cx, g.leaf), iy = pos2vox(p.y, cy, g.leaf), iz = pos2vox(p.z, cz, g.leaf);
    const int rx = vx - ck.x0, ry = vy - ck.y0, rz = vz - ck.z0;
    const long long in = (long long)to1d(ix, iy, iz, g.G) + ck.bias;
    if (rx < 0 || rx >= ck.dx || ry < 0 || ry >= ck.dy || rz < 0 || rz >= ck.dz || in < 0 || in >= (1ll << ck.bits_in)) atomicOr(info + 7, 1u);   // the host's bounds were wrong: reported, never silent
    else k = ((uint32_t)((rz * ck.dy + ry) * ck.dx + rx) << ck.bits_in) | (uint32_t)in;
  }
  key[i] = k;
  idx[i] = (uint32_t)i;
}

__global__ void k_expand_keys(const uint32_t* __restrict__ key32, int n, int G, CompactKey ck, unsigned long long* __restrict__ key64)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t c = key32[i];
  if (c == INVALID_KEY32) { key64[i] = INVALID_KEY; return; }
  const int in = (int)(c & ((1u << ck.bits_in) - 1u)) - ck.bias;
  int rel = (int)(c >> ck.bits_in);
  const int rx = rel % ck.dx; rel /= ck.dx;
  const int ry = rel % ck.dy, rz = rel / ck.dy;
  key64[i] = make_key(to1d(rx + ck.x0, ry + ck.y0, rz + ck.z0, G), in);
}

__device__ __forceinline__ int lower_bound_key(const unsigned long long* __restrict__ keys, int n, unsigned long long k)
{
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (keys[mid] < k) lo = mid + 1; else hi = mid;
  }
  return lo;
}

struct PointRaw { uint4 a, b; };
__device__ __forceinline__ PointRaw load_raw(const lsb_point* p) { PointRaw r; r.a = reinterpret_cast<const uint4*>(p)[0]; r.b = reinterpret_cast<const uint4*>(p)[1]; return r; }
__device__ __forceinline__ void store_raw(lsb_point* p, const PointRaw& r) { reinterpret_cast<uint4*>(p)[0] = r.a; reinterpret_cast<uint4*>(p)[1] = r.b; }
__device__ __forceinline__ float raw_intensity(const PointRaw& r) { return __uint_as_float(r.b.z); }
__device__ __forceinline__ unsigned raw_label(const PointRaw& r) { return r.b.w >> 24; }
__device__ __forceinline__ void raw_set_time_label(PointRaw& r, double t, unsigned label)
{
  const unsigned long long u = (unsigned long long)__double_as_longlong(t);
  r.b.x = (unsigned)(u & 0xffffffffull); r.b.y = (unsigned)(u >> 32);
  r.b.w = (r.b.w & 0x00ffffffu) | (label << 24);
}

// SamplingMode::CENTROID (:266-297) couples the leaves of one Add call: after EVERY point that reaches the end of
// the loop body (in bounds and not skipped by the fixed-point test; "effective" below) the reference re-averages
// every leaf that has received a centroid update so far in this call with that leaf's running mean.  What a leaf
// needs from the rest of the cloud is therefore only HOW MANY effective points arrive between its own ones.
// Whether a point is effective depends on its own leaf alone (the leaf existed with label 1, or `fixed` set the
// label on an earlier point of the call), so: this kernel flags the effective points (one thread per segment head,
// flags written in arrival order), an exclusive scan turns the flags into arrival positions, and k_apply_segments
// replays each leaf's recurrence over those positions.
__global__ void k_centroid_effective(const unsigned long long* __restrict__ skey, const uint32_t* __restrict__ sidx, int n,
                                     const unsigned long long* __restrict__ map_keys, const lsb_point* __restrict__ map_pts, int n_map,
                                     unsigned fixed, uint32_t* __restrict__ eff)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = skey[i];
  if (k == INVALID_KEY) return;            // out of bounds: eff stays 0
  if (i > 0 && skey[i - 1] == k) return;   // not a segment head
  const int pos = lower_bound_key(map_keys, n_map, k);
  bool exists = pos < n_map && map_keys[pos] == k;
  unsigned label = exists ? (reinterpret_cast<const uint4*>(map_pts)[2 * (size_t)pos + 1].w >> 24) : 0u;
  for (int j = i; j < n && skey[j] == k; ++j) {
    uint32_t e = 1u;
    if (!exists) exists = true;
    else if (label == 1u) e = 0u;
    if (e) label = fixed;
    eff[sidx[j]] = e;
  }
}

// `steps` re-averaging passes of RollingGrid.cxx:282-297 on one leaf: x <- (x * c + m) / (c + 1) per coordinate,
// in float.  The map is a contraction towards m: once a pass leaves the three coordinates unchanged every later
// pass does too, so the loop stops there (same values as running all the passes).
__device__ __forceinline__ void centroid_passes(PointRaw& cur, float mx, float my, float mz, uint32_t count, uint32_t steps)
{
  const float c = (float)count, c1 = (float)(count + 1u);
  float x = __uint_as_float(cur.a.x), y = __uint_as_float(cur.a.y), z = __uint_as_float(cur.a.z);
  for (uint32_t t = 0; t < steps; ++t) {
    const float nx = (x * c + mx) / c1, ny = (y * c + my) / c1, nz = (z * c + mz) / c1;
    const bool same = __float_as_uint(nx) == __float_as_uint(x) && __float_as_uint(ny) == __float_as_uint(y) && __float_as_uint(nz) == __float_as_uint(z);
    x = nx; y = ny; z = nz;
    if (same) break;
  }
  cur.a.x = __float_as_uint(x); cur.a.y = __float_as_uint(y); cur.a.z = __float_as_uint(z);
}

// One thread per segment head of the sorted incoming keys: replays the sequential
// insertion rule of RollingGrid::Add for that leaf (:203-311).
// info[0] |= updated, flags[i] = 1 for heads that create a NEW leaf.
// CENTROID only: tpos = arrival position of every point among the effective ones (exclusive scan of eff).
__global__ void k_apply_segments(const unsigned long long* __restrict__ skey, const uint32_t* __restrict__ sidx, int n,
                                 const lsb_point* __restrict__ cloud, unsigned long long* __restrict__ map_keys,
                                 lsb_point* __restrict__ map_pts, uint32_t* __restrict__ map_counts, int n_map, GridGeom g, int sampling,
                                 unsigned fixed, double currentTime, uint32_t* __restrict__ flags, lsb_point* __restrict__ seg_pts,
                                 uint32_t* __restrict__ seg_counts, uint32_t* __restrict__ info, const uint32_t* __restrict__ eff,
                                 const uint32_t* __restrict__ tpos)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  flags[i] = 0;
  const unsigned long long k = skey[i];
  if (k == INVALID_KEY) return;
  if (i > 0 && skey[i - 1] == k) return;   // not a segment head
  const int pos = lower_bound_key(map_keys, n_map, k);
  bool exists = pos < n_map && map_keys[pos] == k;
  const bool existed = exists;
  PointRaw cur;
  uint32_t count = 0;
  if (exists) { cur = load_raw(map_pts + pos); count = map_counts[pos]; }
  bool updated = false, seen = false;
  // CENTROID: running mean of the points averaged into this leaf during this call, and the arrival position
  // of the last point of this leaf that was processed
  float mx = 0.f, my = 0.f, mz = 0.f;
  uint32_t mcount = 0, last_t = 0;
  const uint32_t T = (sampling == LSB_SAMPLING_CENTROID) ? tpos[n - 1] + eff[n - 1] : 0u;
  for (int j = i; j < n && skey[j] == k; ++j) {
    const PointRaw p = load_raw(cloud + sidx[j]);
    if (sampling == LSB_SAMPLING_CENTROID && mcount && eff[sidx[j]]) {
      // the effective points of OTHER leaves that arrived since this leaf's previous one
      centroid_passes(cur, mx, my, mz, count, tpos[sidx[j]] - last_t - 1u);
    }
    if (!exists) { cur = p; exists = true; updated = true; }
    else {
      if (raw_label(cur) == 1) continue;
      if (sampling == LSB_SAMPLING_CENTROID) {
        const float c = (float)mcount, c1 = (float)(mcount + 1u);
        mx = (mx * c + __uint_as_float(p.a.x)) / c1;
        my = (my * c + __uint_as_float(p.a.y)) / c1;
        mz = (mz * c + __uint_as_float(p.a.z)) / c1;
        ++mcount;
      }
      else if (sampling == LSB_SAMPLING_LAST) { cur = p; updated = true; }
      else if (sampling == LSB_SAMPLING_MAX_INTENSITY) {
        if (raw_intensity(p) > raw_intensity(cur)) { cur = p; updated = true; }
      } else if (sampling == LSB_SAMPLING_CENTER_POINT) {
        int vx, vy, vz, ix, iy, iz;
        to3d((int)(k >> 32), g.G, vx, vy, vz);
        // inverse of to1d for the inner key is ambiguous; recompute from the stored point instead
        const float px = __uint_as_float(p.a.x), py = __uint_as_float(p.a.y), pz = __uint_as_float(p.a.z);
        const float cx = (float)vx * g.res + g.ox, cy = (float)vy * g.res + g.oy, cz = (float)vz * g.res + g.oz;
        ix = pos2vox(px, cx, g.leaf); iy = pos2vox(py, cy, g.leaf); iz = pos2vox(pz, cz, g.leaf);
        const float half = g.res / 2.f;
        const Vec3f ctr(cx - half + g.leaf * (float)ix, cy - half + g.leaf * (float)iy, cz - half + g.leaf * (float)iz);
        const Vec3f pn(px, py, pz), po(__uint_as_float(cur.a.x), __uint_as_float(cur.a.y), __uint_as_float(cur.a.z));
        if (norm(pn - ctr) < norm(po - ctr)) { cur = p; updated = true; }
      }
      // LSB_SAMPLING_FIRST: keep the stored point
    }
    if (sampling == LSB_SAMPLING_CENTROID) {
      if (mcount) centroid_passes(cur, mx, my, mz, count, 1u);   // this point's own pass, with the count before its increment
      last_t = tpos[sidx[j]];
    }
    raw_set_time_label(cur, currentTime, fixed);
    if (!seen) { ++count; seen = true; }
  }
  if (sampling == LSB_SAMPLING_CENTROID && mcount) centroid_passes(cur, mx, my, mz, count, T - last_t - 1u);
  if (existed) { store_raw(map_pts + pos, cur); map_counts[pos] = count; }
  else { flags[i] = 1; store_raw(seg_pts + i, cur); seg_counts[i] = count; }
  if (updated) atomicOr(info, 1u);
}

// merge by rank: both key lists are sorted and disjoint
// d_nb != nullptr: the number of inserts is read from device memory (the launch was sized for the worst case before the host knew it)
__global__ void k_merge(const unsigned long long* __restrict__ ka, const lsb_point* __restrict__ pa, const uint32_t* __restrict__ ca, int na,
                        const unsigned long long* __restrict__ kb, const lsb_point* __restrict__ pb, const uint32_t* __restrict__ cb, int nb,
                        unsigned long long* __restrict__ ko, lsb_point* __restrict__ po, uint32_t* __restrict__ co,
                        const uint32_t* __restrict__ d_nb = nullptr)
{
  if (d_nb) nb = (int)*d_nb;
  if (d_nb && nb == 0) return;          // nothing to insert: the map stays where it is
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < na) {
    const unsigned long long k = ka[i];
    const int o = i + lower_bound_key(kb, nb, k);
    ko[o] = k; store_raw(po + o, load_raw(pa + i)); co[o] = ca[i];
  } else if (i < na + nb) {
    const int j = i - na;
    const unsigned long long k = kb[j];
    const int o = j + lower_bound_key(ka, na, k);
    ko[o] = k; store_raw(po + o, load_raw(pb + j)); co[o] = cb[j];
  }
}

// Roll (:136-156): shift the outer index, flag survivors
__global__ void k_roll_keys(unsigned long long* __restrict__ keys, int n, int G, int ox, int oy, int oz, uint32_t* __restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const unsigned long long k = keys[i];
  int x, y, z;
  to3d((int)(k >> 32), G, x, y, z);
  x -= ox; y -= oy; z -= oz;
  const bool in = x >= 0 && x < G && y >= 0 && y < G && z >= 0 && z < G;
  flags[i] = in ? 1u : 0u;
  if (in) keys[i] = ((unsigned long long)(unsigned)to1d(x, y, z, G) << 32) | (k & 0xffffffffull);
}

__global__ void k_compact(const uint32_t* __restrict__ flags, const uint32_t* __restrict__ scan, int n,
                          const unsigned long long* __restrict__ ki, const lsb_point* __restrict__ pi, const uint32_t* __restrict__ ci,
                          unsigned long long* __restrict__ ko, lsb_point* __restrict__ po, uint32_t* __restrict__ co, uint32_t* __restrict__ info)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (flags[i]) {
    const uint32_t o = scan[i];
    if (ko) ko[o] = ki[i];
    store_raw(po + o, load_raw(pi + i));
    if (co) co[o] = ci[i];
  }
  if (i == n - 1) info[2] = scan[i] + flags[i];
}

// submap / Get flags.  mode 0: inside bbox; 1: inside && (count >= minFrames || label == 1);
// 2: inside && (count < minFrames && label != 1); 3: all; 4: count > minFrames
// bb != nullptr: the voxel bounds come from a bounding box in device memory (ordered-int encoding of k_bbox /
// k_transformed_bbox), converted exactly as the host does in SubMapSelect, so the host need not read the box back first
__global__ void k_flag_entries(const unsigned long long* __restrict__ keys, const lsb_point* __restrict__ pts, const uint32_t* __restrict__ counts,
                               int n, int G, int lx, int ly, int lz, int hx, int hy, int hz, int mode, unsigned minFrames, uint32_t* __restrict__ flags,
                               const int* __restrict__ bb = nullptr, GridGeom g = GridGeom())
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (bb) {
    lx = max(pos2vox(ord2f_dev(bb[0]), g.ox, g.res), 0); hx = min(pos2vox(ord2f_dev(bb[3]), g.ox, g.res), G - 1);
    ly = max(pos2vox(ord2f_dev(bb[1]), g.oy, g.res), 0); hy = min(pos2vox(ord2f_dev(bb[4]), g.oy, g.res), G - 1);
    lz = max(pos2vox(ord2f_dev(bb[2]), g.oz, g.res), 0); hz = min(pos2vox(ord2f_dev(bb[5]), g.oz, g.res), G - 1);
  }
  bool f = true;
  if (mode <= 2) {
    int x, y, z;
    to3d((int)(keys[i] >> 32), G, x, y, z);
    f = lx <= x && x <= hx && ly <= y && y <= hy && lz <= z && z <= hz;
  }
  if (f && mode != 0 && mode != 3) {
    const unsigned c = counts[i];
    if (mode == 4) f = c > minFrames;
    else {
      const unsigned label = reinterpret_cast<const uint4*>(pts)[2 * (size_t)i + 1].w >> 24;
      const bool keep = c >= minFrames || label == 1;
      f = (mode == 1) ? keep : !keep;
    }
  }
  flags[i] = f ? 1u : 0u;
}

// voxel / frame-count test of k_flag_entries for one entry (same modes)
__device__ __forceinline__ bool entry_selected(const unsigned long long* __restrict__ keys, const lsb_point* __restrict__ pts,
                                               const uint32_t* __restrict__ counts, int i, int G, int lx, int ly, int lz, int hx, int hy, int hz,
                                               int mode, unsigned minFrames)
{
  bool f = true;
  if (mode <= 2) {
    int x, y, z;
    to3d((int)(keys[i] >> 32), G, x, y, z);
    f = lx <= x && x <= hx && ly <= y && y <= hy && lz <= z && z <= hz;
  }
  if (f && mode != 0 && mode != 3) {
    const unsigned c = counts[i];
    if (mode == 4) f = c > minFrames;
    else {
      const unsigned label = reinterpret_cast<const uint4*>(pts)[2 * (size_t)i + 1].w >> 24;
      const bool keep = c >= minFrames || label == 1;
      f = (mode == 1) ? keep : !keep;
    }
  }
  return f;
}

// The counters a host phase waits for go to pinned host memory straight from the kernel that completes them (the tile that
// holds the grand total): no device-to-host copy is enqueued and the waiting thread polls a sequence word instead of
// synchronising the stream (map.cu: wait_info).  info[0..7] were completed by this or earlier kernels of the stream.
__device__ __forceinline__ void publish_info(const uint32_t* __restrict__ info, volatile uint32_t* __restrict__ h_info, uint32_t seq, bool last_tile)
{
  if (!last_tile || threadIdx.x != 0) return;       // thread 0 of warp 0 wrote the total (sel::tile_offset)
  __threadfence();
#pragma unroll
  for (int q = 0; q < 8; ++q) h_info[q] = __ldcg(info + q);
  __threadfence_system();
  h_info[8] = seq;
}

// Sub-map selection in ONE launch: k_flag_entries + exclusive scan + k_compact (points only) by decoupled look-back
// (select.cuh).  A tile is SEL_ITEMS consecutive entries per thread, so the output keeps the map's order.
constexpr int SEL_ITEMS = 4;
__global__ void __launch_bounds__(TPB) k_select_entries(const unsigned long long* __restrict__ keys, const lsb_point* __restrict__ pts,
                                                        const uint32_t* __restrict__ counts, int n, int G, int lx, int ly, int lz, int hx, int hy,
                                                        int hz, int mode, unsigned minFrames, const int* __restrict__ bb, GridGeom g,
                                                        lsb_point* __restrict__ out, unsigned long long* __restrict__ desc,
                                                        unsigned* __restrict__ ticket, unsigned epoch, unsigned ntiles, uint32_t* __restrict__ info,
                                                        volatile uint32_t* __restrict__ h_info, uint32_t seq)
{
  if (bb) {
    lx = max(pos2vox(ord2f_dev(bb[0]), g.ox, g.res), 0); hx = min(pos2vox(ord2f_dev(bb[3]), g.ox, g.res), G - 1);
    ly = max(pos2vox(ord2f_dev(bb[1]), g.oy, g.res), 0); hy = min(pos2vox(ord2f_dev(bb[4]), g.oy, g.res), G - 1);
    lz = max(pos2vox(ord2f_dev(bb[2]), g.oz, g.res), 0); hz = min(pos2vox(ord2f_dev(bb[5]), g.oz, g.res), G - 1);
  }
  unsigned tile = sel::take_ticket(ticket);
  const int base = (int)tile * (TPB * SEL_ITEMS) + (int)threadIdx.x * SEL_ITEMS;
  bool f[SEL_ITEMS];
  unsigned cnt = 0;
#pragma unroll
  for (int q = 0; q < SEL_ITEMS; ++q) {
    const int i = base + q;
    f[q] = i < n && entry_selected(keys, pts, counts, i, G, lx, ly, lz, hx, hy, hz, mode, minFrames);
    cnt += f[q] ? 1u : 0u;
  }
  unsigned o = sel::tile_offset<TPB>(cnt, desc, ticket, epoch, ntiles, tile, info + 2);
#pragma unroll
  for (int q = 0; q < SEL_ITEMS; ++q)
    if (f[q]) { store_raw(out + o, load_raw(pts + base + q)); ++o; }
  publish_info(info, h_info, seq, tile == ntiles - 1);
}

// k_gather_inserts with its scan, one launch: flagged segment heads -> the (sorted) insert list; info[1] = their number
__global__ void __launch_bounds__(TPB) k_select_inserts(const uint32_t* __restrict__ flags, int n, const unsigned long long* __restrict__ skey,
                                                        const lsb_point* __restrict__ seg_pts, const uint32_t* __restrict__ seg_counts,
                                                        unsigned long long* __restrict__ ins_keys, lsb_point* __restrict__ ins_pts,
                                                        uint32_t* __restrict__ ins_counts, unsigned long long* __restrict__ desc,
                                                        unsigned* __restrict__ ticket, unsigned epoch, unsigned ntiles, uint32_t* __restrict__ info,
                                                        volatile uint32_t* __restrict__ h_info, uint32_t seq)
{
  unsigned tile = sel::take_ticket(ticket);
  const int base = (int)tile * (TPB * SEL_ITEMS) + (int)threadIdx.x * SEL_ITEMS;
  bool f[SEL_ITEMS];
  unsigned cnt = 0;
#pragma unroll
  for (int q = 0; q < SEL_ITEMS; ++q) {
    f[q] = base + q < n && flags[base + q] != 0u;
    cnt += f[q] ? 1u : 0u;
  }
  unsigned o = sel::tile_offset<TPB>(cnt, desc, ticket, epoch, ntiles, tile, info + 1);
#pragma unroll
  for (int q = 0; q < SEL_ITEMS; ++q) {
    if (!f[q]) continue;
    const int i = base + q;
    ins_keys[o] = skey[i];
    store_raw(ins_pts + o, load_raw(seg_pts + i));
    ins_counts[o] = seg_counts[i];
    ++o;
  }
  publish_info(info, h_info, seq, tile == ntiles - 1);
}

// ClearOldPoints (:325-350)
__global__ void k_flag_old(const lsb_point* __restrict__ pts, int n, double currentTime, double threshold, uint32_t* __restrict__ flags)
{
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 h = reinterpret_cast<const uint4*>(pts)[2 * (size_t)i + 1];
  const double t = __longlong_as_double((long long)(((unsigned long long)h.y << 32) | h.x));
  const unsigned label = h.w >> 24;
  flags[i] = (!label && currentTime - t > threshold) ? 0u : 1u;
}

}  // namespace

// ============================================================== host side ====
static inline float ord2f_host(int i) { const int j = (i >= 0) ? i : (i ^ 0x7fffffff); float f; std::memcpy(&f, &j, 4); return f; }

void cloud_bbox_async(const lsb_point* d_cloud, size_t n, DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox, cudaStream_t s, LaunchCounter& lc)
{
  d_bbox.reserve(8, s); h_bbox.reserve(8);
  k_bbox_init<<<1, 32, 0, s>>>((int*)d_bbox.p); LSB_LAUNCH_CHECK(); lc.add();
  k_bbox<<<std::min(div_up(n, TPB), 592u), TPB, 0, s>>>(d_cloud, (int)n, (int*)d_bbox.p); LSB_LAUNCH_CHECK(); lc.add();
  LSB_CUDA(cudaMemcpyAsync(h_bbox.p, d_bbox.p, 6 * sizeof(float), cudaMemcpyDeviceToHost, s));
}

void bbox_decode(const float* h_bbox, float mn[3], float mx[3])
{
  for (int a = 0; a < 3; ++a) {
    int v;
    std::memcpy(&v, h_bbox + a, 4); mn[a] = ord2f_host(v);
    std::memcpy(&v, h_bbox + 3 + a, 4); mx[a] = ord2f_host(v);
  }
}

void cloud_bbox(const lsb_point* d_cloud, size_t n, float mn[3], float mx[3], DevBuf<float>& d_bbox, PinnedBuf<float>& h_bbox,
                cudaStream_t s, LaunchCounter& lc)
{
  cloud_bbox_async(d_cloud, n, d_bbox, h_bbox, s, lc);
  LSB_CUDA(cudaStreamSynchronize(s));
  bbox_decode(h_bbox.p, mn, mx);
}

// waits for the sequence word a kernel of stream s leaves in pinned memory (publish_info); the stream is queried now and then so
// that a failed launch surfaces as an error instead of an endless wait
static void wait_info(const uint32_t* h_info, uint32_t seq, cudaStream_t s)
{
  const volatile uint32_t* flag = h_info + 8;
  unsigned spins = 0;
  while (*flag != seq) {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
    if ((++spins & 0x3fffu) == 0) {
      const cudaError_t rc = cudaStreamQuery(s);
      if (rc == cudaErrorNotReady) continue;
      LSB_CUDA(rc);
      if (*flag != seq) throw CudaError(LSB_E_CUDA, "RollingGrid: the stream is idle but its counters never arrived");
    }
  }
}

// the box a fused box kernel of stream s publishes in pinned memory (slam_kernels.cu: sequence word behind the six bounds)
static void wait_box(const float* h_bbox, int seq, cudaStream_t s)
{
  const volatile int* flag = reinterpret_cast<const volatile int*>(h_bbox) + 6;
  unsigned spins = 0;
  while (*flag != seq) {
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
    if ((++spins & 0x3fffu) == 0) {
      const cudaError_t rc = cudaStreamQuery(s);
      if (rc == cudaErrorNotReady) continue;
      LSB_CUDA(rc);
      if (*flag != seq) throw CudaError(LSB_E_CUDA, "RollingGrid: the stream is idle but the keypoints' box never arrived");
    }
  }
}

static GridGeom make_geom(const RollingGridGpu& m)
{
  GridGeom g;
  g.G = m.GridSize;
  const float shift = (float)(int(m.GridSize / 2) * m.VoxelResolution);
  g.ox = m.Position[0] - shift; g.oy = m.Position[1] - shift; g.oz = m.Position[2] - shift;
  g.res = (float)m.VoxelResolution;
  g.leaf = (float)m.LeafSize;
  return g;
}

void RollingGridGpu::ensure_scratch(size_t n, cudaStream_t s)
{
  flags.reserve(n + 1, s); scan.reserve(n + 1, s);
  d_info.reserve(8, s);
  if (!h_info.p) { h_info.reserve(16); std::memset(h_info.p, 0, 16 * sizeof(uint32_t)); }
  size_t tmp = 0, tmp2 = 0;
  cub::DeviceScan::ExclusiveSum(nullptr, tmp, flags.p, scan.p, (int)n, s);
  cub::DeviceRadixSort::SortPairs(nullptr, tmp2, nk_a.p, nk_b.p, ni_a.p, ni_b.p, (int)n, 0, 64, s);
  cub_tmp.reserve(std::max(tmp, tmp2) + 256, s);
}

void RollingGridGpu::Clear()
{
  NbPoints = 0;
  n_entries = 0;
  ResetTree();
}

void RollingGridGpu::Reset(const float p[3], cudaStream_t)
{
  Clear();
  const float r = (float)VoxelResolution;
  for (int a = 0; a < 3; ++a) Position[a] = std::floor(p[a] / r) * r;
}

// keeps map entries with flags != 0 (order preserved); returns the new count (one sync)
size_t RollingGridGpu::compact_entries(cudaStream_t s, LaunchCounter& lc)
{
  const int n = (int)n_entries;
  if (n == 0) return 0;
  keys_b.reserve(n_entries, s); pts_b.reserve(n_entries, s); counts_b.reserve(n_entries, s);
  size_t tb = cub_tmp.cap;
  LSB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, flags.p, scan.p, n, s)); lc.add(2);
  k_compact<<<div_up(n_entries, TPB), TPB, 0, s>>>(flags.p, scan.p, n, keys.p, pts.p, counts.p, keys_b.p, pts_b.p, counts_b.p, d_info.p);
  LSB_LAUNCH_CHECK(); lc.add();
  LSB_CUDA(cudaMemcpyAsync(h_info.p, d_info.p, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  LSB_CUDA(cudaStreamSynchronize(s));
  keys.swap(keys_b); pts.swap(pts_b); counts.swap(counts_b);
  n_entries = h_info.p[2];
  return n_entries;
}

void RollingGridGpu::Roll(const float mn[3], const float mx[3], cudaStream_t s, LaunchCounter& lc)
{
  const double halfGridSize = static_cast<double>(GridSize) / 2 * VoxelResolution;
  const float h = (float)halfGridSize;
  const float r = (float)VoxelResolution;
  int vo[3];
  for (int a = 0; a < 3; ++a) {
    const float down = mn[a] - (Position[a] - h);
    const float up = mx[a] - (Position[a] + h);
    float off = (up + down) / 2.f;
    const float dmin0 = (down < 0.f) ? down : 0.f;
    const float umax0 = (up > 0.f) ? up : 0.f;
    off = (off < dmin0) ? dmin0 : off;
    off = (off > umax0) ? umax0 : off;
    vo[a] = (int)std::round(off / r);
  }
  if (vo[0] == 0 && vo[1] == 0 && vo[2] == 0) return;
  if (n_entries) {
    ensure_scratch(n_entries, s);
    k_roll_keys<<<div_up(n_entries, TPB), TPB, 0, s>>>(keys.p, (int)n_entries, GridSize, vo[0], vo[1], vo[2], flags.p); LSB_LAUNCH_CHECK(); lc.add();
    NbPoints = (unsigned)compact_entries(s, lc);
  } else NbPoints = 0;
  for (int a = 0; a < 3; ++a) Position[a] += (float)vo[a] * r;
}

// Add (:160-318) in three host phases so that the three maps of a frame can run on their own
// streams with their read-backs overlapped: AddBegin (bbox of the cloud), AddSort (roll, then
// key / sort / replay / gather), AddMerge (rank merge of the new leaves).
void RollingGridGpu::AddBegin(const lsb_point* d_cloud, size_t n, bool roll, cudaStream_t s, LaunchCounter& lc)
{
  box_by_kernel = false;
  if (n == 0) return;
  if (roll) cloud_bbox_async(d_cloud, n, d_bbox, h_bbox, s, lc);
}

void RollingGridGpu::AddSort(const lsb_point* d_cloud, size_t n, bool fixed, double currentTime, bool roll, cudaStream_t s, LaunchCounter& lc)
{
  if (n == 0) return;
  float mn[3], mx[3];
  if (roll) {
    if (box_by_kernel) wait_box(h_bbox.p, box_seq, s);     // published by the tail of the ICP loop (k_finish_keypoints)
    else LSB_CUDA(cudaStreamSynchronize(s));
    box_by_kernel = false;
    bbox_decode(h_bbox.p, mn, mx);
    Roll(mn, mx, s, lc);
  }
  const GridGeom g = make_geom(*this);
  const int N = (int)n;
  // compact 32-bit sort keys when the cloud's voxel box and the inner index range fit (see CompactKey)
  CompactKey ck{};
  bool compact = false;
  int key_bits = 32;
  static const bool compact_on = [] { const char* e = std::getenv("LSB_MAP_COMPACT_KEYS"); return !e || e[0] != '0'; }();   // A/B hook
  if (roll && compact_on && LeafSize > 0. && VoxelResolution > 0.) {
    const float o[3] = {g.ox, g.oy, g.oz};
    int lo[3], hi[3];
    bool any = true;
    for (int a = 0; a < 3; ++a) {
      lo[a] = std::max(pos2vox(mn[a], o[a], g.res), 0);
      hi[a] = std::min(pos2vox(mx[a], o[a], g.res), g.G - 1);
      any = any && lo[a] <= hi[a];
    }
    if (any) {
      ck.x0 = lo[0]; ck.y0 = lo[1]; ck.z0 = lo[2];
      ck.dx = hi[0] - lo[0] + 1; ck.dy = hi[1] - lo[1] + 1; ck.dz = hi[2] - lo[2] + 1;
      const long long imax = (long long)std::ceil(VoxelResolution / (2. * LeafSize)) + 2;       // |inner coordinate| along one axis
      const long long B = imax * ((long long)g.G * g.G + g.G + 1);
      int bits_in = 1, bits_out = 0;
      while ((1ll << bits_in) <= 2 * B) ++bits_in;
      while ((1ll << bits_out) < (long long)ck.dx * ck.dy * ck.dz) ++bits_out;
      if (bits_in + bits_out <= 31 && B < (1ll << 30)) { ck.bits_in = bits_in; ck.bias = (int)B; compact = true; key_bits = bits_in + bits_out; }
    }
  }
  nk_a.reserve(n, s); nk_b.reserve(n, s); ni_a.reserve(n, s); ni_b.reserve(n, s);
  ins_keys.reserve(n, s); ins_pts.reserve(n, s); ins_counts.reserve(n, s);
  pts_b.reserve(n, s); counts_b.reserve(n, s);      // per-segment scratch (seg_pts / seg_counts)
  ensure_scratch(std::max(n, n_entries), s);
  const unsigned G = div_up(n, TPB);
  LSB_CUDA(cudaMemsetAsync(d_info.p, 0, 8 * sizeof(uint32_t), s));
  size_t tb = cub_tmp.cap;
  const uint32_t* sorted_idx = ni_b.p;
  if (compact) {
    // flags / scan are free until the segment replay: they hold the 32-bit keys before and after the sort
    k_point_keys32<<<G, TPB, 0, s>>>(d_cloud, N, g, ck, flags.p, ni_a.p, d_info.p); LSB_LAUNCH_CHECK(); lc.add();
    static const bool own_sort = [] { const char* e = std::getenv("LSB_MAP_OWN_SORT"); return !e || e[0] != '0'; }();   // A/B hook
    if (own_sort && n <= Sort32::max_pairs()) {
      // one launch (sort32.cu); one bit above the significant ones keeps INVALID_KEY32 (all ones) behind every valid key
      const bool in_b = sorter.sort(flags.p, ni_a.p, scan.p, ni_b.p, n, std::min(32, key_bits + 1), s, lc);
      sorted_idx = in_b ? ni_b.p : ni_a.p;
      k_expand_keys<<<G, TPB, 0, s>>>(in_b ? scan.p : flags.p, N, g.G, ck, nk_b.p); LSB_LAUNCH_CHECK(); lc.add();
    } else {
      LSB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tb, flags.p, scan.p, ni_a.p, ni_b.p, N, 0, 32, s)); lc.add(2 + 4);
      k_expand_keys<<<G, TPB, 0, s>>>(scan.p, N, g.G, ck, nk_b.p); LSB_LAUNCH_CHECK(); lc.add();
    }
  } else {
    k_point_keys<<<G, TPB, 0, s>>>(d_cloud, N, g, nk_a.p, ni_a.p); LSB_LAUNCH_CHECK(); lc.add();
    // significant bits only: the outer index is < G^3 and INVALID_KEY (all ones) still sorts last as long as 2^bits > G^3
    int outer_bits = 1;
    while ((1ll << outer_bits) <= (long long)g.G * g.G * g.G) ++outer_bits;
    const int end_bit = std::min(64, 32 + outer_bits);
    LSB_CUDA(cub::DeviceRadixSort::SortPairs(cub_tmp.p, tb, nk_a.p, nk_b.p, ni_a.p, ni_b.p, N, 0, end_bit, s)); lc.add(2 + (end_bit + 7) / 8);
  }
  if (Sampling == LSB_SAMPLING_CENTROID) {
    c_eff.reserve(n, s); c_tpos.reserve(n, s);
    LSB_CUDA(cudaMemsetAsync(c_eff.p, 0, n * sizeof(uint32_t), s));
    k_centroid_effective<<<G, TPB, 0, s>>>(nk_b.p, sorted_idx, N, keys.p, pts.p, (int)n_entries, fixed ? 1u : 0u, c_eff.p); LSB_LAUNCH_CHECK(); lc.add();
    tb = cub_tmp.cap;
    LSB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, c_eff.p, c_tpos.p, N, s)); lc.add(2);
  }
  k_apply_segments<<<G, TPB, 0, s>>>(nk_b.p, sorted_idx, N, d_cloud, keys.p, pts.p, counts.p, (int)n_entries, g, Sampling, fixed ? 1u : 0u,
                                     currentTime, flags.p, pts_b.p, counts_b.p, d_info.p, c_eff.p, c_tpos.p);
  LSB_LAUNCH_CHECK(); lc.add();
  {
    const unsigned tiles = div_up(n, (size_t)TPB * SEL_ITEMS);
    const unsigned epoch = sel_state.prepare(tiles, s);
    info_seq = info_seq + 1 ? info_seq + 1 : 1;
    k_select_inserts<<<tiles, TPB, 0, s>>>(flags.p, N, nk_b.p, pts_b.p, counts_b.p, ins_keys.p, ins_pts.p, ins_counts.p, sel_state.desc.p,
                                           sel_state.ticket.p, epoch, tiles, d_info.p, h_info.p, info_seq);
    LSB_LAUNCH_CHECK(); lc.add();
  }
  // The merge by rank is enqueued at once, sized for the worst case (every point a new leaf), and reads the number of inserts
  // from device memory: it runs while the count is on its way to the host (AddMerge), which only does the bookkeeping.
  // Output: a second set of map arrays (merge_*: keys_b / pts_b / counts_b are the roll's and the replay's scratch).
  {
    const size_t total_max = n_entries + n;
    merge_keys.reserve(total_max, s); merge_pts.reserve(total_max, s); merge_counts.reserve(total_max, s);
    k_merge<<<div_up(total_max, TPB), TPB, 0, s>>>(keys.p, pts.p, counts.p, (int)n_entries, ins_keys.p, ins_pts.p, ins_counts.p, 0,
                                                   merge_keys.p, merge_pts.p, merge_counts.p, d_info.p + 1);
    LSB_LAUNCH_CHECK(); lc.add();
  }
}

void RollingGridGpu::AddMerge(size_t n, cudaStream_t s, LaunchCounter& lc)
{
  if (n == 0) return;
  wait_info(h_info.p, info_seq, s);
  if (h_info.p[7]) throw CudaError(LSB_E_CUDA, "RollingGrid::Add: a leaf key fell outside the compact key range (internal bound error)");
  const bool updated = h_info.p[0] != 0;
  const size_t n_ins = h_info.p[1];
  if (n_ins) {          // the merged map was written by the launch AddSort enqueued: it becomes the map
    keys.swap(merge_keys); pts.swap(merge_pts); counts.swap(merge_counts);
    n_entries += n_ins;
    NbPoints += (unsigned)n_ins;
  }
  if (updated) ResetTree();
}

void RollingGridGpu::Add(const lsb_point* d_cloud, size_t n, bool fixed, double currentTime, bool roll, cudaStream_t s, LaunchCounter& lc)
{
  AddBegin(d_cloud, n, roll, s, lc);
  AddSort(d_cloud, n, fixed, currentTime, roll, s, lc);
  AddMerge(n, s, lc);
}

void RollingGridGpu::ClearOldPoints(double currentTime, cudaStream_t s, LaunchCounter& lc)
{
  if (!n_entries) return;
  ensure_scratch(n_entries, s);
  k_flag_old<<<div_up(n_entries, TPB), TPB, 0, s>>>(pts.p, (int)n_entries, currentTime, DecayingThreshold, flags.p); LSB_LAUNCH_CHECK(); lc.add();
  compact_entries(s, lc);   // NbPoints is NOT decremented (reference behaviour, :325-350)
}

size_t RollingGridGpu::Get(bool clean, lsb_point* d_out, cudaStream_t s, LaunchCounter& lc)
{
  if (!n_entries) return 0;
  const int n = (int)n_entries;
  if (!clean) {
    LSB_CUDA(cudaMemcpyAsync(d_out, pts.p, n_entries * sizeof(lsb_point), cudaMemcpyDeviceToDevice, s));
    return n_entries;
  }
  ensure_scratch(n_entries, s);
  k_flag_entries<<<div_up(n_entries, TPB), TPB, 0, s>>>(keys.p, pts.p, counts.p, n, GridSize, 0, 0, 0, 0, 0, 0, 4, MinFramesPerVoxel, flags.p);
  LSB_LAUNCH_CHECK(); lc.add();
  size_t tb = cub_tmp.cap;
  LSB_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp.p, tb, flags.p, scan.p, n, s)); lc.add(2);
  k_compact<<<div_up(n_entries, TPB), TPB, 0, s>>>(flags.p, scan.p, n, nullptr, pts.p, nullptr, nullptr, d_out, nullptr, d_info.p);
  LSB_LAUNCH_CHECK(); lc.add();
  LSB_CUDA(cudaMemcpyAsync(h_info.p, d_info.p, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
  LSB_CUDA(cudaStreamSynchronize(s));
  return h_info.p[2];
}

void RollingGridGpu::build_grid(const float* lo, const float* hi, cudaStream_t s, LaunchCounter& lc)
{
  tree_valid = true;
  if (n_submap == 0) { grid.clear(); return; }
  grid.build(submap.p, n_submap, knn_cell_factor * (float)LeafSize, lo, hi, s, lc);
}

void RollingGridGpu::BuildSubMapKdTree(cudaStream_t s, LaunchCounter& lc)
{
  submap.reserve(std::max<size_t>(n_entries, 1), s);
  n_submap = Get(false, submap.p, s, lc);
  build_grid(nullptr, nullptr, s, lc);
}

// BuildSubMapKdTree(mn, mx, minNbPoints) (:362-442) in two host phases (SubMapSelect enqueues the
// first selection pass and its count read-back, SubMapFinish consumes it and builds the kNN grid)
// so that the three maps overlap their read-backs on their own streams.
void RollingGridGpu::submap_pass(int mode, size_t offset, cudaStream_t s, LaunchCounter& lc, const int* d_bb)
{
  const int n = (int)n_entries;
  const unsigned tiles = div_up(n_entries, (size_t)TPB * SEL_ITEMS);
  const unsigned epoch = sel_state.prepare(tiles, s);
  info_seq = info_seq + 1 ? info_seq + 1 : 1;
  k_select_entries<<<tiles, TPB, 0, s>>>(keys.p, pts.p, counts.p, n, GridSize, sub_lo[0], sub_lo[1], sub_lo[2], sub_hi[0], sub_hi[1], sub_hi[2],
                                         mode, MinFramesPerVoxel, d_bb, make_geom(*this), submap.p + offset, sel_state.desc.p,
                                         sel_state.ticket.p, epoch, tiles, d_info.p, h_info.p, info_seq);
  LSB_LAUNCH_CHECK(); lc.add();
}

void RollingGridGpu::SubMapSelect(const float mn[3], const float mx[3], int minNbPoints, cudaStream_t s, LaunchCounter& lc)
{
  const GridGeom g = make_geom(*this);
  const float o[3] = {g.ox, g.oy, g.oz};
  for (int a = 0; a < 3; ++a) {
    sub_lo[a] = std::max(pos2vox(mn[a], o[a], g.res), 0);
    sub_hi[a] = std::min(pos2vox(mx[a], o[a], g.res), GridSize - 1);
  }
  n_submap = 0;
  sub_min_nb = minNbPoints;
  sub_filter = !(minNbPoints < 0 || MinFramesPerVoxel <= 1);
  sub_from_device = false;
  submap.reserve(std::max<size_t>(n_entries, 1), s);
  if (n_entries) {
    ensure_scratch(n_entries, s);
    submap_pass(sub_filter ? 1 : 0, 0, s, lc);
  }
}

// The same with the bounding box still in device memory (d_bbox, its copy to h_bbox already enqueued by the caller): the
// selection runs without the host knowing the bounds; SubMapFinish reads box and count back with ONE synchronisation.
void RollingGridGpu::SubMapSelectDevice(int minNbPoints, cudaStream_t s, LaunchCounter& lc)
{
  n_submap = 0;
  sub_min_nb = minNbPoints;
  sub_filter = !(minNbPoints < 0 || MinFramesPerVoxel <= 1);
  sub_from_device = true;
  submap.reserve(std::max<size_t>(n_entries, 1), s);
  if (n_entries) {
    ensure_scratch(n_entries, s);
    submap_pass(sub_filter ? 1 : 0, 0, s, lc, reinterpret_cast<const int*>(d_bbox.p));
  }
}

void RollingGridGpu::SubMapFinish(cudaStream_t s, LaunchCounter& lc)
{
  const GridGeom g = make_geom(*this);
  const float o[3] = {g.ox, g.oy, g.oz};
  auto grid_bounds = [&](float blo[3], float bhi[3]) {
    // every submap point lies in an outer voxel of [lo, hi]: host-known kNN grid bounds (with slack)
    for (int a = 0; a < 3; ++a) {
      blo[a] = o[a] + ((float)sub_lo[a] - 0.5f) * g.res - 0.01f * g.res;
      bhi[a] = o[a] + ((float)sub_hi[a] + 0.5f) * g.res + 0.01f * g.res;
    }
    return sub_lo[0] <= sub_hi[0] && sub_lo[1] <= sub_hi[1] && sub_lo[2] <= sub_hi[2];
  };
  auto bounds_from_box = [&] {
    float mn[3], mx[3];
    bbox_decode(h_bbox.p, mn, mx);
    for (int a = 0; a < 3; ++a) {
      sub_lo[a] = std::max(pos2vox(mn[a], o[a], g.res), 0);
      sub_hi[a] = std::min(pos2vox(mx[a], o[a], g.res), GridSize - 1);
    }
  };
  float blo[3], bhi[3];
  if (sub_from_device && n_entries && !sub_filter && box_seq > 0) {
    // The box arrives first (the head kernel of the chain published it): the kNN grid over the sub-map is enqueued at once,
    // sized for the worst case (the whole map selected) with the sub-map's size read from device memory, and runs while the
    // size is on its way to the host.
    wait_box(h_bbox.p, box_seq, s);
    bounds_from_box();
    if (grid_bounds(blo, bhi)) {
      tree_valid = true;
      grid.build(submap.p, n_entries, knn_cell_factor * (float)LeafSize, blo, bhi, s, lc, d_info.p + 2);
      wait_info(h_info.p, info_seq, s);
      n_submap = h_info.p[2];
      grid.set_count(n_submap);
      return;
    }
  }
  // the selection kernel leaves the count in pinned memory (publish_info); the box was published by an earlier kernel of the stream
  if (n_entries) wait_info(h_info.p, info_seq, s);
  else if (sub_from_device) LSB_CUDA(cudaStreamSynchronize(s));
  if (sub_from_device) bounds_from_box();
  if (n_entries) {
    n_submap = h_info.p[2];
    if (sub_filter && (int)n_submap < sub_min_nb) {
      submap_pass(2, n_submap, s, lc);
      wait_info(h_info.p, info_seq, s);
      n_submap += h_info.p[2];
    }
  }
  const bool bounded = grid_bounds(blo, bhi);
  build_grid(bounded ? blo : nullptr, bounded ? bhi : nullptr, s, lc);
}

void RollingGridGpu::BuildSubMapKdTree(const float mn[3], const float mx[3], int minNbPoints, cudaStream_t s, LaunchCounter& lc)
{
  SubMapSelect(mn, mx, minNbPoints, s, lc);
  SubMapFinish(s, lc);
}

void RollingGridGpu::SetGridSize(int size, cudaStream_t s, LaunchCounter& lc)
{
  GridSize = size;
  const size_t n = n_entries;
  if (n) {
    DevBuf<lsb_point> prev;
    prev.reserve(n, s);
    Get(false, prev.p, s, lc);
    Clear();
    Add(prev.p, n, false, -1., true, s, lc);
    LSB_CUDA(cudaStreamSynchronize(s));
  } else Clear();
}

void RollingGridGpu::SetVoxelResolution(double resolution, cudaStream_t s, LaunchCounter& lc)
{
  VoxelResolution = int(resolution / LeafSize) * LeafSize;
  const float r = (float)VoxelResolution;
  for (int a = 0; a < 3; ++a) Position[a] = std::floor(Position[a] / r) * r;
  const size_t n = n_entries;
  if (n) {
    DevBuf<lsb_point> prev;
    prev.reserve(n, s);
    Get(false, prev.p, s, lc);
    Clear();
    Add(prev.p, n, false, -1., true, s, lc);
    LSB_CUDA(cudaStreamSynchronize(s));
  } else Clear();
}

}  // namespace lsb

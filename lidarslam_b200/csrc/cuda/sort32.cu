// Stable radix sort of (32-bit key, 32-bit value) pairs in ONE cooperative launch, for the leaf keys of a keypoint cloud in
// the map update (map.cu: a few 10^3 .. 10^5 pairs, 20 - 31 significant key bits).  cub::DeviceRadixSort needs ~12 stream
// operations for such a sort (histogram, scan, a memset and a onesweep kernel per 8-bit digit): ~95 us of launch and tail
// latencies for 81 k pairs on the frame's critical chain.  Here: at most three digits of up to 11 bits, all passes inside one
// kernel with grid barriers in between.
//   per pass, phase A: every CTA ranks the 4096 pairs of its tile digit by digit in arrival order (per-warp counters in shared
//     memory, lanes of a round grouped with __match_any_sync) and publishes its digit histogram;
//   grid barrier; phase B: digit base = pairs with a smaller digit anywhere + pairs with the same digit in earlier tiles;
//     a pair goes to base + pairs with the same digit in earlier warps of the tile + its rank in its warp; grid barrier.
// Stable by construction: tiles, warps, rounds and lanes are all taken in input order.
#include "sort32.cuh"
#include <algorithm>

namespace lsb {

namespace {

constexpr int S_TPB = 256, S_ITEMS = 16, S_TILE = S_TPB * S_ITEMS, S_WARPS = S_TPB / 32;
constexpr int S_MAX_BITS = 11, S_MAX_BINS = 1 << S_MAX_BITS;

__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned target)
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

struct SortPlan { int passes; int shift[3]; int bits[3]; };

__global__ void __launch_bounds__(S_TPB) k_sort_pairs32(uint32_t* __restrict__ ka, uint32_t* __restrict__ va, uint32_t* __restrict__ kb,
                                                        uint32_t* __restrict__ vb, int n, SortPlan plan, uint32_t* __restrict__ hist,
                                                        unsigned* __restrict__ bar)
{
  extern __shared__ uint32_t smem[];
  uint32_t* wc = smem;                               // [S_WARPS][bins] per-warp digit counters, then exclusive over the warps
  uint32_t* binbase = smem + S_WARPS * S_MAX_BINS;   // [bins]
  __shared__ uint32_t s_scan[S_WARPS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ncta = (int)gridDim.x, cta = (int)blockIdx.x;
  const int tile0 = cta * S_TILE + warp * (S_ITEMS * 32) + lane;
  unsigned nbar = 0;
  uint32_t* kin = ka; uint32_t* vin = va; uint32_t* kout = kb; uint32_t* vout = vb;
  for (int pass = 0; pass < plan.passes; ++pass) {
    const int bins = 1 << plan.bits[pass], shift = plan.shift[pass];
    const uint32_t mask = (uint32_t)bins - 1u;
    uint32_t* h = hist + (size_t)(pass & 1) * (size_t)ncta * S_MAX_BINS;
    for (int b = threadIdx.x; b < S_WARPS * bins; b += S_TPB) wc[(b / bins) * S_MAX_BINS + (b % bins)] = 0u;
    __syncthreads();
    uint32_t key[S_ITEMS], val[S_ITEMS], rank[S_ITEMS];
    uint32_t* mywc = wc + warp * S_MAX_BINS;
#pragma unroll
    for (int j = 0; j < S_ITEMS; ++j) {
      const int i = tile0 + j * 32;
      const bool valid = i < n;
      key[j] = valid ? kin[i] : 0xffffffffu;
      val[j] = valid ? vin[i] : 0u;
    }
#pragma unroll
    for (int j = 0; j < S_ITEMS; ++j) {
      const bool valid = tile0 + j * 32 < n;
      const uint32_t d = (key[j] >> shift) & mask;
      // lanes of this round with the same digit (invalid lanes form their own group and are not counted)
      const unsigned grp = __match_any_sync(0xffffffffu, valid ? d : 0xffffffffu);
      const uint32_t before = valid ? mywc[d] : 0u;
      rank[j] = before + (uint32_t)__popc(grp & ((1u << lane) - 1u));
      __syncwarp();
      if (valid && lane == __ffs(grp) - 1) mywc[d] = before + (uint32_t)__popc(grp);
      __syncwarp();
    }
    __syncthreads();
    // tile histogram; the per-warp counters become exclusive prefixes over the warps
    for (int b = threadIdx.x; b < bins; b += S_TPB) {
      uint32_t run = 0;
#pragma unroll
      for (int w = 0; w < S_WARPS; ++w) { const uint32_t c = wc[w * S_MAX_BINS + b]; wc[w * S_MAX_BINS + b] = run; run += c; }
      h[(size_t)cta * S_MAX_BINS + b] = run;
    }
    grid_barrier(bar, (++nbar) * (unsigned)ncta);
    // digit bases: thread t owns the consecutive bins [t * per, (t + 1) * per)
    const int per = (bins + S_TPB - 1) / S_TPB;
    uint32_t tot[S_MAX_BINS / S_TPB], pre[S_MAX_BINS / S_TPB];
    uint32_t mine = 0;
#pragma unroll
    for (int u = 0; u < S_MAX_BINS / S_TPB; ++u) {
      tot[u] = pre[u] = 0u;
      const int b = (int)threadIdx.x * per + u;
      if (u < per && b < bins) {
        for (int c0 = 0; c0 < ncta; c0 += 8) {        // eight independent loads in flight per batch
          uint32_t v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) v[e] = (c0 + e < ncta) ? __ldcg(h + (size_t)(c0 + e) * S_MAX_BINS + b) : 0u;
#pragma unroll
          for (int e = 0; e < 8; ++e) { tot[u] += v[e]; pre[u] += (c0 + e < cta) ? v[e] : 0u; }
        }
      }
      mine += tot[u];
    }
    // exclusive scan of `mine` over the block
    uint32_t inc = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) { const uint32_t o = __shfl_up_sync(0xffffffffu, inc, off); if (lane >= off) inc += o; }
    if (lane == 31) s_scan[warp] = inc;
    __syncthreads();
    uint32_t base = inc - mine;
    for (int w = 0; w < warp; ++w) base += s_scan[w];
#pragma unroll
    for (int u = 0; u < S_MAX_BINS / S_TPB; ++u) {
      const int b = (int)threadIdx.x * per + u;
      if (u < per && b < bins) { binbase[b] = base + pre[u]; base += tot[u]; }
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < S_ITEMS; ++j) {
      if (tile0 + j * 32 < n) {
        const uint32_t d = (key[j] >> shift) & mask;
        const uint32_t dst = binbase[d] + mywc[d] + rank[j];
        kout[dst] = key[j];
        vout[dst] = val[j];
      }
    }
    if (pass + 1 < plan.passes) grid_barrier(bar, (++nbar) * (unsigned)ncta);
    uint32_t* t = kin; kin = kout; kout = t;
    t = vin; vin = vout; vout = t;
  }
  // the last CTA out leaves the barrier counters zero for the next launch
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(bar + 1, 1u) == (unsigned)ncta - 1u) { bar[0] = 0u; bar[1] = 0u; __threadfence(); }
  }
}

}  // namespace

size_t Sort32::max_pairs() { return (size_t)64 * S_TILE; }

bool Sort32::sort(uint32_t* ka, uint32_t* va, uint32_t* kb, uint32_t* vb, size_t n, int key_bits, cudaStream_t s, LaunchCounter& lc)
{
  // returns true when the sorted pairs are in (kb, vb), false when they are in (ka, va)
  if (n == 0) return false;
  if (n > max_pairs() || key_bits < 1 || key_bits > 32) throw CudaError(LSB_E_INVALID, "Sort32: unsupported size");
  SortPlan plan;
  plan.passes = (key_bits + S_MAX_BITS - 1) / S_MAX_BITS;
  int left = key_bits, at = 0;
  for (int p = 0; p < 3; ++p) { plan.shift[p] = 0; plan.bits[p] = 1; }
  for (int p = 0; p < plan.passes; ++p) {
    const int b = (left + (plan.passes - p) - 1) / (plan.passes - p);
    plan.shift[p] = at; plan.bits[p] = b;
    at += b; left -= b;
  }
  const unsigned ncta = div_up(n, (size_t)S_TILE);
  constexpr size_t kSmem = (size_t)(S_WARPS + 1) * S_MAX_BINS * sizeof(uint32_t);
  if (!configured) {
    LSB_CUDA(cudaFuncSetAttribute(k_sort_pairs32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem));
    configured = true;
  }
  hist.reserve((size_t)2 * 64 * S_MAX_BINS, s);
  if (!bar.p) { bar.reserve(2, s); LSB_CUDA(cudaMemsetAsync(bar.p, 0, 2 * sizeof(unsigned), s)); }
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeCooperative;
  attr.val.cooperative = 1;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ncta);
  cfg.blockDim = dim3(S_TPB);
  cfg.dynamicSmemBytes = kSmem;
  cfg.stream = s;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  LSB_CUDA(cudaLaunchKernelEx(&cfg, k_sort_pairs32, ka, va, kb, vb, (int)n, plan, hist.p, bar.p));
  lc.add();
  return (plan.passes & 1) != 0;
}

}  // namespace lsb

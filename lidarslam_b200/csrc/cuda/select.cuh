// Single-pass device-wide stream compaction support: the exclusive offset of a tile's selected items by decoupled look-back
// (one launch instead of flag kernel + scan init + scan + scatter on the latency-bound map chains).  Tiles take a dynamic ticket,
// publish (epoch, status, value) descriptors in one 64-bit word and look back a warp's width at a time.  The epoch changes
// with every launch, so descriptors of earlier launches read as "not there yet" and nothing has to be cleared between
// launches; the ticket counter is reset by the last tile.  One SelectState serves launches that are ordered on a stream.
#pragma once
#include "dev_util.h"

namespace lsb {

struct SelectState {
  DevBuf<unsigned long long> desc;
  DevBuf<unsigned> ticket;        // [1]
  unsigned epoch = 0;
  // call before a launch of `tiles` tiles; returns the epoch to pass to the kernel
  unsigned prepare(size_t tiles, cudaStream_t s)
  {
    if (tiles + 1 > desc.cap) {
      desc.reserve(tiles + 1, s);
      LSB_CUDA(cudaMemsetAsync(desc.p, 0, desc.cap * sizeof(unsigned long long), s));
    }
    if (!ticket.p) { ticket.reserve(4, s); LSB_CUDA(cudaMemsetAsync(ticket.p, 0, 4 * sizeof(unsigned), s)); }
    epoch = (epoch + 1) & 0x3fffffffu;
    if (epoch == 0) {          // wrapped: stale descriptors could match again
      LSB_CUDA(cudaMemsetAsync(desc.p, 0, desc.cap * sizeof(unsigned long long), s));
      epoch = 1;
    }
    return epoch;
  }
};

#ifdef __CUDACC__
namespace sel {

constexpr unsigned ST_AGG = 1u, ST_PREFIX = 2u;

__device__ __forceinline__ unsigned long long pack(unsigned epoch, unsigned status, unsigned value)
{
  return ((unsigned long long)((epoch << 2) | status) << 32) | value;
}
__device__ __forceinline__ unsigned long long ld_desc(const unsigned long long* p)
{
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_desc(unsigned long long* p, unsigned long long v)
{
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Called by every thread of the block (blockDim.x = TPB, a multiple of 32, at most 1024).  `count` = this thread's number of
// selected items.  Returns the tile id in `tile` (all threads) and the exclusive offset of this THREAD's items in the output.
// `total_out`: the last tile stores the grand total there (may be nullptr).  Tiles cover the input in ticket order: the caller
// derives its items from `tile`, not from blockIdx.x.
template <int TPB>
__device__ __forceinline__ unsigned tile_offset(unsigned count, unsigned long long* desc, unsigned* ticket, unsigned epoch, unsigned ntiles,
                                                unsigned& tile, unsigned* total_out)
{
  __shared__ unsigned s_warp[TPB / 32];
  __shared__ unsigned s_excl;
  const unsigned lane = threadIdx.x & 31u, wid = threadIdx.x >> 5;
  // inclusive scan inside the warp
  unsigned inc = count;
#pragma unroll
  for (int off = 1; off < 32; off <<= 1) {
    const unsigned o = __shfl_up_sync(0xffffffffu, inc, off);
    if ((int)lane >= off) inc += o;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  unsigned wbase = 0, agg = 0;
#pragma unroll
  for (int w = 0; w < TPB / 32; ++w) {
    const unsigned v = s_warp[w];
    if (w < (int)wid) wbase += v;
    agg += v;
  }
  if (wid == 0) {
    const unsigned t = tile;          // valid in every thread (set by the caller from take_ticket)
    unsigned excl = 0;
    if (t == 0) {
      if (lane == 0) st_desc(desc, pack(epoch, ST_PREFIX, agg));
    } else {
      if (lane == 0) st_desc(desc + t, pack(epoch, ST_AGG, agg));
      int j = (int)t - 1 - (int)lane;            // this lane's predecessor in the current window
      while (true) {
        unsigned long long d = 0;
        bool ready = true;
        if (j >= 0) {
          d = ld_desc(desc + j);
          const unsigned hi = (unsigned)(d >> 32);
          ready = (hi >> 2) == epoch && (hi & 3u) != 0u;
        }
        if (!__all_sync(0xffffffffu, ready)) continue;             // spin until the whole window is published
        const unsigned hi = (unsigned)(d >> 32);
        const bool is_prefix = j >= 0 && (hi & 3u) == ST_PREFIX;
        const unsigned pm = __ballot_sync(0xffffffffu, is_prefix || j < 0);   // j < 0: before tile 0 (contributes nothing, ends the walk)
        const unsigned val = (j >= 0) ? (unsigned)d : 0u;
        // lanes up to and including the nearest prefix (smallest lane index with the bit set) contribute
        const int stop = pm ? (__ffs(pm) - 1) : 32;
        unsigned c = ((int)lane <= stop) ? val : 0u;
#pragma unroll
        for (int off = 16; off; off >>= 1) c += __shfl_xor_sync(0xffffffffu, c, off);
        excl += c;
        if (pm) break;
        j -= 32;
      }
      if (lane == 0) st_desc(desc + t, pack(epoch, ST_PREFIX, excl + agg));
    }
    if (lane == 0) {
      s_excl = excl;
      if (t == ntiles - 1) {
        if (total_out) *total_out = excl + agg;
        *ticket = 0u;                 // every tile has taken its ticket by now
      }
    }
  }
  __syncthreads();
  return s_excl + wbase + (inc - count);
}

// Dynamic tile id of the block (call once, before tile_offset): tiles are numbered in the order blocks start, which makes the
// look-back deadlock-free whatever order the hardware schedules blocks in.
__device__ __forceinline__ unsigned take_ticket(unsigned* ticket)
{
  __shared__ unsigned s_t;
  if (threadIdx.x == 0) s_t = atomicAdd(ticket, 1u);
  __syncthreads();
  return s_t;
}

}  // namespace sel
#endif

}  // namespace lsb

// Stable single-launch radix sort of (32-bit key, 32-bit value) pairs (sort32.cu).
#pragma once
#include "dev_util.h"

namespace lsb {

struct Sort32 {
  DevBuf<uint32_t> hist;       // [2][64 tiles][2048 bins] digit histograms of the tiles, double-buffered by pass parity
  DevBuf<unsigned> bar;        // [arrivals, exits] of the grid barrier: zero between launches
  bool configured = false;
  static size_t max_pairs();   // larger inputs: use cub::DeviceRadixSort
  // Sorts n pairs by the low `key_bits` bits of the key, stable.  (ka, va) is the input and is overwritten; returns true when
  // the result is in (kb, vb), false when it is in (ka, va).  Launches on s are ordered; one Sort32 per stream.
  bool sort(uint32_t* ka, uint32_t* va, uint32_t* kb, uint32_t* vb, size_t n, int key_bits, cudaStream_t s, LaunchCounter& lc);
};

}  // namespace lsb

// LCP overlap estimator arithmetic shared by the CUDA kernels and the CPU oracle
// (reference slam_lib/src/ConfidenceEstimators.cxx:27-65, Slam.cxx:1370-1388).
//   probability of a sampled point against one map = exp(-d2 / (2 (leaf/3)^2)), d2 = squared
//   distance (float) to its nearest sub-map point; best over the maps in use; mean over the samples.
// Canonical sum order (the reference's OpenMP reduction order is unspecified): the radix-32 tree
// of lm_core.h in FLOAT - fold32 of 32 consecutive samples, fold32 of 32 group sums (1024-sample
// chunk), fold32 of 32 chunk sums, block sums added in order; missing leaves are +0.0f.
#pragma once
#include "detmath.h"

namespace lsb {

#define LSB_LCP_CHUNK 1024

// index of the n-th sampled point: cloud->at(n / subsamplingRatio) (float division, truncation)
LSB_HD size_t lcp_sample_index(int n, float ratio) { return (size_t)((float)n / ratio); }

// (leaf / 3.f)^2 as the reference computes it: double leaf / float 3 -> double, squared, stored in a float
LSB_HD float lcp_sq_threshold(double leaf)
{
  const double t = leaf / (double)3.f;
  return (float)(t * t);
}

// beyond this squared distance the probability is exactly 0.f (exp(-110) < half the smallest subnormal)
LSB_HD float lcp_sq_cutoff(float sqThreshold) { return 2.f * sqThreshold * 110.f; }

LSB_HD float lcp_probability(float nnSqDist, float sqThreshold)
{
  if (!(nnSqDist <= lcp_sq_cutoff(sqThreshold))) return 0.f;
  return dm::expf(-nnSqDist / (2.f * sqThreshold));
}

}  // namespace lsb

// CPU ORACLE (test infrastructure) — exact kNN (stand-in for nanoflann behind
// KDTreePCLAdaptor.h:57-105) and KeypointsMatcher restatement
// (reference slam_lib/src/KeypointsMatcher.cxx, line ranges per function).
#pragma once
#include "orc_common.h"
#include <array>

namespace orc {

// Exact k-d tree, leaf size 16, L2 on float[3] with d2 = ((dx^2 + dy^2) + dz^2)
// in float (nanoflann L2_Simple_Adaptor order).  nanoflann's own tie order follows
// its traversal; the canonical rule here (and on the GPU) is ascending (d2, index).
struct KdTree {
  const Cloud* cloud = nullptr;
  struct Node { int lo, hi, left, right, dim; float split; };
  std::vector<Node> nodes;
  std::vector<int> perm;

  void Reset(const Cloud* c);
  bool Empty() const { return !cloud || cloud->empty(); }
  // returns number found (min(k, size)); idx/d2 ascending by (d2, idx)
  int Knn(const float q[3], int k, int* idx, float* d2) const;
  int Knn(const double q[3], int k, int* idx, float* d2) const
  {
    const float f[3] = {(float)q[0], (float)q[1], (float)q[2]};   // KDTreePCLAdaptor.h:96-101
    return Knn(f, k, idx, d2);
  }

 private:
  int Build(int lo, int hi);
  void Search(int node, const float q[3], int k, int& n, int* idx, float* d2) const;
};

// One match = residual record r = A (R X + t - P), Tukey(sat) scaled by weight
struct MatchRecord {
  double A[9];   // row-major
  double P[3];
  double X[3];
  double weight;
};

struct MatchingResults {
  std::vector<uint8_t> Rejections;
  std::vector<double> Weights;
  std::vector<MatchRecord> Residuals;   // index-aligned; valid where Rejections == SUCCESS
  std::array<int, LSB_MATCH_N_STATUS> Histogram{};
  unsigned NbMatches() const { return (unsigned)Histogram[LSB_MATCH_SUCCESS]; }
  void Reset(size_t n)
  {
    Rejections.assign(n, LSB_MATCH_UNKNOWN);
    Weights.assign(n, 0.);
    Residuals.assign(n, MatchRecord{});
    Histogram.fill(0);
  }
};

struct Matcher {
  lsb_match_params Params;
  int NbThreads = 1;
  Iso3d PosePrior;
  Matcher(const lsb_match_params& p, const Iso3d& prior, int nbThreads) : Params(p), NbThreads(nbThreads), PosePrior(prior) {}

  MatchingResults BuildMatchResiduals(const Cloud& curr, const KdTree& prev, int type) const;   // :33-74

  struct MatchInfo { uint8_t Status; double Weight; MatchRecord Rec; };
  MatchInfo BuildLineMatch(const KdTree& t, const Point& p) const;    // :106-187
  MatchInfo BuildPlaneMatch(const KdTree& t, const Point& p) const;   // :190-273
  MatchInfo BuildBlobMatch(const KdTree& t, const Point& p) const;    // :276-346
  void GetPerRingLineNeighbors(const KdTree& t, const double pos[3], unsigned k, std::vector<int>& idx, std::vector<float>& d2) const;                // :349-405
  void GetRansacLineNeighbors(const KdTree& t, const double pos[3], unsigned k, double maxDistInlier, std::vector<int>& idx, std::vector<float>& d2) const;   // :408-480
};

}  // namespace orc

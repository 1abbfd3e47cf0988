// CPU ORACLE (test infrastructure) — kNN + KeypointsMatcher restatement.
// Reference: slam_lib/src/KeypointsMatcher.cxx, slam_lib/include/LidarSlam/KDTreePCLAdaptor.h
#include "orc_matcher.h"
#include <algorithm>
#include <numeric>

namespace orc {

//==============================================================================
// KdTree
//==============================================================================
void KdTree::Reset(const Cloud* c)
{
  cloud = c;
  nodes.clear();
  perm.clear();
  if (!c || c->empty()) return;
  perm.resize(c->size());
  std::iota(perm.begin(), perm.end(), 0);
  nodes.reserve(c->size() / 8 + 2);
  Build(0, (int)c->size());
}

int KdTree::Build(int lo, int hi)
{
  const int id = (int)nodes.size();
  nodes.push_back(Node{lo, hi, -1, -1, 0, 0.f});
  if (hi - lo <= 16) return id;
  float mn[3] = {1e30f, 1e30f, 1e30f}, mx[3] = {-1e30f, -1e30f, -1e30f};
  for (int i = lo; i < hi; ++i) {
    const Point& p = (*cloud)[perm[i]];
    const float v[3] = {p.x, p.y, p.z};
    for (int d = 0; d < 3; ++d) { mn[d] = std::min(mn[d], v[d]); mx[d] = std::max(mx[d], v[d]); }
  }
  int dim = 0;
  for (int d = 1; d < 3; ++d) if (mx[d] - mn[d] > mx[dim] - mn[dim]) dim = d;
  const int mid = (lo + hi) / 2;
  auto coord = [&](int i) { const Point& p = (*cloud)[i]; return dim == 0 ? p.x : (dim == 1 ? p.y : p.z); };
  std::nth_element(perm.begin() + lo, perm.begin() + mid, perm.begin() + hi, [&](int a, int b) { return coord(a) < coord(b); });
  const float split = coord(perm[mid]);
  const int l = Build(lo, mid);
  const int r = Build(mid, hi);
  nodes[id].left = l; nodes[id].right = r; nodes[id].dim = dim; nodes[id].split = split;
  return id;
}

void KdTree::Search(int node, const float q[3], int k, int& n, int* idx, float* d2) const
{
  const Node& nd = nodes[node];
  if (nd.left < 0) {
    for (int i = nd.lo; i < nd.hi; ++i) {
      const int pi = perm[i];
      const Point& p = (*cloud)[pi];
      const float dx = q[0] - p.x, dy = q[1] - p.y, dz = q[2] - p.z;
      const float d = (dx * dx + dy * dy) + dz * dz;
      if (n == k && !(d < d2[k - 1] || (d == d2[k - 1] && pi < idx[k - 1]))) continue;
      int j = (n < k) ? n : k - 1;
      while (j > 0 && (d < d2[j - 1] || (d == d2[j - 1] && pi < idx[j - 1]))) { d2[j] = d2[j - 1]; idx[j] = idx[j - 1]; --j; }
      d2[j] = d; idx[j] = pi;
      if (n < k) ++n;
    }
    return;
  }
  // all points of the left child have coord <= split, of the right child >= split
  const float diff = q[nd.dim] - nd.split;
  const int first = (diff < 0.f) ? nd.left : nd.right;
  const int second = (diff < 0.f) ? nd.right : nd.left;
  Search(first, q, k, n, idx, d2);
  // diff*diff is a lower bound (in float) of any distance in the far child; keep ties
  if (n < k || !(diff * diff > d2[k - 1])) Search(second, q, k, n, idx, d2);
}

int KdTree::Knn(const float q[3], int k, int* idx, float* d2) const
{
  if (Empty() || k <= 0) return 0;
  int n = 0;
  Search(0, q, k, n, idx, d2);
  return n;
}

//==============================================================================
// Matcher
//==============================================================================
MatchingResults Matcher::BuildMatchResiduals(const Cloud& curr, const KdTree& prev, int type) const
{
  MatchingResults res;
  res.Reset(curr.size());
  if (!curr.empty() && !prev.Empty()) {
    std::array<int, LSB_MATCH_N_STATUS> hist{};
#pragma omp parallel num_threads(NbThreads)
    {
      std::array<int, LSB_MATCH_N_STATUS> h{};
#pragma omp for schedule(guided, 8)
      for (int i = 0; i < (int)curr.size(); ++i) {
        MatchInfo m;
        switch (type) {
          case LSB_EDGE: m = BuildLineMatch(prev, curr[i]); break;
          case LSB_PLANE: m = BuildPlaneMatch(prev, curr[i]); break;
          case LSB_BLOB: m = BuildBlobMatch(prev, curr[i]); break;
          default: m = MatchInfo{LSB_MATCH_UNKNOWN, 0., MatchRecord{}};
        }
        res.Rejections[i] = m.Status;
        res.Weights[i] = m.Weight;
        res.Residuals[i] = m.Rec;
        h[m.Status]++;
      }
#pragma omp critical
      for (int s = 0; s < LSB_MATCH_N_STATUS; ++s) hist[s] += h[s];
    }
    res.Histogram = hist;
  }
  return res;
}

static inline Matcher::MatchInfo Fail(uint8_t status) { return Matcher::MatchInfo{status, 0., MatchRecord{}}; }

static inline MatchRecord MakeRecord(const Mat3d& A, const Vec3d& mean, const Vec3d& X, double w)
{
  MatchRecord r;
  for (int i = 0; i < 9; ++i) r.A[i] = A.m[i];
  r.P[0] = mean.x; r.P[1] = mean.y; r.P[2] = mean.z;
  r.X[0] = X.x; r.X[1] = X.y; r.X[2] = X.z;
  r.weight = w;
  return r;
}

//-----------------------------------------------------------------------------
Matcher::MatchInfo Matcher::BuildLineMatch(const KdTree& t, const Point& p) const
{
  if (Params.edge_nb_neighbors < 2 || Params.edge_min_nb_neighbors < 2) return Fail(LSB_MATCH_BAD_MODEL_PARAMETRIZATION);
  const Vec3d base((double)p.x, (double)p.y, (double)p.z);
  const Vec3d world = lsb::apply(PosePrior, base);
  const double pos[3] = {world.x, world.y, world.z};
  std::vector<int> idx;
  std::vector<float> d2;
  if (Params.single_edge_per_ring) GetPerRingLineNeighbors(t, pos, Params.edge_nb_neighbors, idx, d2);
  else GetRansacLineNeighbors(t, pos, Params.edge_nb_neighbors, Params.edge_max_model_error, idx, d2);
  if (idx.size() < Params.edge_min_nb_neighbors) return Fail(LSB_MATCH_NOT_ENOUGH_NEIGHBORS);
  if ((double)d2.back() > Params.max_neighbors_distance * Params.max_neighbors_distance) return Fail(LSB_MATCH_NEIGHBORS_TOO_FAR);
  Vec3d mean; Mat3d V; double ev[3];
  MeanAndPCA<double>(*t.cloud, idx.data(), (int)idx.size(), mean, V, ev);
  const Vec3d n = V.col(2);
  Mat3d A;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) A(i, j) = ((i == j) ? 1.0 : 0.0) - n[i] * n[j];
  if (!std::isfinite(A(0, 0))) return Fail(LSB_MATCH_INVALID_NUMERICAL);
  const double mse = ev[0] + ev[1];
  if (mse >= Params.edge_max_model_error * Params.edge_max_model_error) return Fail(LSB_MATCH_MSE_TOO_LARGE);
  const double w = (mse <= 1e-6) ? 1. : 1. - std::sqrt(mse) / Params.edge_max_model_error;
  return MatchInfo{LSB_MATCH_SUCCESS, w, MakeRecord(A, mean, base, w)};
}

//-----------------------------------------------------------------------------
Matcher::MatchInfo Matcher::BuildPlaneMatch(const KdTree& t, const Point& p) const
{
  if (Params.plane_nb_neighbors < 3) return Fail(LSB_MATCH_BAD_MODEL_PARAMETRIZATION);
  const Vec3d base((double)p.x, (double)p.y, (double)p.z);
  const Vec3d world = lsb::apply(PosePrior, base);
  const double pos[3] = {world.x, world.y, world.z};
  const int k = (int)Params.plane_nb_neighbors;
  std::vector<int> idx(k);
  std::vector<float> d2(k);
  const int n = t.Knn(pos, k, idx.data(), d2.data());
  if (n < k) return Fail(LSB_MATCH_NOT_ENOUGH_NEIGHBORS);
  if ((double)d2[n - 1] > Params.max_neighbors_distance * Params.max_neighbors_distance) return Fail(LSB_MATCH_NEIGHBORS_TOO_FAR);
  Vec3d mean; Mat3d V; double ev[3];
  MeanAndPCA<double>(*t.cloud, idx.data(), n, mean, V, ev);
  if (ev[1] / ev[2] < Params.planarity_threshold) return Fail(LSB_MATCH_BAD_PCA_STRUCTURE);
  const Vec3d nn = V.col(0);
  Mat3d A;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) A(i, j) = nn[i] * nn[j];
  if (!std::isfinite(A(0, 0))) return Fail(LSB_MATCH_INVALID_NUMERICAL);
  const double mse = ev[0];
  if (mse >= Params.plane_max_model_error * Params.plane_max_model_error) return Fail(LSB_MATCH_MSE_TOO_LARGE);
  const double w = (mse <= 1e-6) ? 1. : 1. - std::sqrt(mse) / Params.plane_max_model_error;
  return MatchInfo{LSB_MATCH_SUCCESS, w, MakeRecord(A, mean, base, w)};
}

//-----------------------------------------------------------------------------
Matcher::MatchInfo Matcher::BuildBlobMatch(const KdTree& t, const Point& p) const
{
  if (Params.blob_nb_neighbors < 4) return Fail(LSB_MATCH_BAD_MODEL_PARAMETRIZATION);
  const Vec3d base((double)p.x, (double)p.y, (double)p.z);
  const Vec3d world = lsb::apply(PosePrior, base);
  const double pos[3] = {world.x, world.y, world.z};
  const int k = (int)Params.blob_nb_neighbors;
  std::vector<int> idx(k);
  std::vector<float> d2(k);
  const int n = t.Knn(pos, k, idx.data(), d2.data());
  if (n < k) return Fail(LSB_MATCH_NOT_ENOUGH_NEIGHBORS);
  if ((double)d2[n - 1] > Params.max_neighbors_distance * Params.max_neighbors_distance) return Fail(LSB_MATCH_NEIGHBORS_TOO_FAR);
  Vec3d mean; Mat3d V; double ev[3];
  MeanAndPCA<double>(*t.cloud, idx.data(), n, mean, V, ev);
  if (ev[0] <= 0. || ev[1] <= 0.) return Fail(LSB_MATCH_BAD_PCA_STRUCTURE);
  const double s[3] = {1.0 / std::sqrt(ev[0]), 1.0 / std::sqrt(ev[1]), 1.0 / std::sqrt(ev[2])};
  Mat3d A;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j)
      A(i, j) = lsb::sum3((V(i, 0) * s[0]) * V(j, 0), (V(i, 1) * s[1]) * V(j, 1), (V(i, 2) * s[2]) * V(j, 2));
  if (!std::isfinite(A(0, 0)) || !std::isfinite(s[0] * (s[1] * s[2]))) return Fail(LSB_MATCH_INVALID_NUMERICAL);
  return MatchInfo{LSB_MATCH_SUCCESS, 1.0, MakeRecord(A, mean, base, 1.0)};
}

//-----------------------------------------------------------------------------
void Matcher::GetPerRingLineNeighbors(const KdTree& t, const double pos[3], unsigned knearest, std::vector<int>& outIdx, std::vector<float>& outD2) const
{
  std::vector<int> idx(knearest);
  std::vector<float> d2(knearest);
  const int n = t.Knn(pos, (int)knearest, idx.data(), d2.data());
  if (n == 0) return;
  const Cloud& c = *t.cloud;
  const int closestLaserId = (int)c[idx[0]].laser_id;
  int idMin = std::numeric_limits<int>::max(), idMax = std::numeric_limits<int>::min();
  for (int k = 0; k < n; ++k) { const int s = c[idx[k]].laser_id; idMin = std::min(idMin, s); idMax = std::max(idMax, s); }
  std::vector<uint8_t> took(idMax - idMin + 1, 0);
  took[closestLaserId - idMin] = 1;
  const int maxScanLineDiff = 4;
  for (int id = idMin; id <= idMax; ++id)
    if (std::abs(closestLaserId - id) > maxScanLineDiff) took[id - idMin] = 1;
  outIdx.clear(); outD2.clear();
  for (int k = 0; k < n; ++k) {
    const int s = c[idx[k]].laser_id - idMin;
    if (!took[s]) { took[s] = 1; outIdx.push_back(idx[k]); outD2.push_back(d2[k]); }
  }
}

//-----------------------------------------------------------------------------
void Matcher::GetRansacLineNeighbors(const KdTree& t, const double pos[3], unsigned knearest, double maxDistInlier, std::vector<int>& outIdx, std::vector<float>& outD2) const
{
  std::vector<int> idx(knearest);
  std::vector<float> d2(knearest);
  const int n = t.Knn(pos, (int)knearest, idx.data(), d2.data());
  if (n < 2) return;
  const Cloud& c = *t.cloud;
  const float sqMax = (float)(maxDistInlier * maxDistInlier);
  const Vec3f P1 = P3f(c[idx[0]]);
  int best = -1; size_t bestCount = 0;
  std::vector<std::vector<int>> inliers(n - 1);
  for (int pt = 1; pt < n; ++pt) {
    const Vec3f dir = lsb::normalized(P3f(c[idx[pt]]) - P1);
    std::vector<int>& in = inliers[pt - 1];
    for (int cd = 1; cd < n; ++cd) {
      if (cd == pt) in.push_back(cd);
      else if (lsb::sqnorm(lsb::cross(P3f(c[idx[cd]]) - P1, dir)) < sqMax) in.push_back(cd);
    }
  }
  for (int k = 0; k < n - 1; ++k)
    if (inliers[k].size() > bestCount) { bestCount = inliers[k].size(); best = k; }
  outIdx.clear(); outD2.clear();
  outIdx.push_back(idx[0]); outD2.push_back(d2[0]);
  for (int in : inliers[best]) { outIdx.push_back(idx[in]); outD2.push_back(d2[in]); }
}

}  // namespace orc

// CPU ORACLE (test infrastructure) — SpinningSensorKeypointExtractor restatement.
// Reference: slam_lib/src/SpinningSensorKeypointExtractor.cxx (line ranges per function).
#include "orc_extractor.h"

namespace orc {

namespace {

// LineFitting (SSKE.cxx:33-115), all arithmetic in float
struct LineFitting {
  Vec3f Direction, Position;
  float MaxDistance = 0.02f;                       // :52
  float MaxAngle = (float)(40. * 0.017453293);     // :55  DEG2RAD(40.) (PCL macro)

  float SquaredDistanceToPoint(const Vec3f& p) const   // :111-114
  {
    return lsb::sqnorm(lsb::cross(p - Position, Direction));
  }

  bool FitPCA(const Cloud& cloud, const int* idx, int n)   // :59-84
  {
    Mat3f evecs;
    float evals[3];
    MeanAndPCA<float>(cloud, idx, n, Position, evecs, evals);
    Direction = evecs.col(2);
    const float sqMaxDistance = MaxDistance * MaxDistance;
    for (int i = 0; i < n; ++i)
      if (SquaredDistanceToPoint(P3f(cloud[idx[i]])) > sqMaxDistance) return false;
    return true;
  }

  bool FitPCAAndCheckConsistency(const Cloud& cloud, const int* idx, int n)   // :87-108
  {
    const float maxSinAngle = lsb::dm::sinf(MaxAngle);
    const Vec3f U = lsb::normalized(P3f(cloud[idx[n - 1]]) - P3f(cloud[idx[0]]));
    for (int i = 0; i < n - 1; ++i) {
      const Vec3f V = lsb::normalized(P3f(cloud[idx[i + 1]]) - P3f(cloud[idx[i]]));
      const float sinAngle = lsb::norm(lsb::cross(U, V));
      if (sinAngle > maxSinAngle) return false;
    }
    return FitPCA(cloud, idx, n);
  }
};

// Utils::SortIdx(v, false) (Utilities.h:96-110).  std::sort is unstable; the
// oracle's canonical order is (value descending, index ascending).
std::vector<int> SortIdxDesc(const std::vector<float>& v)
{
  std::vector<int> idx(v.size());
  std::iota(idx.begin(), idx.end(), 0);
  std::sort(idx.begin(), idx.end(), [&v](int a, int b) { return v[a] > v[b] || (v[a] == v[b] && a < b); });
  return idx;
}

}  // namespace

Extractor::Extractor() { lsb_default_ke_params(&P); }

//-----------------------------------------------------------------------------
void Extractor::ComputeKeyPoints(const Point* pts, size_t n)
{
  AzimuthalResolution = P.azimuthal_resolution;
  ConvertAndSortScanLines(pts, n);
  PrepareDataForNextFrame();
  InvalidateNotUsablePoints();
  ComputeCurvature();
  SetKeyPointsLabels();
  P.azimuthal_resolution = AzimuthalResolution;
}

//-----------------------------------------------------------------------------
void Extractor::ConvertAndSortScanLines(const Point* pts, size_t n)
{
  ScanSize = n;
  for (auto& c : ScanLines) c.clear();
  for (auto& c : ScanLineSrc) c.clear();
  for (size_t i = 0; i < n; ++i) {
    const Point& p = pts[i];
    while (p.laser_id >= ScanLines.size()) { ScanLines.emplace_back(); ScanLineSrc.emplace_back(); }
    ScanLines[p.laser_id].push_back(p);
    ScanLineSrc[p.laser_id].push_back((int)i);
  }
  NbLaserRings = (unsigned)ScanLines.size();   // never shrinks (:156-164)
  if (AzimuthalResolution < 1e-6 || M_PI / 4. < AzimuthalResolution) EstimateAzimuthalResolution();
}

//-----------------------------------------------------------------------------
void Extractor::PrepareDataForNextFrame()
{
  for (int k = 0; k < 3; ++k) Keypoints[k].clear();
  Angles.resize(NbLaserRings); Saliency.resize(NbLaserRings); DepthGap.resize(NbLaserRings);
  IntensityGap.resize(NbLaserRings); IsPointValid.resize(NbLaserRings); Label.resize(NbLaserRings);
  for (unsigned s = 0; s < NbLaserRings; ++s) {
    const size_t np = ScanLines[s].size();
    IsPointValid[s].assign(np, 7);
    Label[s].assign(np, 0);
    Angles[s].assign(np, 0.f); Saliency[s].assign(np, 0.f); DepthGap[s].assign(np, 0.f); IntensityGap[s].assign(np, 0.f);
  }
}

//-----------------------------------------------------------------------------
void Extractor::InvalidateNotUsablePoints()
{
  const int W = P.neighbor_width;
  const float angleBeamNormal = (float)((double)(90 - P.min_beam_surface_angle) / 180. * M_PI);   // :210
  float azimuthalResolution = AzimuthalResolution;
  if (azimuthalResolution < 1e-6 || M_PI / 4 < azimuthalResolution) azimuthalResolution = (float)(0.2 / 180. * M_PI);  // :215-219
  const float maxPosDiffCoeff = lsb::dm::sinf(azimuthalResolution) / lsb::dm::cosf(azimuthalResolution + angleBeamNormal);  // :222

#pragma omp parallel for num_threads(NbThreads) schedule(guided)
  for (int s = 0; s < (int)NbLaserRings; ++s) {
    const Cloud& c = ScanLines[s];
    const int Npts = (int)c.size();
    std::vector<uint8_t>& valid = IsPointValid[s];
    if (IsScanLineAlmostEmpty(Npts)) {
      for (int i = 0; i < Npts; ++i) valid[i] = 0;
      continue;
    }
    for (int i = 0; i < W; ++i) { valid[i] = 0; valid[Npts - 1 - i] = 0; }
    for (int index = W; index < Npts - W; ++index) {
      const Vec3f cur = P3f(c[index]);
      const float L = lsb::norm(cur);
      if (L < P.min_distance_to_sensor) valid[index] = 0;
      const float a = L * maxPosDiffCoeff;
      const float maxPosDiff = (a < 0.02f) ? 0.02f : a;     // std::max(a, 0.02f)
      const float sqMaxPosDiff = maxPosDiff * maxPosDiff;
      const Vec3f nxt = P3f(c[index + 1]);
      if (lsb::sqnorm(nxt - cur) > sqMaxPosDiff) {
        if (L < lsb::norm(nxt)) {
          valid[index + 1] = 0;
          for (int i = index + 1; i < index + W; ++i) {
            if (lsb::sqnorm(P3f(c[i + 1]) - P3f(c[i])) > sqMaxPosDiff) break;
            valid[i + 1] = 0;
          }
        } else {
          valid[index] = 0;
          for (int i = index - 1; i > index - W; --i) {
            if (lsb::sqnorm(P3f(c[i + 1]) - P3f(c[i])) > sqMaxPosDiff) break;
            valid[i] = 0;
          }
        }
      }
    }
  }
}

//-----------------------------------------------------------------------------
void Extractor::ComputeCurvature()
{
  const int W = P.neighbor_width;
  const float sqDistToLineThreshold = P.dist_to_line_threshold * P.dist_to_line_threshold;
  const float sqDepthDistCoeff = 0.25f;
  const float minDepthGapDist = 1.5f;
  const float FMAX = std::numeric_limits<float>::max();

#pragma omp parallel for num_threads(NbThreads) schedule(guided)
  for (int s = 0; s < (int)NbLaserRings; ++s) {
    const Cloud& c = ScanLines[s];
    const int Npts = (int)c.size();
    if (IsScanLineAlmostEmpty(Npts)) continue;
    std::vector<int> left(W), right(W), far;
    far.reserve(2 * W);
    for (int index = W; index + W < Npts; ++index) {
      if (IsPointValid[s][index] == 0) continue;
      const Vec3f central = P3f(c[index]);
      const float ig = c[index + 1].intensity - c[index - 1].intensity;
      IntensityGap[s][index] = (ig < 0.f) ? -ig : ig;
      for (int j = index - 1; j >= index - W; --j) left[index - 1 - j] = j;
      for (int j = index + 1; j <= index + W; ++j) right[j - index - 1] = j;
      LineFitting leftLine, rightLine;
      const bool leftFlat = leftLine.FitPCAAndCheckConsistency(c, left.data(), W);
      const bool rightFlat = rightLine.FitPCAAndCheckConsistency(c, right.data(), W);
      float distLeft = 0.f, distRight = 0.f;
      if (leftFlat && rightFlat) {
        distLeft = leftLine.SquaredDistanceToPoint(central);
        distRight = rightLine.SquaredDistanceToPoint(central);
        if (distLeft < sqDistToLineThreshold && distRight < sqDistToLineThreshold)
          Angles[s][index] = lsb::norm(lsb::cross(leftLine.Direction, rightLine.Direction));
      } else if (!leftFlat && rightFlat) {
        distLeft = FMAX;
        for (int id : left) { const float d = rightLine.SquaredDistanceToPoint(P3f(c[id])); distLeft = (d < distLeft) ? d : distLeft; }
        distLeft *= sqDepthDistCoeff;
      } else if (leftFlat && !rightFlat) {
        distRight = FMAX;
        for (int id : right) { const float d = leftLine.SquaredDistanceToPoint(P3f(c[id])); distRight = (d < distRight) ? d : distRight; }
        distRight *= sqDepthDistCoeff;
      } else {
        const float sqCurrDepth = lsb::sqnorm(central);
        bool gapL = false, gapR = false;
        far.clear();
        for (int id : left) {
          const float d = lsb::sqnorm(P3f(c[id])) - sqCurrDepth;
          if (((d < 0.f) ? -d : d) > minDepthGapDist) { gapL = true; far.push_back(id); }
          else if (gapL) break;
        }
        for (int id : right) {
          const float d = lsb::sqnorm(P3f(c[id])) - sqCurrDepth;
          if (((d < 0.f) ? -d : d) > minDepthGapDist) { gapR = true; far.push_back(id); }
          else if (gapR) break;
        }
        if (far.size() > (unsigned)W) {
          LineFitting farLine;
          farLine.FitPCA(c, far.data(), (int)far.size());
          Saliency[s][index] = farLine.SquaredDistanceToPoint(central);
        }
      }
      DepthGap[s][index] = (distLeft < distRight) ? distRight : distLeft;   // std::max(distLeft, distRight)
    }
  }
}

//-----------------------------------------------------------------------------
void Extractor::SetKeyPointsLabels()
{
  const int W = P.neighbor_width;
  const float sqEdgeSaliencythreshold = P.edge_saliency_threshold * P.edge_saliency_threshold;
  const float sqEdgeDepthGapThreshold = P.edge_depth_gap_threshold * P.edge_depth_gap_threshold;

#pragma omp parallel for num_threads(NbThreads) schedule(guided)
  for (int s = 0; s < (int)NbLaserRings; ++s) {
    const int Npts = (int)ScanLines[s].size();
    if (IsScanLineAlmostEmpty(Npts)) continue;
    std::vector<uint8_t>& valid = IsPointValid[s];
    std::vector<uint8_t>& label = Label[s];
    const std::vector<int> sDepth = SortIdxDesc(DepthGap[s]);
    const std::vector<int> sAngles = SortIdxDesc(Angles[s]);
    const std::vector<int> sSal = SortIdxDesc(Saliency[s]);
    const std::vector<int> sInt = SortIdxDesc(IntensityGap[s]);

    auto addEdges = [&](const std::vector<int>& sorted, const std::vector<float>& values, float threshold, int half) {   // :499-524
      for (int index : sorted) {
        if (values[index] < threshold) break;
        if (!(valid[index] & 1)) continue;
        label[index] |= 1;
        const int b = std::max(0, index - half), e = std::min(Npts - 1, index + half);
        for (int j = b; j <= e; ++j) valid[j] &= ~1;
      }
    };
    addEdges(sDepth, DepthGap[s], sqEdgeDepthGapThreshold, W - 1);
    addEdges(sAngles, Angles[s], P.edge_sin_angle_threshold, W);
    addEdges(sSal, Saliency[s], sqEdgeSaliencythreshold, W - 1);
    addEdges(sInt, IntensityGap[s], P.edge_intensity_gap_threshold, 1);

    for (int k = Npts - 1; k >= 0; --k) {   // :536-563
      const int index = sAngles[k];
      const float sinAngle = Angles[s][index];
      if (sinAngle > P.plane_sin_angle_threshold) break;
      if (!(valid[index] & 2) || (double)sinAngle < 1e-6) continue;
      label[index] |= 2;
      const int b = std::max(0, index - 4), e = std::min(Npts - 1, index + 4);
      for (int j = b; j <= e; ++j) valid[j] &= ~2;
    }
    for (int index = 0; index < Npts; index += 3)   // :568-572
      if (valid[index] & 4) label[index] |= 4;
  }

  for (unsigned s = 0; s < NbLaserRings; ++s) {   // :575-589
    const Cloud& c = ScanLines[s];
    for (size_t i = 0; i < c.size(); ++i)
      for (int k = 0; k < 3; ++k)
        if (Label[s][i] & (1 << k)) {
          IsPointValid[s][i] |= (1 << k);
          Keypoints[k].push_back(c[i]);
        }
  }
}

//-----------------------------------------------------------------------------
void Extractor::EstimateAzimuthalResolution()
{
  std::vector<float> angles;
  angles.reserve(ScanSize);
  for (const Cloud& c : ScanLines) {
    for (size_t i = 1; i < c.size(); ++i) {
      const float x1 = c[i - 1].x, y1 = c[i - 1].y, x2 = c[i].x, y2 = c[i].y;
      const float d = x1 * x2 + y1 * y2;
      const float n1 = std::sqrt(x1 * x1 + y1 * y1), n2 = std::sqrt(x2 * x2 + y2 * y2);
      float angle = lsb::dm::acosf(d / (n1 * n2));
      if (angle < 0.f) angle = -angle;
      if (angle > 1e-4) angles.push_back(angle);
    }
  }
  if (angles.size() < 100) return;   // :616-620
  std::sort(angles.begin(), angles.end());
  size_t maxInliersIdx = angles.size();
  float maxAngle = (float)(5. / 180. * M_PI);
  float medianAngle = 0.f;
  while ((double)maxAngle > 1.8 * (double)medianAngle) {
    maxInliersIdx = std::upper_bound(angles.begin(), angles.begin() + maxInliersIdx, maxAngle) - angles.begin();
    medianAngle = angles[maxInliersIdx / 2];
    maxAngle = (float)std::min((double)medianAngle * 2., (double)maxAngle / 1.8);
  }
  AzimuthalResolution = medianAngle;
}

//-----------------------------------------------------------------------------
std::vector<float> Extractor::DebugArray(int field) const
{
  std::vector<float> v(ScanSize, 0.f);
  for (unsigned s = 0; s < NbLaserRings; ++s)
    for (size_t i = 0; i < ScanLineSrc[s].size(); ++i) {
      float val;
      switch (field) {
        case 0: val = Angles[s][i]; break;
        case 1: val = Saliency[s][i]; break;
        case 2: val = DepthGap[s][i]; break;
        case 3: val = IntensityGap[s][i]; break;
        case 4: case 5: case 6: val = (Label[s][i] >> (field - 4)) & 1; break;
        default: val = (IsPointValid[s][i] >> (field - 7)) & 1; break;
      }
      v[ScanLineSrc[s][i]] = val;
    }
  return v;
}

std::vector<uint8_t> Extractor::LabelsInputOrder() const
{
  std::vector<uint8_t> v(ScanSize, 0);
  for (unsigned s = 0; s < NbLaserRings; ++s)
    for (size_t i = 0; i < ScanLineSrc[s].size(); ++i) v[ScanLineSrc[s][i]] = Label[s][i];
  return v;
}

}  // namespace orc

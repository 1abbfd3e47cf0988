// CPU ORACLE (test infrastructure) — SpinningSensorKeypointExtractor restatement.
// Follows reference slam_lib/src/SpinningSensorKeypointExtractor.cxx line ranges
// cited at each function.  See orc_common.h for the oracle's status.
#pragma once
#include "orc_common.h"
#include <algorithm>
#include <numeric>

namespace orc {

struct Extractor {
  lsb_ke_params P;
  int NbThreads = 1;
  float AzimuthalResolution = 0.f;       // SSKE.h:157
  unsigned NbLaserRings = 0;

  // per ring (scan line) storage
  std::vector<Cloud> ScanLines;
  std::vector<std::vector<int>> ScanLineSrc;      // index of each ring point in the input scan
  std::vector<std::vector<float>> Angles, DepthGap, Saliency, IntensityGap;
  std::vector<std::vector<uint8_t>> IsPointValid, Label;  // bit k = Keypoint k
  Cloud Keypoints[3];
  size_t ScanSize = 0;

  Extractor();
  void ComputeKeyPoints(const Point* pts, size_t n);                 // SSKE.cxx:118-136
  // GetDebugArray (SSKE.cxx:640-680): field as in lsb_get_ke_debug
  std::vector<float> DebugArray(int field) const;
  // per input point: bit0 EDGE bit1 PLANE bit2 BLOB
  std::vector<uint8_t> LabelsInputOrder() const;

 private:
  void ConvertAndSortScanLines(const Point* pts, size_t n);          // :139-171
  void PrepareDataForNextFrame();                                     // :174-204
  void InvalidateNotUsablePoints();                                   // :207-308
  void ComputeCurvature();                                            // :311-471
  void SetKeyPointsLabels();                                          // :474-590
  void EstimateAzimuthalResolution();                                 // :593-637
  bool IsScanLineAlmostEmpty(int n) const { return n < 2 * P.neighbor_width + 1; }  // SSKE.h:122
};

}  // namespace orc

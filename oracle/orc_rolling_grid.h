// CPU ORACLE (test infrastructure) — RollingGrid restatement
// (reference slam_lib/src/RollingGrid.cxx, slam_lib/include/LidarSlam/RollingGrid.h).
// The reference iterates std::unordered_map (implementation-defined order); the
// oracle's canonical order is ascending (outer 1-D index, inner 1-D index as
// signed int), which fixes the submap point order and therefore kNN tie-breaks.
#pragma once
#include "orc_matcher.h"
#include <map>

namespace orc {

struct Arr3f { float x, y, z; };
struct Arr3i { int x, y, z; };

struct RollingGrid {
  struct Voxel { Point point; unsigned count = 0; };
  using SamplingVG = std::map<int, Voxel>;
  using RollingVG = std::map<int, SamplingVG>;

  int GridSize = 50;               // RollingGrid.h:173
  double VoxelResolution = 10.;    // :176
  double LeafSize = 0.2;           // :179
  RollingVG Voxels;
  Arr3f VoxelGridPosition{0, 0, 0};
  unsigned NbPoints = 0;
  KdTree Tree;
  Cloud SubMap;
  bool TreeValid = false;          // KdTree.GetInputCloud() non-empty
  unsigned MinFramesPerVoxel = 0;
  int Sampling = LSB_SAMPLING_MAX_INTENSITY;
  double DecayingThreshold = -1;

  RollingGrid() { Reset(); }
  void Reset(Arr3f position = Arr3f{0, 0, 0});                 // RollingGrid.cxx:40-48
  void Clear();                                                  // :51-56
  void SetGridSize(int size);                                    // :59-70
  void SetVoxelResolution(double resolution);                    // :73-88
  Cloud Get(bool clean = false) const;                           // :95-114
  unsigned Size() const { return NbPoints; }
  void Roll(Arr3f minPoint, Arr3f maxPoint);                     // :117-157
  void Add(const Cloud& pc, bool fixed = false, double currentTime = -1., bool roll = true);   // :160-318
  void ClearOldPoints(double currentTime);                       // :325-350
  void BuildSubMapKdTree();                                      // :353-359
  void BuildSubMapKdTree(Arr3f minPoint, Arr3f maxPoint, int minNbPoints = -1);   // :362-442
  bool IsSubMapKdTreeValid() const { return TreeValid && !SubMap.empty(); }
  bool IsTimeThreshold() const { return DecayingThreshold > 0; }

  int To1d(Arr3i v) const { return v.z * GridSize * GridSize + v.y * GridSize + v.x; }   // :449-452
  Arr3i To3d(int id) const                                                                 // :455-463
  {
    const int z = id / (GridSize * GridSize);
    id -= z * GridSize * GridSize;
    const int y = id / GridSize;
    id -= y * GridSize;
    return Arr3i{id, y, z};
  }
  // Utils::PositionToVoxel<Array3f> (RollingGrid.h:38-42): float arithmetic (Eigen
  // demotes the double resolution to the array's float scalar), round half away.
  static Arr3i PositionToVoxel(Arr3f p, Arr3f origin, double resolution)
  {
    const float r = (float)resolution;
    return Arr3i{(int)std::round((p.x - origin.x) / r), (int)std::round((p.y - origin.y) / r), (int)std::round((p.z - origin.z) / r)};
  }
  void ResetTree() { TreeValid = false; Tree.Reset(nullptr); }
};

}  // namespace orc

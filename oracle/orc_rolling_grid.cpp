// CPU ORACLE (test infrastructure) — RollingGrid restatement.
// Reference: slam_lib/src/RollingGrid.cxx (line ranges per function).
#include "orc_rolling_grid.h"
#include <set>
#include <algorithm>

namespace orc {

void RollingGrid::Reset(Arr3f p)
{
  Clear();
  const float r = (float)VoxelResolution;
  VoxelGridPosition = Arr3f{std::floor(p.x / r) * r, std::floor(p.y / r) * r, std::floor(p.z / r) * r};
}

void RollingGrid::Clear()
{
  NbPoints = 0;
  Voxels.clear();
  ResetTree();
}

void RollingGrid::SetGridSize(int size)
{
  GridSize = size;
  Cloud prev = Get();
  Clear();
  if (!prev.empty()) Add(prev);
}

void RollingGrid::SetVoxelResolution(double resolution)
{
  VoxelResolution = int(resolution / LeafSize) * LeafSize;
  const float r = (float)VoxelResolution;
  VoxelGridPosition = Arr3f{std::floor(VoxelGridPosition.x / r) * r, std::floor(VoxelGridPosition.y / r) * r, std::floor(VoxelGridPosition.z / r) * r};
  Cloud prev = Get();
  Clear();
  if (!prev.empty()) Add(prev);
}

Cloud RollingGrid::Get(bool clean) const
{
  Cloud pc;
  pc.reserve(NbPoints);
  for (const auto& out : Voxels)
    for (const auto& in : out.second)
      if (!clean || in.second.count > MinFramesPerVoxel) pc.push_back(in.second.point);
  return pc;
}

void RollingGrid::Roll(Arr3f mn, Arr3f mx)
{
  const double halfGridSize = static_cast<double>(GridSize) / 2 * VoxelResolution;
  const float h = (float)halfGridSize;
  const float r = (float)VoxelResolution;
  auto axis = [&](float pos, float lo, float hi) {
    const float down = lo - (pos - h);
    const float up = hi - (pos + h);
    float off = (up + down) / 2.f;
    const float dmin0 = (down < 0.f) ? down : 0.f;   // downOffset.min(0)
    const float umax0 = (up > 0.f) ? up : 0.f;       // upOffset.max(0)
    off = (off < dmin0) ? dmin0 : off;               // offset.max(downOffset.min(0))
    off = (off > umax0) ? umax0 : off;               // .min(upOffset.max(0))
    return (int)std::round(off / r);
  };
  const Arr3i vo{axis(VoxelGridPosition.x, mn.x, mx.x), axis(VoxelGridPosition.y, mn.y, mx.y), axis(VoxelGridPosition.z, mn.z, mx.z)};
  if (vo.x == 0 && vo.y == 0 && vo.z == 0) return;
  unsigned newNb = 0;
  RollingVG nv;
  for (auto& out : Voxels) {
    const Arr3i id = To3d(out.first);
    const Arr3i ni{id.x - vo.x, id.y - vo.y, id.z - vo.z};
    if (ni.x >= 0 && ni.x < GridSize && ni.y >= 0 && ni.y < GridSize && ni.z >= 0 && ni.z < GridSize) {
      newNb += (unsigned)out.second.size();
      nv[To1d(ni)] = std::move(out.second);
    }
  }
  NbPoints = newNb;
  Voxels.swap(nv);
  VoxelGridPosition.x += (float)vo.x * r;
  VoxelGridPosition.y += (float)vo.y * r;
  VoxelGridPosition.z += (float)vo.z * r;
}

void RollingGrid::Add(const Cloud& pc, bool fixed, double currentTime, bool roll)
{
  if (pc.empty()) return;
  if (roll) {
    Arr3f mn{pc[0].x, pc[0].y, pc[0].z}, mx = mn;   // pcl::getMinMax3D
    for (const Point& p : pc) {
      mn.x = std::min(mn.x, p.x); mn.y = std::min(mn.y, p.y); mn.z = std::min(mn.z, p.z);
      mx.x = std::max(mx.x, p.x); mx.y = std::max(mx.y, p.y); mx.z = std::max(mx.z, p.z);
    }
    Roll(mn, mx);
  }
  const float r = (float)VoxelResolution;
  const float shift = (float)(int(GridSize / 2) * VoxelResolution);
  const Arr3f origin{VoxelGridPosition.x - shift, VoxelGridPosition.y - shift, VoxelGridPosition.z - shift};
  std::set<std::pair<int, int>> seen;
  std::map<std::pair<int, int>, Voxel> meanPts;   // CENTROID bookkeeping (:184)
  bool updated = false;
  for (const Point& point : pc) {
    const Arr3f pp{point.x, point.y, point.z};
    const Arr3i vout = PositionToVoxel(pp, origin, VoxelResolution);
    if (!(vout.x >= 0 && vout.x < GridSize && vout.y >= 0 && vout.y < GridSize && vout.z >= 0 && vout.z < GridSize)) continue;
    const Arr3f centerIn{(float)vout.x * r + origin.x, (float)vout.y * r + origin.y, (float)vout.z * r + origin.z};
    const Arr3i vin = PositionToVoxel(pp, centerIn, LeafSize);
    const int idxOut = To1d(vout);
    const int idxIn = To1d(vin);
    auto itOut = Voxels.find(idxOut);
    const bool exists = itOut != Voxels.end() && itOut->second.count(idxIn);
    if (!exists) {
      Voxels[idxOut][idxIn].point = point;
      ++NbPoints;
      updated = true;
    } else {
      Voxel& voxel = Voxels[idxOut][idxIn];
      if (voxel.point.label == 1) continue;
      switch (Sampling) {
        case LSB_SAMPLING_FIRST: break;
        case LSB_SAMPLING_LAST: voxel.point = point; updated = true; break;
        case LSB_SAMPLING_MAX_INTENSITY:
          if (point.intensity > voxel.point.intensity) { voxel.point = point; updated = true; }
          break;
        case LSB_SAMPLING_CENTER_POINT: {
          const float half = (float)(VoxelResolution / 2.f), leaf = (float)LeafSize;
          const Vec3f ctr(centerIn.x - half + leaf * (float)vin.x, centerIn.y - half + leaf * (float)vin.y, centerIn.z - half + leaf * (float)vin.z);
          if (lsb::norm(P3f(point) - ctr) < lsb::norm(P3f(voxel.point) - ctr)) { voxel.point = point; updated = true; }
          break;
        }
        case LSB_SAMPLING_CENTROID: {
          Voxel& v = meanPts[{idxOut, idxIn}];
          const float c = (float)v.count, c1 = (float)(v.count + 1);
          v.point.x = (v.point.x * c + point.x) / c1;
          v.point.y = (v.point.y * c + point.y) / c1;
          v.point.z = (v.point.z * c + point.z) / c1;
          ++v.count;
          break;
        }
      }
    }
    if (Sampling == LSB_SAMPLING_CENTROID) {   // :282-297 (re-averages on every added point, as the reference does)
      for (auto& kv : meanPts) {
        Voxel& voxel = Voxels[kv.first.first][kv.first.second];
        const float c = (float)voxel.count, c1 = (float)(voxel.count + 1);
        voxel.point.x = (voxel.point.x * c + kv.second.point.x) / c1;
        voxel.point.y = (voxel.point.y * c + kv.second.point.y) / c1;
        voxel.point.z = (voxel.point.z * c + kv.second.point.z) / c1;
      }
    }
    Voxel& voxel = Voxels[idxOut][idxIn];
    voxel.point.time = currentTime;
    voxel.point.label = fixed ? 1 : 0;
    if (seen.insert({idxOut, idxIn}).second) ++voxel.count;
  }
  if (updated) ResetTree();
}

void RollingGrid::ClearOldPoints(double currentTime)
{
  for (auto itOut = Voxels.begin(); itOut != Voxels.end();) {
    for (auto itIn = itOut->second.begin(); itIn != itOut->second.end();) {
      const Voxel& v = itIn->second;
      if (!v.point.label && currentTime - v.point.time > DecayingThreshold) itIn = itOut->second.erase(itIn);
      else ++itIn;
    }
    if (itOut->second.empty()) itOut = Voxels.erase(itOut);
    else ++itOut;
  }
}

void RollingGrid::BuildSubMapKdTree()
{
  SubMap = Get();
  Tree.Reset(&SubMap);
  TreeValid = true;
}

void RollingGrid::BuildSubMapKdTree(Arr3f mn, Arr3f mx, int minNbPoints)
{
  const float shift = (float)(int(GridSize / 2) * VoxelResolution);
  const Arr3f origin{VoxelGridPosition.x - shift, VoxelGridPosition.y - shift, VoxelGridPosition.z - shift};
  Arr3i lo = PositionToVoxel(mn, origin, VoxelResolution), hi = PositionToVoxel(mx, origin, VoxelResolution);
  lo = Arr3i{std::max(lo.x, 0), std::max(lo.y, 0), std::max(lo.z, 0)};
  hi = Arr3i{std::min(hi.x, GridSize - 1), std::min(hi.y, GridSize - 1), std::min(hi.z, GridSize - 1)};
  SubMap.clear();
  SubMap.reserve(NbPoints);
  auto inside = [&](int key) {
    const Arr3i id = To3d(key);
    return lo.x <= id.x && id.x <= hi.x && lo.y <= id.y && id.y <= hi.y && lo.z <= id.z && id.z <= hi.z;
  };
  if (minNbPoints < 0 || MinFramesPerVoxel <= 1) {
    for (const auto& out : Voxels)
      if (inside(out.first))
        for (const auto& in : out.second) SubMap.push_back(in.second.point);
  } else {
    for (const auto& out : Voxels)
      if (inside(out.first))
        for (const auto& in : out.second)
          if (in.second.count >= MinFramesPerVoxel || in.second.point.label == 1) SubMap.push_back(in.second.point);
    if ((int)SubMap.size() < minNbPoints) {
      for (const auto& out : Voxels)
        if (inside(out.first))
          for (const auto& in : out.second)
            if (in.second.count < MinFramesPerVoxel && in.second.point.label != 1) SubMap.push_back(in.second.point);
    }
  }
  Tree.Reset(&SubMap);
  TreeValid = true;
}

}  // namespace orc

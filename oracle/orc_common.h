// CPU ORACLE — test infrastructure, NOT the product.
// Dependency-free C++17/OpenMP restatement of the LidarSlam per-frame hot path
// (reference files cited function by function).  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may call into this directory.
//
// PARITY UNPINNED: the reference cannot be built here (Eigen/PCL/Ceres/nanoflann are
// absent, no network) and ships no golden vectors (ros_wrapping/tests/README.md:3-4),
// so this restatement is checked line-by-line against the reference sources and
// against independent numpy/scipy computations (tests/test_oracle_*.py), not against
// outputs of the reference binary.
//
// The third-party arithmetic (PCL covariance + eigen33, Eigen quaternion/slerp,
// deterministic trig) comes from lidarslam_b200/csrc/common/*.h, which is shared
// with the CUDA kernels on purpose: GPU and oracle must execute the same IEEE
// operation sequence to be bit-comparable.  All algorithm layers are written
// independently here, sequentially, following the reference's control flow.
#pragma once
#include <cstdint>
#include <cstddef>
#include <vector>
#include <cmath>
#include <limits>
#include "../include/lidarslam_b200.h"
#include "../lidarslam_b200/csrc/common/pca.h"

namespace orc {

using Point = lsb_point;
using Cloud = std::vector<Point>;
using lsb::Vec3f; using lsb::Vec3d; using lsb::Mat3f; using lsb::Mat3d; using lsb::Iso3d;

inline Vec3f P3f(const Point& p) { return Vec3f(p.x, p.y, p.z); }

// Utils::ComputeMeanAndPCA (Utilities.h:247-262) over cloud[indices] in order
template <typename T>
inline void MeanAndPCA(const Cloud& cloud, const int* idx, int n, lsb::Vec3<T>& mean, lsb::Mat3<T>& evecs, T* evals)
{
  lsb::CovAccum<T> acc;
  for (int i = 0; i < n; ++i) { const Point& p = cloud[idx[i]]; acc.add(p.x, p.y, p.z); }
  lsb::Mat3<T> cov;
  acc.finish(mean, cov);
  lsb::eigen33(cov, evecs, evals);
}

inline double deg2rad(double d) { return d / 180. * M_PI; }

}  // namespace orc

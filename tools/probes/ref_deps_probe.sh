#!/bin/bash
# Looks for the libraries the reference's slam_lib needs (slam_lib/CMakeLists.txt:14-42: Eigen3, PCL, Ceres,
# nanoflann) on the machine it runs on and writes what it finds.  BASELINE.md section 2 asks for this on
# the GPU box: if all four were present the reference library could be compiled there as the CPU arm.
out=${1:-gpurun_out/r02_ref_deps_probe.txt}
mkdir -p "$(dirname "$out")"
{
  echo "host: $(uname -srm)  cores: $(nproc)  date: $(date -u +%FT%TZ)"
  nvidia-smi --query-gpu=name,driver_version --format=csv,noheader 2>/dev/null
  echo "== headers / libraries (find over the whole file system, /proc and /sys skipped)"
  for pat in "Eigen" "eigen3" "nanoflann.hpp" "ceres" "libceres*" "pcl" "pcl-*" "libpcl_*" "PCLConfig.cmake" "CeresConfig.cmake" "Eigen3Config.cmake" "nanoflannConfig.cmake"; do
    n=$(find / -xdev \( -path /proc -o -path /sys \) -prune -o -name "$pat" -print 2>/dev/null | grep -v "^$GRAFT_REPO_ROOT" | head -5)
    echo "-- $pat: ${n:-none}"
  done
  echo "== pkg-config"
  for p in eigen3 ceres nanoflann pcl_common pcl_common-1.10 pcl_common-1.12; do
    if command -v pkg-config >/dev/null 2>&1; then pkg-config --exists $p && echo "$p: $(pkg-config --modversion $p)" || echo "$p: not found"; else echo "pkg-config itself: not installed"; break; fi
  done
  echo "== package manager"
  (dpkg -l 2>/dev/null | grep -Ei "eigen|libpcl|ceres|nanoflann" || echo "dpkg: no eigen/pcl/ceres/nanoflann packages")
  echo "== compile test: #include <Eigen/Dense>, <pcl/point_cloud.h>, <ceres/ceres.h>, <nanoflann.hpp>"
  for h in "Eigen/Dense" "pcl/point_cloud.h" "ceres/ceres.h" "nanoflann.hpp"; do
    echo "#include <$h>" | g++ -x c++ -fsyntax-only - >/dev/null 2>&1 && echo "$h: compiles" || echo "$h: fatal error, header not found"
  done
  echo "== reference tree on this machine"
  ls -d /root/reference 2>/dev/null || echo "/root/reference: absent"
  echo "== verdict"
  echo "reference library build: unavailable on this box unless all four 'compiles' above"
} > "$out" 2>&1
cat "$out"

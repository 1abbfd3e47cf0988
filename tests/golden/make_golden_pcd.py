"""Writes the committed PCD fixtures of tests/test_pcd.py::test_golden_files with the numpy restatement
(oracle/pcd_py.py): 64 LidarPoints in the three PCD data formats + the cloud itself (lidar_points.npy).
Run from the repository root: python tests/golden/make_golden_pcd.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from lidarslam_b200 import capi  # noqa: E402
from oracle import pcd_py  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
r = np.random.default_rng(20261104)
n = 64
p = np.zeros(n, dtype=capi.POINT_DTYPE)
p["x"], p["y"], p["z"] = (r.normal(0, 30, n).astype(np.float32) for _ in range(3))
p["w"] = 1.0
p["time"] = 1.7e9 + r.uniform(0, 1e3, n)
p["intensity"] = r.uniform(0, 255, n).astype(np.float32)
p["laser_id"] = r.integers(0, 128, n)
p["device_id"] = r.integers(0, 2, n)
p["label"] = r.integers(0, 2, n)
p["z"][5] = np.nan
np.save(os.path.join(HERE, "lidar_points.npy"), p)
for fmt, name in enumerate(("ascii", "binary", "binary_compressed")):
    pcd_py.write(os.path.join(HERE, "lidar_points_%s.pcd" % name), p, fmt)
print("wrote", n, "points")

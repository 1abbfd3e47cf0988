"""The C-ABI library loads and exports every symbol include/lidarslam_b200.h declares; the
product fails loudly without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "lidarslam_b200.h")).read()
    return sorted(set(re.findall(r"LSB_API[^;]*?\b(lsb_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(built):
    from lidarslam_b200 import capi
    lib = capi.load_library()
    names = _declared()
    assert len(names) >= 45
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_point_layout_matches_lidarpoint():
    from lidarslam_b200 import capi
    d = capi.POINT_DTYPE
    assert d.itemsize == 32
    assert [d.fields[n][1] for n in ("x", "y", "z", "w", "time", "intensity", "laser_id", "device_id", "label")] == [0, 4, 8, 12, 16, 24, 28, 30, 31]


def test_defaults_follow_reference(built):
    from lidarslam_b200 import capi
    p = capi.default_params()
    assert list(p.use_keypoints) == [1, 1, 0] and p.ego_motion == 1 and p.undistortion == 2      # Slam.h:412,417
    assert (p.ego_motion_icp_max_iter, p.localization_icp_max_iter, p.localization_lm_max_iter) == (4, 3, 15)
    assert (p.localization_edge_nb_neighbors, p.localization_plane_nb_neighbors, p.localization_blob_nb_neighbors) == (10, 5, 10)
    assert p.min_nb_matched_keypoints == 20 and p.kf_distance_threshold == 0.5 and p.kf_angle_threshold == 5.0
    k = capi.default_ke_params()
    assert k.neighbor_width == 4 and abs(k.edge_sin_angle_threshold - 0.86) < 1e-7 and k.edge_intensity_gap_threshold == 50.0


def test_no_cpu_fallback(built):
    from lidarslam_b200 import capi
    lib = capi.load_library()
    if lib.lsb_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    rc = lib.lsb_create(C.byref(h), C.c_int(0))
    assert rc == -4 and not h.value                       # LSB_E_NO_DEVICE
    lib.lsb_last_error.restype = C.c_char_p
    assert b"no CPU fallback" in lib.lsb_last_error()
    from lidarslam_b200.slam import Slam
    with pytest.raises(RuntimeError):
        Slam()


def test_product_does_not_import_oracle():
    """Nothing under lidarslam_b200/ imports, includes, links or loads anything under oracle/."""
    pkg = os.path.join(ROOT, "lidarslam_b200")
    bad = re.compile(r"(^\s*(from|import)\s+oracle\b)|(import\s+pyorc)|(#include\s+\"[^\"]*oracle/)|(liborc)|(CDLL\([^)]*oracle)", re.M)
    for d, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(d, f), errors="ignore").read()
                assert not bad.search(txt), os.path.join(d, f)

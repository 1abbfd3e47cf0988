"""ctypes front-end of the synthetic spinning-LiDAR generator (tools/synth/synth_lidar.cpp).
Test/bench infrastructure: produces the BASELINE.json sensor geometries."""
import ctypes as C
import os
import subprocess

import numpy as np

from lidarslam_b200.capi import POINT_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class Sensor(C.Structure):
    _fields_ = [("n_beams", C.c_int), ("n_cols", C.c_int), ("elev_min_deg", C.c_double), ("elev_max_deg", C.c_double),
                ("device_id", C.c_int), ("mount_xyz", C.c_double * 3), ("mount_rpy", C.c_double * 3)]


def sensor(name, device_id=0, mount_xyz=(0, 0, 0), mount_rpy=(0, 0, 0), n_cols=None):
    table = {
        "vlp16": (16, 1800, -15.0, 15.0),
        "hdl64": (64, 2083, -24.8, 2.0),
        "os128": (128, 2048, -22.5, 22.5),
        "vls128": (128, 1800, -25.0, 15.0),
        "pandar64": (64, 1800, -25.0, 15.0),
        "tiny": (16, 512, -15.0, 15.0),
    }
    b, c, lo, hi = table[name]
    s = Sensor(b, n_cols or c, lo, hi, device_id)
    s.mount_xyz[:] = mount_xyz
    s.mount_rpy[:] = mount_rpy
    return s


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def load():
    global _LIB
    if _LIB is None:
        p = os.path.join(_HERE, "libsynth.so")
        if not os.path.exists(p):
            build()
        _LIB = C.CDLL(p)
        _LIB.synth_create.restype = C.c_void_p
    return _LIB


class Sequence:
    def __init__(self, scene_seed=1234, sequence_id=0, speed=8.0, period=0.1, range_noise=0.01):
        self.lib = load()
        self.h = C.c_void_p(self.lib.synth_create(C.c_uint64(scene_seed), C.c_uint64(sequence_id)))
        self.lib.synth_set_motion(self.h, C.c_double(speed), C.c_double(period), C.c_double(range_noise))

    def frame(self, sens, idx):
        """returns (points[n_beams*n_cols], gt_pose 4x4 BASE-in-WORLD at stamp, stamp_us)"""
        out = np.zeros(sens.n_beams * sens.n_cols, dtype=POINT_DTYPE)
        T = np.zeros(16)
        stamp = C.c_uint64()
        self.lib.synth_frame(self.h, C.byref(sens), C.c_int(idx), out.ctypes.data_as(C.c_void_p),
                             T.ctypes.data_as(C.c_void_p), C.byref(stamp))
        return out, T.reshape(4, 4), stamp.value

    def __del__(self):
        try:
            self.lib.synth_destroy(self.h)
        except Exception:
            pass

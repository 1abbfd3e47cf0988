"""Probe: does lsb_destroy give a handle's memory back to the device's stream-ordered pool?"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from lidarslam_b200.slam import Slam  # noqa: E402
from tools.synth import synth  # noqa: E402

rt = ctypes.CDLL("libcudart.so.12")


def pool():
    p, v, out = ctypes.c_void_p(), ctypes.c_uint64(), []
    rt.cudaDeviceGetDefaultMemPool(ctypes.byref(p), 0)
    for attr in (5, 7):
        rt.cudaMemPoolGetAttribute(p, attr, ctypes.byref(v))
        out.append(v.value >> 20)
    return tuple(out)


torch.zeros(1, device="cuda")
sens = synth.sensor("hdl64")
seq = synth.Sequence(scene_seed=5)
frames = [seq.frame(sens, j) for j in range(4)]
for rep in range(4):
    print("rep", rep, "before create", pool())
    h = Slam(device=0)
    print("   after create", pool())
    for j, (pts, _, stamp) in enumerate(frames):
        h.AddFrame(pts, stamp, j + 1)
    print("   after 4 frames", pool())
    if rep == 1:
        h.GetMap(1), h.GetRegisteredFrame()
        print("   after getters", pool())
    h.close()
    print("   after destroy", pool())

"""Python handle on the C ABI with the method names of LidarSlam::Slam
(reference Slam.h:98-394).  Test / bench convenience only: the product is the
shared library; this file adds no computation."""
import ctypes as C

import numpy as np

from . import capi


class Slam(capi.CApi):
    def __init__(self, device=0):
        super().__init__(capi.load_library(), "lsb_", (device,))
        self._dev_bufs = []

    # ---- device-resident input (bench: inputs already in HBM) -----------------
    def DevUpload(self, pts):
        pts = capi.as_points(pts)
        f = self.lib.lsb_dev_alloc
        f.restype = C.c_void_p
        d = f(self.h, C.c_size_t(max(pts.nbytes, 32)))
        if not d:
            raise RuntimeError("lsb_dev_alloc failed")
        self._check(self.lib.lsb_dev_upload(self.h, C.c_void_p(d), capi._ptr(pts), C.c_size_t(pts.nbytes)))
        self._dev_bufs.append(d)
        return d

    def DevFree(self, d):
        self.lib.lsb_dev_free(self.h, C.c_void_p(d))
        self._dev_bufs.remove(d)

    def AddFramesDev(self, dev_ptrs, sizes, stamps_us, seqs=None):
        n = len(dev_ptrs)
        ptrs = (C.c_void_p * n)(*dev_ptrs)
        sz = (C.c_size_t * n)(*[int(s) for s in sizes])
        stamps = (C.c_uint64 * n)(*[int(s) for s in stamps_us])
        sq = (C.c_uint32 * n)(*[int(s) for s in (seqs or [0] * n)])
        return self._check(self.lib.lsb_add_frames_dev(self.h, C.c_int(n), ptrs, sz, stamps, sq))

    # ---- pipelining (include/lidarslam_b200.h: no counterpart in the reference) --------
    def _frame_args(self, ptrs, sizes, stamps_us, seqs):
        n = len(ptrs)
        return (C.c_int(n), (C.c_void_p * n)(*ptrs), (C.c_size_t * n)(*[int(s) for s in sizes]),
                (C.c_uint64 * n)(*[int(s) for s in stamps_us]), (C.c_uint32 * n)(*[int(s) for s in (seqs or [0] * n)]))

    def PrefetchFramesDev(self, dev_ptrs, sizes, stamps_us, seqs=None):
        return self._check(self.lib.lsb_prefetch_frames_dev(self.h, *self._frame_args(dev_ptrs, sizes, stamps_us, seqs)))

    def PrefetchFrames(self, frames, stamps_us, seqs=None):
        """frames: numpy point arrays that stay alive and unchanged until the matching AddFrames returns"""
        frames = [capi.as_points(f) for f in frames]
        self._staged_keep = getattr(self, "_staged_keep", []) + [frames]
        self._staged_keep = self._staged_keep[-4:]
        return self._check(self.lib.lsb_prefetch_frames(self.h, *self._frame_args([f.ctypes.data for f in frames], [len(f) for f in frames], stamps_us, seqs)))

    def SetPipelining(self, defer_map_update):
        self._check(self.lib.lsb_set_pipelining(self.h, C.c_int(1 if defer_map_update else 0)))

    def Synchronize(self):
        self._check(self.lib.lsb_synchronize(self.h))

    def TimerBegin(self):
        self._check(self.lib.lsb_timer_begin(self.h))

    def TimerEnd(self):
        ms = C.c_float()
        self._check(self.lib.lsb_timer_end(self.h, C.byref(ms)))
        return ms.value

    def FlushL2(self):
        self._check(self.lib.lsb_dev_flush_l2(self.h))

    def close(self):
        if self.h:
            for d in list(self._dev_bufs):
                self.lib.lsb_dev_free(self.h, C.c_void_p(d))
            self._dev_bufs = []
        super().close()


def pinned_points(n):
    """numpy view on a pinned host buffer of n points (fast H2D for AddFrames)."""
    lib = capi.load_library()
    f = lib.lsb_host_alloc
    f.restype = C.c_void_p
    nbytes = max(n, 1) * capi.POINT_DTYPE.itemsize
    p = f(C.c_size_t(nbytes))
    if not p:
        raise RuntimeError("lsb_host_alloc failed")
    buf = (C.c_char * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=capi.POINT_DTYPE, count=n)
    return arr, p


def pinned_free(p):
    lib = capi.load_library()
    lib.lsb_host_free(C.c_void_p(p))

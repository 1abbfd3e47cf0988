"""lsb_knn_search_noalloc: the exact kNN entry point that allocates nothing (static __device__ scratch; no counterpart in
the reference).  Checked through the C ABI against the existing entry point
lsb_knn_search on two calls in a row: indices and distances bit-exact (index / integer work; the distances are the same
float operations in the same order).  This file is named to be collected LAST."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _call(lib, torch, cloud_d, m, q, k):
    """one call of the new entry point on device buffers (torch owns them; the entry point itself allocates nothing)"""
    q_d = torch.from_numpy(np.ascontiguousarray(q, dtype=np.float32)).cuda()
    idx_d = torch.full((len(q), k), -7, dtype=torch.int32, device="cuda")
    d2_d = torch.full((len(q), k), -7.0, dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.lsb_knn_search_noalloc(C.c_int(0), C.c_void_p(cloud_d.data_ptr() if m else None), C.c_size_t(m), C.c_void_p(q_d.data_ptr()),
                                    C.c_size_t(len(q)), C.c_int(k), C.c_void_p(idx_d.data_ptr()), C.c_void_p(d2_d.data_ptr()),
                                    C.c_void_p(stream))
    torch.cuda.synchronize()
    return rc, idx_d.cpu().numpy(), d2_d.cpu().numpy()


def test_knn_search_noalloc_matches_knn_search(gpu, synth):
    import torch
    from lidarslam_b200 import capi
    lib = capi.load_library()
    lib.lsb_knn_search_noalloc.restype = C.c_int
    lib.lsb_last_error.restype = C.c_char_p
    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    pts, _, _ = synth.Sequence(scene_seed=1234).frame(synth.sensor("vlp16"), 0)
    rng = np.random.default_rng(17)
    cloud = pts[rng.permutation(len(pts))[:6000]].copy()
    cloud[10:20] = cloud[0]                                                      # ties at equal distance: smaller index first
    xyz = np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)
    qa = np.concatenate([xyz[:1500] + rng.normal(0, 0.3, (1500, 3)).astype(np.float32), [[500., 500., 50.], [-1000., 3., 2.]], xyz[:5]]).astype(np.float32)
    qb = (xyz[2000:4048] + rng.normal(0, 1.0, (2048, 3)).astype(np.float32)).astype(np.float32)   # the maximum number of queries
    g = gpu()
    cloud_d = torch.from_numpy(cloud.view(np.uint8).reshape(-1)).cuda()
    # two calls in a row on the shared scratch array (different sizes, different k: the second finds the first one's contents)
    for q, k in ((qa, 10), (qb, 16), (qa, 1), (qa, 10)):
        ref_idx, ref_d2 = g.KnnSearch(cloud, q, k)
        rc, idx, d2 = _call(lib, torch, cloud_d, len(cloud), q, k)
        assert rc == 0, lib.lsb_last_error()
        assert np.array_equal(idx, ref_idx), "indices differ at %s" % np.argwhere(idx != ref_idx)[:5]
        assert np.array_equal(d2.view(np.uint32), ref_d2.view(np.uint32))
    # fewer points than k, and an empty cloud: -1 / +inf beyond the number found, as lsb_knn_search
    small = cloud[:7].copy()
    small_d = torch.from_numpy(small.view(np.uint8).reshape(-1)).cuda()
    ref_idx, ref_d2 = g.KnnSearch(small, qa[:64], 10)
    rc, idx, d2 = _call(lib, torch, small_d, 7, qa[:64], 10)
    assert rc == 0 and np.array_equal(idx, ref_idx) and np.array_equal(d2.view(np.uint32), ref_d2.view(np.uint32))
    assert (idx[:, 7:] == -1).all() and np.isinf(d2[:, 7:]).all()
    rc, idx, d2 = _call(lib, torch, small_d, 0, qa[:8], 3)
    assert rc == 0 and (idx == -1).all() and np.isinf(d2).all()
    # sizes beyond the static array are refused with a clear error, nothing is written
    too_many = np.zeros((2049, 3), np.float32)
    rc, idx, _ = _call(lib, torch, cloud_d, len(cloud), too_many, 4)
    assert rc == -3 and b"2048" in lib.lsb_last_error() and (idx == -7).all()                    # LSB_E_CAPACITY
    rc, idx, _ = _call(lib, torch, cloud_d, len(cloud), qa[:4], 17)
    assert rc == -3 and b"LSB_NOALLOC_MAX_K" in lib.lsb_last_error() and (idx == -7).all()

"""GPU parity tests, stage by stage, through the C ABI (kernel-level entry points)
against the CPU oracle on identical seeded inputs.  Bit-exact for labels, keypoints,
neighbour indices, map contents; pose within 1e-6 (north star)."""
import numpy as np
import pytest

from lidarslam_b200 import capi

pytestmark = pytest.mark.gpu


def _frame(synth, sensor="tiny", idx=0, seed=1234):
    seq = synth.Sequence(scene_seed=seed)
    return seq.frame(synth.sensor(sensor), idx)


@pytest.mark.parametrize("sensor", ["tiny", "vlp16", "hdl64", "os128", "vls128", "pandar64"])
def test_extractor_bit_exact(gpu, oracle_lib, synth, sensor):
    pts, _, _ = _frame(synth, sensor)
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    lg, cg, kg = g.ExtractKeypoints(pts)
    lo, co, ko = o.ExtractKeypoints(pts)
    assert g.GetExtractorParams().azimuthal_resolution == o.GetExtractorParams().azimuthal_resolution
    for f, name in enumerate(["sin_angle", "saliency", "depth_gap", "intensity_gap", "edge", "plane", "blob", "edge_valid", "plane_valid", "blob_valid"]):
        a, b = g.GetKeDebug(f), o.GetKeDebug(f)
        bad = np.flatnonzero(a.view(np.uint32) != b.view(np.uint32))
        assert bad.size == 0, "%s differs at %d points, first %s: gpu %s oracle %s" % (name, bad.size, bad[:5], a[bad[:5]], b[bad[:5]])
    assert cg == co
    assert np.array_equal(lg, lo)
    for k in range(3):
        assert kg[k].tobytes() == ko[k].tobytes()
    assert co[0] > 0 and co[1] > 0 and co[2] > 0


def test_extractor_edge_cases(gpu, oracle_lib, synth):
    pts, _, _ = _frame(synth, "tiny")
    g, o = gpu(), oracle_lib.Oracle(nb_threads=2)
    # ragged rings: drop most of ring 3 (almost empty scan line), shuffle arrival order of ring 5, duplicate points
    keep = ~((pts["laser_id"] == 3) & (np.arange(len(pts)) % 97 != 0))
    p2 = pts[keep].copy()
    p2 = np.concatenate([p2, p2[:40]])
    lg, cg, kg = g.ExtractKeypoints(p2)
    lo, co, ko = o.ExtractKeypoints(p2)
    assert cg == co and np.array_equal(lg, lo)
    for k in range(3):
        assert kg[k].tobytes() == ko[k].tobytes()
    # too small input: every ring is "almost empty"
    lg, cg, _ = g.ExtractKeypoints(pts[:40])
    lo, co, _ = o.ExtractKeypoints(pts[:40])
    assert cg == co == [0, 0, 0] and np.array_equal(lg, lo)


@pytest.mark.parametrize("k", [1, 5, 10])
def test_knn_bit_exact(gpu, oracle_lib, synth, k):
    pts, _, _ = _frame(synth, "vlp16")
    rng = np.random.default_rng(7)
    cloud = pts[rng.permutation(len(pts))[:6000]].copy()
    q = np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)[:3000] + rng.normal(0, 0.3, (3000, 3)).astype(np.float32)
    q = np.concatenate([q, [[500., 500., 50.], [-1000., 3., 2.]], q[:5]]).astype(np.float32)   # far outliers + exact duplicates
    cloud[10:20] = cloud[0]                                                                      # ties at equal distance
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    ig, dg = g.KnnSearch(cloud, q, k)
    io, do = o.KnnSearch(cloud, q, k)
    assert np.array_equal(ig, io)
    assert np.array_equal(dg.view(np.uint32), do.view(np.uint32))
    # fewer points than k
    ig, dg = g.KnnSearch(cloud[:3], q[:50], 5)
    io, do = o.KnnSearch(cloud[:3], q[:50], 5)
    assert np.array_equal(ig, io) and np.array_equal(dg.view(np.uint32), do.view(np.uint32))


@pytest.mark.gpu
@pytest.mark.parametrize("k", [5, 10])
def test_knn_bit_exact_large_query_set(gpu, oracle_lib, synth, k):
    """More than 9.5 k queries take the one-thread-per-query stream kernel (K8) instead of the 4-lane one;
    sparse queries exercise the sphere-bounded warp pass and the brute-force fallback."""
    pts, _, _ = _frame(synth, "hdl64")
    rng = np.random.default_rng(11)
    perm = rng.permutation(len(pts))
    cloud = pts[perm[:50000]].copy()
    qp = pts[perm[50000:50000 + 30000]]
    q = np.stack([qp["x"], qp["y"], qp["z"]], 1).astype(np.float32)
    q[::7] += rng.normal(0, 2.0, (len(q[::7]), 3)).astype(np.float32)      # pushed off the surfaces: sparse neighbourhoods
    q[::501] += np.float32(40.0)                                            # isolated
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    ig, dg = g.KnnSearch(cloud, q, k)
    io, do = o.KnnSearch(cloud, q, k)
    assert np.array_equal(ig, io)
    assert np.array_equal(dg.view(np.uint32), do.view(np.uint32))


@pytest.mark.parametrize("case", ["lattice", "collinear", "identical", "clusters", "huge_extent", "tiny"])
def test_knn_degenerate_clouds(gpu, oracle_lib, case):
    """Exact kNN with (d2, index) ordering on clouds that stress ties, empty cells and the fall-back passes; results against the
    oracle's kd-tree and an O(NQ x M) numpy scan with the same float arithmetic ((dx*dx + dy*dy) + dz*dz in f32)."""
    rng = np.random.default_rng(5)
    if case == "lattice":          # integer lattice: dozens of exact ties per query, decided by the smaller index
        g3 = np.stack(np.meshgrid(np.arange(12), np.arange(12), np.arange(6), indexing="ij"), -1).reshape(-1, 3).astype(np.float32) * 0.5
        xyz = g3[rng.permutation(len(g3))]
        q = np.concatenate([xyz[:200], xyz[:200] + np.float32(0.25)])
    elif case == "collinear":
        t = rng.uniform(-30, 30, 3000).astype(np.float32)
        xyz = np.stack([t, 0.5 * t, np.zeros_like(t)], 1)
        q = np.stack([rng.uniform(-35, 35, 500), rng.uniform(-20, 20, 500), rng.uniform(-3, 3, 500)], 1).astype(np.float32)
    elif case == "identical":      # every target point at the same place
        xyz = np.tile(np.float32([3.0, -2.0, 1.0]), (500, 1))
        q = rng.normal(0, 5, (100, 3)).astype(np.float32)
    elif case == "clusters":       # dense blobs far apart: most cells empty, isolated queries between them
        c = rng.uniform(-80, 80, (12, 3)).astype(np.float32)
        xyz = (c[rng.integers(0, 12, 8000)] + rng.normal(0, 0.2, (8000, 3))).astype(np.float32)
        q = np.concatenate([xyz[:1500] + rng.normal(0, 0.05, (1500, 3)).astype(np.float32), rng.uniform(-100, 100, (500, 3)).astype(np.float32)])
    elif case == "huge_extent":    # the cell size has to grow to respect the cell budget
        xyz = rng.uniform(-3000, 3000, (5000, 3)).astype(np.float32)
        q = rng.uniform(-3500, 3500, (800, 3)).astype(np.float32)
    else:                          # fewer points than k, one point, queries on top of the points
        xyz = rng.normal(0, 1, (7, 3)).astype(np.float32)
        q = np.concatenate([xyz, rng.normal(0, 10, (20, 3)).astype(np.float32)])
    cloud = np.zeros(len(xyz), dtype=capi.POINT_DTYPE)
    cloud["x"], cloud["y"], cloud["z"], cloud["w"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], 1.0
    q = np.ascontiguousarray(q, dtype=np.float32)
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    for k in (1, 5, 10, 32):
        ig, dg = g.KnnSearch(cloud, q, k)
        io, do = o.KnnSearch(cloud, q, k)
        assert np.array_equal(ig, io), (case, k)
        assert np.array_equal(dg.view(np.uint32), do.view(np.uint32)), (case, k)
        # independent scan: distances in float32 with the reference's summation order, ties to the smaller index
        d = q[:, None, :] - xyz[None, :, :]
        d2 = (d[..., 0] * d[..., 0] + d[..., 1] * d[..., 1]) + d[..., 2] * d[..., 2]
        kk = min(k, len(xyz))
        order = np.lexsort((np.broadcast_to(np.arange(len(xyz)), d2.shape), d2), axis=1)[:, :kk]
        assert np.array_equal(ig[:, :kk], order.astype(np.int32)), (case, k)
        assert np.array_equal(dg[:, :kk], np.take_along_axis(d2, order, 1)), (case, k)


def test_undistort_bit_exact(gpu, oracle_lib, synth):
    pts, _, _ = _frame(synth, "tiny")
    H0 = np.eye(4)
    c, s = np.cos(0.02), np.sin(0.02)
    H1 = np.array([[c, -s, 0, 0.8], [s, c, 0, 0.05], [0, 0, 1, -0.01], [0, 0, 0, 1.]])
    g, o = gpu(), oracle_lib.Oracle()
    a = g.Undistort(pts, H0, H1, -0.1, 0.0)
    b = o.Undistort(pts, H0, H1, -0.1, 0.0)
    assert a.tobytes() == b.tobytes()
    assert a.tobytes() != pts.tobytes()


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_map_add_bit_exact(gpu, oracle_lib, synth, mode):
    g, o = gpu(), oracle_lib.Oracle()
    for s in (g, o):
        for k in range(3):
            s.SetVoxelGridSamplingMode(k, mode)
    seq = synth.Sequence()
    sens = synth.sensor("vlp16")
    for i in range(3):
        pts, T, _ = seq.frame(sens, i * 40)
        w = pts.copy()
        xyz = np.stack([pts["x"], pts["y"], pts["z"]], 1).astype(np.float64) @ T[:3, :3].T + T[:3, 3]
        w["x"], w["y"], w["z"] = xyz[:, 0].astype(np.float32), xyz[:, 1].astype(np.float32), xyz[:, 2].astype(np.float32)
        for k, leaf in ((0, 0.3), (1, 0.6)):
            g.MapAdd(k, w[k::2], False, 0.1 * i)
            o.MapAdd(k, w[k::2], False, 0.1 * i)
            a, b = g.GetMap(k), o.GetMap(k)
            assert len(a) == len(b) and len(a) > 1000
            assert a.tobytes() == b.tobytes(), "map %d differs after add %d" % (k, i)


def test_map_centroid_sampling(gpu, oracle_lib):
    """SamplingMode::CENTROID (RollingGrid.cxx:266-297): the per-point re-averaging couples the leaves of one Add call
    through the arrival order.  Dense clusters (many points per leaf), leaves touched over several calls (count > 1),
    fixed points (later points of a fixed leaf are skipped and do not advance the re-averaging), out-of-grid points."""
    rng = np.random.default_rng(5)

    def cloud(n, spread, centres):
        p = np.zeros(n, dtype=capi.POINT_DTYPE)
        c = np.asarray(centres, np.float64)[rng.integers(0, len(centres), n)]
        xyz = (c + rng.normal(0, spread, (n, 3))).astype(np.float32)
        p["x"], p["y"], p["z"], p["w"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], 1.0
        p["intensity"] = rng.uniform(0, 255, n).astype(np.float32)
        p["laser_id"] = rng.integers(0, 16, n)
        return p

    g, o = gpu(), oracle_lib.Oracle()
    for s in (g, o):
        s.SetVoxelGridSamplingMode(1, 4)
    centres = rng.uniform(-30, 30, (40, 3))
    steps = [(4000, 0.4, False), (4000, 0.4, False), (3000, 0.8, True), (5000, 0.5, False), (2000, 300.0, False), (4000, 0.3, False)]
    for i, (n, spread, fixed) in enumerate(steps):
        c = cloud(n, spread, centres)
        g.MapAdd(1, c, fixed, 0.1 * i)
        o.MapAdd(1, c, fixed, 0.1 * i)
        a, b = g.GetMap(1), o.GetMap(1)
        assert len(a) == len(b) and len(a) > 500
        assert a.tobytes() == b.tobytes(), "centroid map differs after add %d" % i
    assert (a["label"] == 1).any() and (a["label"] == 0).any()
    # the stored points are averages, not input points
    assert g.GetMap(1, True).tobytes() == o.GetMap(1, True).tobytes()


def test_map_roll_drops_voxels(gpu, oracle_lib):
    """A 8^3 grid of 5 m voxels (40 m across): clouds that fit the grid make it roll to their bounding box and drop the voxels
    that leave it (RollingGrid.cxx:117-157); points outside the grid are skipped (:190-196); re-binning on SetVoxelGridSize /
    SetVoxelGridResolution (:59-88)."""
    rng = np.random.default_rng(21)

    def cloud(centre, half, n):
        p = np.zeros(n, dtype=capi.POINT_DTYPE)
        xyz = (rng.uniform(-half, half, (n, 3)) + np.asarray(centre)).astype(np.float32)
        p["x"], p["y"], p["z"], p["w"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], 1.0
        p["intensity"] = rng.uniform(0, 255, n).astype(np.float32)
        return p

    g, o = gpu(), oracle_lib.Oracle()
    for s in (g, o):
        s.SetVoxelGridSize(8)
        s.SetVoxelGridResolution(5.0)
    sizes = []
    steps = [((0, 0, 0), 14.0), ((22, 3, -2), 14.0), ((47, -20, 4), 12.0), ((47, -20, 4), 30.0), ((-5, 0, 0), 6.0)]
    for i, (centre, half) in enumerate(steps):
        c = cloud(centre, half, 6000)
        g.MapAdd(1, c, False, 0.1 * i)
        o.MapAdd(1, c, False, 0.1 * i)
        a, b = g.GetMap(1), o.GetMap(1)
        assert a.tobytes() == b.tobytes(), "map differs after add %d" % i
        sizes.append(len(a))
        for ax in ("x", "y", "z"):
            assert a[ax].max() - a[ax].min() <= 45.0 + 1e-3          # never wider than the grid (+ one voxel of slack)
    assert sizes[0] > 3000 and sizes[1] < sizes[0] + 6000           # the second add rolled the grid and dropped part of the first cloud
    assert (g.GetMap(1)["x"] < 20).any()                             # the last cloud pulled the grid back
    # re-binning: a larger grid keeps everything, a coarser resolution re-snaps the voxels
    for s in (g, o):
        s.SetVoxelGridSize(20)
        s.SetVoxelGridResolution(10.0)
    assert g.GetMap(1).tobytes() == o.GetMap(1).tobytes()
    c = cloud((60, 0, 0), 40.0, 8000)
    g.MapAdd(1, c, True, 1.0)
    o.MapAdd(1, c, True, 1.0)
    a = g.GetMap(1)
    assert a.tobytes() == o.GetMap(1).tobytes() and (a["label"] == 1).any()


def test_match_and_solve(gpu, oracle_lib, synth):
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    seq = synth.Sequence()
    sens = synth.sensor("vlp16")
    p0, T0, _ = seq.frame(sens, 0)
    p1, T1, _ = seq.frame(sens, 1)
    _, _, k0 = o.ExtractKeypoints(p0)
    _, _, k1 = o.ExtractKeypoints(p1)
    prior = np.linalg.inv(T0) @ T1
    prior[:3, 3] += [0.05, -0.03, 0.01]
    for single in (0, 1):
        mp = capi.MatchParams(single, 5.0, 8 if single else 10, 3 if single else 4, 0.2, 5, 0.04, 0.2, 10, 2.0)
        rg = g.MatchAndSolve(k1, k0, mp, 15, prior)
        ro = o.MatchAndSolve(k1, k0, mp, 15, prior)
        assert rg["rc"] == ro["rc"] == 0
        assert np.array_equal(rg["status"], ro["status"])
        assert np.array_equal(rg["weight"].view(np.uint64), ro["weight"].view(np.uint64))
        assert (rg["status"] == 0).sum() > 500
        assert rg["iterations"] == ro["iterations"] and rg["accepted"] == ro["accepted"]
        # same operations in the same order on both sides (radix-32 sum tree, warp-spread Cholesky): bit-identical
        assert np.array_equal(rg["pose"], ro["pose"])
        assert rg["cost"] == ro["cost"]
        assert np.allclose(np.array(rg["err"].covariance), np.array(ro["err"].covariance), rtol=1e-6, atol=1e-12)


def _pool_mib():
    """(reserved, used) MiB of device 0's default stream-ordered memory pool, asked from the CUDA runtime directly."""
    import ctypes
    rt = ctypes.CDLL("libcudart.so.12")
    pool, v, out = ctypes.c_void_p(), ctypes.c_uint64(), []
    assert rt.cudaDeviceGetDefaultMemPool(ctypes.byref(pool), 0) == 0
    for attr in (5, 7):   # cudaMemPoolAttrReservedMemCurrent, cudaMemPoolAttrUsedMemCurrent
        assert rt.cudaMemPoolGetAttribute(pool, attr, ctypes.byref(v)) == 0
        out.append(v.value >> 20)
    return out


def test_every_handle_tops_up_the_memory_pool(gpu, synth):
    """Buffer growth in the middle of a frame must be served by the pool, never by the driver (5 - 125 ms a time,
    profiles/r01_e2e_stalls.txt): every new handle leaves at least LSB_POOL_PREWARM_MB (4096) unused in the pool,
    also when earlier handles of the process are alive and hold memory."""
    import os
    want = int(os.environ.get("LSB_POOL_PREWARM_MB", "4096"))
    if want == 0:
        pytest.skip("pool reservation switched off")
    import torch
    torch.zeros(1, device="cuda")                    # libcudart of the torch wheel needs a context to answer
    import gc
    gc.collect()                                     # handles of earlier tests give their memory back now, not in between
    handles = []
    for i in range(3):
        _, used_before = _pool_mib()
        h = gpu(device=0)
        handles.append(h)
        reserved, used = _pool_mib()                 # the handle's own first buffers already come out of the head-room
        assert reserved - used_before >= want - 1, (i, reserved, used_before, used)
        seq = synth.Sequence(scene_seed=77 + i)
        for j in range(3):                           # the handle takes memory out of the pool
            pts, _, stamp = seq.frame(synth.sensor("hdl64"), j)
            h.AddFrame(pts, stamp, j + 1)
    r2, u2 = _pool_mib()
    assert u2 > 0 and r2 >= u2
    for h in handles:
        h.close()


@pytest.mark.parametrize("n,bits", [(1, 5), (31, 11), (4096, 12), (4097, 22), (5000, 23), (81233, 29), (100000, 32), (262144, 28)])
def test_sort_pairs32_is_a_stable_sort(gpu, n, bits):
    """sort32.cu (the map update's key sort): stable order of numpy's argsort(kind="stable") on the low `bits` bits,
    tiles / passes / partial last tile, keys with few distinct values (long runs) and the all-ones key of rejected points."""
    import ctypes as C
    g = gpu()
    rng = np.random.default_rng(n + bits)
    mask = (1 << bits) - 1
    keys = rng.integers(0, 1 << 32, size=n, dtype=np.uint64).astype(np.uint32)
    if n > 100:
        keys[rng.integers(0, n, size=n // 3)] = keys[0]                 # runs of equal keys
        keys[rng.integers(0, n, size=n // 50)] = 0xFFFFFFFF
        keys[: n // 4] &= 0xFF                                          # small values: empty high digits
    vals = np.arange(n, dtype=np.uint32)
    order = np.argsort(keys & np.uint32(mask), kind="stable")
    k, v = keys.copy(), vals.copy()
    rc = g.lib.lsb_sort_pairs32(g.h, k.ctypes.data_as(C.c_void_p), v.ctypes.data_as(C.c_void_p), C.c_size_t(n), C.c_int(bits))
    assert rc == 0, g.last_error()
    assert np.array_equal(v, vals[order])
    assert np.array_equal(k, keys[order])

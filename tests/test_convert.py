"""Input adaptors (reference ros_wrapping/lidar_conversions: VelodyneToLidarNode.cxx:50-111,
RobosenseToLidarNode.cxx:58-137, Utilities.h:58-108).  CPU: the oracle's sequential restatement against a
pure-Python transcription of the estimator.  GPU: the parallel device form (ring bucketing + first descent,
max-scan for the dual-return filter) bit-identical to the oracle, and a converted frame fed to AddFrames."""
import math

import numpy as np
import pytest

from lidarslam_b200 import capi


def _velodyne_from(pts, time_valid, rng=None, ring_perm=None):
    v = np.zeros(len(pts), dtype=capi.VELODYNE_DTYPE)
    for f in ("x", "y", "z", "intensity"):
        v[f] = pts[f]
    v["ring"] = pts["laser_id"] if ring_perm is None else np.asarray(ring_perm)[pts["laser_id"]]
    v["time"] = pts["time"].astype(np.float32) if time_valid else 0.0
    return v


def _py_velodyne_times(v, laser, rpm, first_packet):
    """Utilities.h:58-108 + VelodyneToLidarNode.cxx:103-104, point by point (atan2 in float like std::atan2(float, float))."""
    prev, init, out = {}, None, np.zeros(len(v))
    for i in range(len(v)):
        adv = (math.pi - float(np.float32(math.atan2(float(v["y"][i]), float(v["x"][i]))))) / (2 * math.pi)
        if not prev:
            init = adv
        fa = math.fmod(1.0 + math.fmod(adv - init, 1.0), 1.0)
        if fa < prev.get(int(laser[i]), 0.0):
            fa += 1.0
        prev[int(laser[i])] = fa
        out[i] = (fa if first_packet else fa - 1) / rpm * 60.0
    return out


def test_oracle_velodyne_against_python(oracle_lib, synth):
    o = oracle_lib.Oracle(nb_threads=1)
    pts, _, _ = synth.Sequence().frame(synth.sensor("tiny"), 2)
    perm = np.random.default_rng(0).permutation(16)
    inv = np.argsort(perm)
    # valid time field: copied; laser ids through the mapping; device id set
    v = _velodyne_from(pts, True, ring_perm=perm)
    got = o.ConvertVelodyne(v, laser_id_mapping=inv, device_id=3)
    assert np.array_equal(got["laser_id"], pts["laser_id"]) and (got["device_id"] == 3).all() and (got["w"] == 1).all()
    assert np.array_equal(got["time"], pts["time"].astype(np.float32).astype(np.float64))
    assert got["x"].tobytes() == pts["x"].tobytes() and got["intensity"].tobytes() == pts["intensity"].tobytes()
    # no time field: azimuth advancement, both stamping conventions
    v = _velodyne_from(pts, False)
    for first_packet in (False, True):
        got = o.ConvertVelodyne(v, rpm=600.0, timestamp_first_packet=first_packet)
        want = _py_velodyne_times(v, v["ring"], 600.0, first_packet)
        assert np.abs(got["time"] - want).max() < 2e-9        # float atan2: libm vs the oracle's 1-ulp atan2f
        lo, hi = (0.0, 0.1 + 1e-3) if first_packet else (-0.1, 1e-3)
        assert got["time"].min() >= lo - 1e-12 and got["time"].max() <= hi
    # the scan is clockwise and complete: time grows along each ring, ending near the stamp
    t = got["time"][v["ring"] == 5]
    assert (np.diff(t) > -1e-9).all()
    assert len(o.ConvertVelodyne(v[:0])) == 0


def _rs_cloud(synth, dual=True, seed=3):
    """Organised 16 x 512 cloud (ring-major) with NaN returns and duplicated (dual-return) points."""
    pts, _, _ = synth.Sequence().frame(synth.sensor("tiny"), 1)
    H, W = 16, 512
    grid = pts.reshape(W, H).T.copy()                            # synth emits column-major firing order
    r = np.zeros(H * W, dtype=capi.XYZI_DTYPE)
    for f in ("x", "y", "z", "intensity"):
        r[f] = grid[f].ravel()
    rng = np.random.default_rng(seed)
    bad = rng.random(H * W) < 0.05
    r["x"][bad] = np.nan
    r["z"][rng.random(H * W) < 0.01] = np.inf
    if dual:
        for i in rng.integers(1, H * W, 400):                    # second return identical to the first ...
            r["x"][i], r["y"][i], r["z"][i] = r["x"][i - 1], r["y"][i - 1], r["z"][i - 1]
        for i in rng.integers(3, H * W, 200):                    # ... also across a gap of invalid points and in runs
            r["x"][i - 1] = np.nan
            r["x"][i], r["y"][i], r["z"][i] = r["x"][i - 2], r["y"][i - 2], r["z"][i - 2]
    return r, W, H


def test_oracle_robosense_against_python(oracle_lib, synth):
    o = oracle_lib.Oracle(nb_threads=1)
    r, W, H = _rs_cloud(synth)
    got = o.ConvertRobosense(r, W, H, rpm=600.0)
    rs16 = [0, 1, 2, 3, 4, 5, 6, 7, 15, 14, 13, 12, 11, 10, 9, 8]
    want = []
    for i in range(len(r)):
        p = r[i]
        if not (np.isfinite(p["x"]) and np.isfinite(p["y"]) and np.isfinite(p["z"])):
            continue
        if want and (p["x"], p["y"], p["z"]) == want[-1][:3]:
            continue
        want.append((p["x"], p["y"], p["z"], p["intensity"], rs16[i // W], ((i % W) / W - 1) / 600.0 * 60.0))
    assert len(got) == len(want) < len(r)
    for k, f in enumerate(("x", "y", "z", "intensity", "laser_id", "time")):
        assert np.array_equal(got[f], np.array([w[k] for w in want]).astype(got[f].dtype)), f
    # other ring counts: identity mapping unless a table is given
    g2 = o.ConvertRobosense(r[:8 * W], W, 8)
    assert set(np.unique(g2["laser_id"])) == set(range(8))
    g3 = o.ConvertRobosense(r, W, H, laser_id_mapping=list(range(15, -1, -1)), device_id=1)
    assert g3["laser_id"][0] == 15 and (g3["device_id"] == 1).all()


@pytest.mark.gpu
def test_gpu_adaptors_match_oracle(gpu, oracle_lib, synth):
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    rng = np.random.default_rng(11)
    for name, nframe in (("tiny", 2), ("vlp16", 1), ("os128", 0)):
        pts, _, stamp = synth.Sequence().frame(synth.sensor(name), nframe)
        for time_valid, first_packet, mapping in ((True, False, None), (False, False, None), (False, True, "perm")):
            perm = rng.permutation(int(pts["laser_id"].max()) + 1) if mapping else None
            v = _velodyne_from(pts, time_valid, ring_perm=perm)
            kw = dict(laser_id_mapping=None if perm is None else np.argsort(perm), device_id=2, rpm=600.0, timestamp_first_packet=first_packet)
            a, b = g.ConvertVelodyne(v, **kw), o.ConvertVelodyne(v, **kw)
            assert len(a) == len(b) == len(pts) and a.tobytes() == b.tobytes(), (name, time_valid, first_packet)
    # a scan that spins past 360 degrees and starts mid-ring order: shuffled ring interleaving, repeated first azimuth
    pts, _, _ = synth.Sequence().frame(synth.sensor("vlp16"), 3)
    over = np.concatenate([pts, pts[:4000]])
    v = _velodyne_from(over, False)
    a, b = g.ConvertVelodyne(v), o.ConvertVelodyne(v)
    assert a.tobytes() == b.tobytes() and a["time"].max() > 0          # advancement beyond 1
    # Robosense: NaN / inf returns, dual returns in runs and across gaps
    for dual in (True, False):
        r, W, H = _rs_cloud(synth, dual=dual)
        a, b = g.ConvertRobosense(r, W, H), o.ConvertRobosense(r, W, H)
        assert len(a) == len(b) and a.tobytes() == b.tobytes()
    allnan = np.zeros(64, dtype=capi.XYZI_DTYPE)
    allnan["x"] = np.nan
    assert len(g.ConvertRobosense(allnan, 16, 4)) == 0 == len(o.ConvertRobosense(allnan, 16, 4))
    assert len(g.ConvertVelodyne(v[:0])) == 0


@pytest.mark.gpu
def test_gpu_converted_frames_drive_slam(gpu, oracle_lib, synth):
    """Raw driver clouds -> device conversion -> AddFrames without a host round trip == the oracle on the same clouds."""
    g, o = gpu(), oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    seq, sens = synth.Sequence(scene_seed=1234), synth.sensor("vlp16")
    for i in range(5):
        pts, _, stamp = seq.frame(sens, i)
        v = _velodyne_from(pts, False)                                   # times rebuilt from the azimuth
        for s in (g, o):
            s.ConvertVelodyne(v, rpm=600.0)
        assert g.AddConvertedFrame(stamp, i + 1) == o.AddConvertedFrame(stamp, i + 1)
        Tg, _ = g.GetWorldTransform()
        To, _ = o.GetWorldTransform()
        assert np.abs(Tg - To).max() <= 1e-6
        for k in (0, 1):
            assert g.GetKeypoints(k).tobytes() == o.GetKeypoints(k).tobytes()
    assert np.linalg.norm(Tg[:3, 3]) > 0.3                               # the sequence did move

"""Pipelined use of the C ABI (deferred map update, lsb_prefetch_frames): same results, bit for bit, as the fully
synchronous call the other GPU tests compare with the oracle - poses, keypoints, match statuses, maps, statistics."""
import numpy as np
import pytest

from lidarslam_b200 import capi

pytestmark = pytest.mark.gpu


def _blobs(s, ego=1):
    p = s.GetParams()
    p.use_keypoints[2] = 1
    p.ego_motion = ego
    p.overlap_sampling_ratio = 0.25
    s.SetParams(p)


def _frames(synth, sensor, n, extra=None):
    seq = synth.Sequence(scene_seed=1234)
    sens = [synth.sensor(sensor, device_id=0)]
    if extra:
        sens.append(synth.sensor(extra, device_id=1, mount_xyz=(0.6, 0, -0.35), mount_rpy=(0, np.deg2rad(20), np.pi / 2)))
    out = []
    for i in range(n):
        fs, st = [], []
        for s in sens:
            pts, _, stamp = seq.frame(s, i)
            fs.append(pts)
            st.append(stamp)
        out.append((fs, st))
    return out


def _snapshot(h, with_maps):
    d = {"T": h.GetWorldTransform()[0].copy(), "rel": h.GetRelativeTransform().copy(), "overlap": np.float32(h.GetOverlapEstimation()).tobytes()}
    for k in capi.KEYPOINT_TYPES:
        d["kp%d" % k] = h.GetKeypoints(k).tobytes()
        d["st%d" % k] = h.GetMatchDebug(1, k)[0].tobytes()
        if with_maps:
            d["map%d" % k] = h.GetMap(k).tobytes()
    st = h.GetFrameStats()
    d["stats"] = (list(st.n_keypoints), list(st.n_submap), list(st.n_matches), list(st.n_map), st.is_keyframe, st.loc_lm_iters)
    return d, st


def _same(a, b, what):
    for key in a:
        if isinstance(a[key], np.ndarray):
            assert np.array_equal(a[key], b[key]), "%s: %s differs" % (what, key)
        else:
            assert a[key] == b[key], "%s: %s differs" % (what, key)


@pytest.mark.parametrize("sensor,ego,n", [("vlp16", 1, 14), ("tiny", 3, 12), ("os128", 1, 6)])
def test_deferred_map_update_and_prefetch_match_the_synchronous_call(gpu, synth, sensor, ego, n):
    frames = _frames(synth, sensor, n)
    ref = gpu()
    ref.SetPipelining(False)
    _blobs(ref, ego)
    want = []
    for i, (fs, st) in enumerate(frames):
        ref.AddFrames(fs, st, [i + 1])
        snap, stats = _snapshot(ref, True)
        assert stats.map_update_deferred == 0 and stats.prefetched == 0
        want.append(snap)

    # deferred map update, inspected every frame (the accessors join the update)
    d = gpu()
    _blobs(d, ego)
    deferred = 0
    for i, (fs, st) in enumerate(frames):
        d.AddFrames(fs, st, [i + 1])
        T = d.GetWorldTransform()[0]
        assert np.array_equal(T, want[i]["T"]), "frame %d: pose differs before the join" % i
        snap, stats = _snapshot(d, True)
        deferred += stats.map_update_deferred
        if stats.is_keyframe and stats.map_update_deferred:
            assert stats.ms_map_update > 0.0
        _same(want[i], snap, "deferred, frame %d" % i)
    assert deferred > 0

    # prefetch one frame ahead + deferred update, maps only looked at in the end (nothing joins in between but the path itself)
    p = gpu()
    _blobs(p, ego)
    p.PrefetchFrames(*frames[0], [1])
    hits = 0
    for i, (fs, st) in enumerate(frames):
        if i + 1 < n:
            p.PrefetchFrames(*frames[i + 1], [i + 2])
        p.AddFrames(fs, st, [i + 1])
        assert np.array_equal(p.GetWorldTransform()[0], want[i]["T"]), "prefetch, frame %d: pose differs" % i
        if i in (n // 2, n - 1):
            snap, stats = _snapshot(p, True)
            hits += stats.prefetched
            _same(want[i], snap, "prefetch, frame %d" % i)
    assert hits == 2
    az = [int(h.lib.lsb_get_azimuth_estimations(h.h, 0)) for h in (p, ref)]
    assert az == [1, 1], az
    assert p.GetExtractorParams().azimuthal_resolution == ref.GetExtractorParams().azimuthal_resolution
    # extractor debug arrays belong to the frame that was added last, not to a staged one
    assert np.array_equal(p.GetKeDebug(0, 0), ref.GetKeDebug(0, 0))


def test_prefetch_from_device_buffers_dual_sensor_and_mismatch(gpu, synth):
    n = 8
    frames = _frames(synth, "vlp16", n, extra="tiny")
    c, s_ = np.cos(np.pi / 2), np.sin(np.pi / 2)
    cy, sy = np.cos(np.deg2rad(20)), np.sin(np.deg2rad(20))
    T = np.eye(4)
    T[:3, :3] = np.array([[c, -s_, 0], [s_, c, 0], [0, 0, 1.]]) @ np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    T[:3, 3] = [0.6, 0, -0.35]

    def cfg(s):
        _blobs(s)
        s.SetExtractorParams(s.GetExtractorParams(0), 1)
        s.SetBaseToLidarOffset(T, 1)
    ref = gpu()
    ref.SetPipelining(False)
    cfg(ref)
    for i, (fs, st) in enumerate(frames):
        ref.AddFrames(fs, st, [i + 1, i + 1])
    want, _ = _snapshot(ref, True)

    p = gpu()
    cfg(p)
    dev = [[p.DevUpload(f) for f in fs] for fs, _ in frames]
    sizes = [[len(f) for f in fs] for fs, _ in frames]
    p.PrefetchFramesDev(dev[0], sizes[0], frames[0][1], [1, 1])
    for i, (fs, st) in enumerate(frames):
        if i + 1 < n and i not in (3, 5):
            p.PrefetchFramesDev(dev[i + 1], sizes[i + 1], frames[i + 1][1], [i + 2, i + 2])
        if i == 2:
            # frames 2 and 3 are staged; a third one (two calls ahead: uploaded now, extracted when frame 3's call begins) is
            # accepted, a fourth is refused, and what is staged survives the refusal
            p.PrefetchFramesDev(dev[i + 2], sizes[i + 2], frames[i + 2][1], [i + 3, i + 3])
            with pytest.raises(RuntimeError):
                p.PrefetchFramesDev(dev[i + 3], sizes[i + 3], frames[i + 3][1], [i + 4, i + 4])
        # frame 6 is never staged: adding it drops the staged frame 7, which then arrives unstaged as well
        p.AddFramesDev(dev[i], sizes[i], st, [i + 1, i + 1])
    got, _ = _snapshot(p, True)
    _same(want, got, "device prefetch")
    assert np.array_equal(p.GetTrajectory()[1], ref.GetTrajectory()[1])


def test_setters_drop_staged_inputs_and_whole_loop_timer(gpu, synth):
    n = 6
    frames = _frames(synth, "vlp16", n)
    ref = gpu()
    ref.SetPipelining(False)
    for i, (fs, st) in enumerate(frames):
        if i == 3:
            ke = ref.GetExtractorParams()
            ke.neighbor_width = 5
            ref.SetExtractorParams(ke)
        ref.AddFrames(fs, st, [i + 1])
    want, _ = _snapshot(ref, True)

    p = gpu()
    p.TimerBegin()
    p.PrefetchFrames(*frames[0], [1])
    for i, (fs, st) in enumerate(frames):
        if i == 3:
            ke = p.GetExtractorParams()
            ke.neighbor_width = 5
            p.SetExtractorParams(ke)          # frame 3 was staged with the old width: the staged keypoints must not be used
        if i + 1 < n:
            p.PrefetchFrames(*frames[i + 1], [i + 2])
        p.AddFrames(fs, st, [i + 1])
    ms = p.TimerEnd()
    assert ms > 0.0
    got, _ = _snapshot(p, True)
    _same(want, got, "setter between prefetch and add")
    p.Synchronize()
    p.close()


_MODE_SCRIPT = r"""
import hashlib, sys
sys.path.insert(0, %r)
from lidarslam_b200.slam import Slam
from lidarslam_b200 import capi
from tools.synth import synth
seq, sens = synth.Sequence(scene_seed=1234), synth.sensor("vlp16")
frames = [seq.frame(sens, i) for i in range(10)]
h = Slam(device=0)
p = h.GetParams(); p.use_keypoints[2] = 1; h.SetParams(p)
m = hashlib.sha256()
h.PrefetchFrames([frames[0][0]], [frames[0][2]], [1])
for i, (pts, _, stamp) in enumerate(frames):
    if i + 1 < len(frames):
        h.PrefetchFrames([frames[i + 1][0]], [frames[i + 1][2]], [i + 2])
    h.AddFrames([pts], [stamp], [i + 1])
    m.update(h.GetWorldTransform()[0].tobytes())
for k in capi.KEYPOINT_TYPES:
    m.update(h.GetMap(k).tobytes()); m.update(h.GetKeypoints(k).tobytes())
print("HASH", m.hexdigest())
"""


def test_host_thread_policies_give_the_same_results(gpu):
    """LSB_HOST_MODE 0 / 1 / 2 (helpers spinning / sleeping / everything on the caller's thread, nothing deferred) and
    LSB_HOST_LANES=0: the host-side scheduling changes, the results do not (each policy in its own process: read at start-up)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    got = {}
    for name, env in (("mode0", {"LSB_HOST_MODE": "0"}), ("mode1", {"LSB_HOST_MODE": "1"}), ("mode2", {"LSB_HOST_MODE": "2"}),
                      ("lanes0", {"LSB_HOST_LANES": "0"}), ("sync", {"LSB_DEFER_MAP": "0"}), ("cubsort", {"LSB_MAP_OWN_SORT": "0"})):
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, "-c", _MODE_SCRIPT % root], env=e, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, name + ": " + r.stdout + r.stderr
        got[name] = [l for l in r.stdout.splitlines() if l.startswith("HASH")][0]
    assert len(set(got.values())) == 1, got

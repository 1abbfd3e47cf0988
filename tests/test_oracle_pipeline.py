"""Oracle as a working SLAM (CPU): the ICP solve recovers a known perturbation, the pipeline
tracks the synthetic ground-truth trajectory, soft-fail paths behave like the reference."""
import numpy as np

from lidarslam_b200 import capi


def _rel(T0, T1):
    return np.linalg.inv(T0) @ T1


def test_icp_solve_recovers_known_transform(oracle_lib, synth):
    """Keypoints of one frame matched against themselves moved by a known rigid transform:
    KeypointsMatcher + LocalOptimizer (KeypointsMatcher.cxx:33-74, LocalOptimizer.cxx:74-102)
    must pull the perturbed prior back to that transform."""
    o = oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    seq = synth.Sequence()
    pts, _, _ = seq.frame(synth.sensor("vlp16"), 3)
    _, _, kp = o.ExtractKeypoints(pts)
    kp[2] = kp[2][:0]                                   # edges + planes only
    c, s = np.cos(0.01), np.sin(0.01)
    T = np.array([[c, -s, 0, 0.30], [s, c, 0, -0.10], [0, 0, 1, 0.05], [0, 0, 0, 1.]])
    tgt = []
    for k in range(3):
        w = kp[k].copy()
        xyz = np.stack([w["x"], w["y"], w["z"]], 1).astype(np.float64) @ T[:3, :3].T + T[:3, 3]
        w["x"], w["y"], w["z"] = xyz[:, 0], xyz[:, 1], xyz[:, 2]
        tgt.append(w)
    prior = T.copy()
    prior[:3, 3] += [0.08, -0.05, 0.03]
    pose = prior
    for it in range(4):                                 # ICP iterations (Slam.cxx:1071)
        mp = capi.MatchParams(0, 5.0, 10, 4, 0.2, 5, 0.04, 0.2, 10, 2.0 - 0.5 * it)
        r = o.MatchAndSolve(kp, tgt, mp, 15, pose)
        assert r["rc"] == 0 and (r["status"] == 0).sum() > 1000
        pose = r["pose"]
    assert np.linalg.norm(pose[:3, 3] - T[:3, 3]) < 5e-3
    assert np.abs(pose[:3, :3] - T[:3, :3]).max() < 1e-3
    cov = np.array(r["err"].covariance).reshape(6, 6)
    assert np.allclose(cov, cov.T, atol=1e-12) and (np.linalg.eigvalsh(cov) > -1e-12).all()


def test_sequence_tracks_ground_truth(oracle_lib, synth):
    o = oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    seq = synth.Sequence()
    sens = synth.sensor("vlp16")
    T0 = None
    errs = []
    for i in range(25):
        pts, Tgt, stamp = seq.frame(sens, i)
        rc = o.AddFrame(pts, stamp, i + 1)
        if i == 0:
            assert rc == 2                                # LSB_W_LOC_SKIPPED: empty maps (Slam.cxx:1098-1107)
            T0 = Tgt
        else:
            assert rc == 0
        T, t = o.GetWorldTransform()
        assert abs(t - stamp * 1e-6) < 1e-9
        errs.append(np.linalg.norm(T[:3, 3] - _rel(T0, Tgt)[:3, 3]))
    assert errs[0] < 1e-12
    assert max(errs) < 0.10, errs                         # metres, over 25 frames at 8 m/s with 1 cm range noise
    st = o.GetFrameStats()
    assert st.n_matches[0] > 100 and st.n_matches[1] > 1000 and st.loc_icp_iters == 3
    assert o.GetTotalMatchedKeypoints() == st.n_matches[0] + st.n_matches[1]


def test_soft_failures(oracle_lib, synth):
    o = oracle_lib.Oracle(nb_threads=2)
    seq = synth.Sequence()
    pts, _, stamp = seq.frame(synth.sensor("tiny"), 0)
    assert o.AddFrame(pts, stamp, 1) == 2
    assert o.AddFrame(pts, stamp, 2) == 1                 # duplicate stamp: frame ignored (Slam.cxx:727-731)
    assert o.AddFrame(pts[:0], stamp + 100000, 3) == 1    # empty frame (Slam.cxx:720-724)
    pts2, _, stamp2 = seq.frame(synth.sensor("tiny"), 1)
    assert o.AddFrame(pts2, stamp2, 7) & 32               # dropped frames warning (Slam.cxx:734-740)
    assert len(o.GetKeypoints(2)) == 0                    # blobs dropped unless UseBlobs (Slam.cxx:790-800)
    o.Reset()
    assert len(o.GetMap(0)) == 0 and len(o.GetMap(1)) == 0


def test_overlap_estimator_against_numpy(oracle_lib, synth):
    """LCP estimator (ConfidenceEstimators.cxx:27-65) of the oracle == a numpy/cKDTree restatement."""
    from scipy.spatial import cKDTree
    o = oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    p = o.GetParams()
    p.overlap_sampling_ratio = 0.33
    p.use_keypoints[2] = 1
    o.SetParams(p)
    seq, sens = synth.Sequence(), synth.sensor("vlp16")
    for i in range(3):
        pts, _, stamp = seq.frame(sens, i)
        o.AddFrame(pts, stamp, i + 1)
        ov = o.GetOverlapEstimation()
        if i == 0:
            assert ov == -1.0                    # no sub-map yet
            continue
        reg = o.GetRegisteredFrame()
        n = int(np.float32(len(reg)) * np.float32(0.33))
        idx = (np.arange(n, dtype=np.float32) / np.float32(0.33)).astype(np.int64)
        q = np.stack([reg["x"], reg["y"], reg["z"]], 1)[idx].astype(np.float64)
        best = np.zeros(n, np.float32)
        for k, leaf in ((0, 0.3), (1, 0.6), (2, 0.3)):
            sub = o.GetTargetSubMap(k)
            if len(sub) == 0:
                continue
            d, _ = cKDTree(np.stack([sub["x"], sub["y"], sub["z"]], 1).astype(np.float64)).query(q)
            thr = np.float32((leaf / 3.0) ** 2)
            best = np.maximum(best, np.exp(-(d.astype(np.float32) ** 2) / (2 * thr)).astype(np.float32))
        assert abs(ov - best.mean(dtype=np.float64)) < 2e-6
        assert 0.3 < ov < 1.0


def test_sensor_constraints(oracle_lib, synth):
    """SURVEY 8f-3 (SensorConstraints.cxx, Slam.cxx:256-261, 346-352, 1123-1131): wheel-odometer and IMU gravity blocks
    in the localization solve; time mismatch warning; gravity reference = mean of the fullest direction bin."""
    import _sensors
    sens = synth.sensor("vlp16")
    n = 8
    free, cons = (oracle_lib.Oracle(nb_threads=oracle_lib.max_threads()) for _ in range(2))
    cons.SetWheelOdomWeight(1e4)
    cons.SetGravityWeight(1e4)
    wheel, grav = _sensors.feed([cons], synth.Sequence(), sens, n, noise=1e-3)
    seq = synth.Sequence()
    T0 = None
    for i in range(n):
        pts, Tgt, stamp = seq.frame(sens, i)
        rf, rc = free.AddFrame(pts, stamp, i + 1), cons.AddFrame(pts, stamp, i + 1)
        assert rf == rc, (i, rf, rc)
        T0 = Tgt if i == 0 else T0
        Tf, _ = free.GetWorldTransform()
        Tc, _ = cons.GetWorldTransform()
        gt = _rel(T0, Tgt)
        if i >= 2:
            # the heavily weighted odometer block pins |t| to the measured chord (1 mm noise), the free run only to ICP accuracy
            assert abs(np.linalg.norm(Tc[:3, 3]) - (wheel[i + 1][1] - wheel[1][1])) < 5e-3
            assert np.abs(Tc - Tf).max() > 1e-7               # the blocks did enter the solve
            assert np.linalg.norm(Tc[:3, 3] - gt[:3, 3]) < 0.10
    g = cons.GetGravityReference()
    dirs = np.array([a / np.linalg.norm(a) for _, a in grav])
    assert abs(np.linalg.norm(g) - 1) < 1e-12 and np.abs(g - dirs.mean(0) / np.linalg.norm(dirs.mean(0))).max() < 2e-2
    assert (free.GetGravityReference() == 0).all()
    # stamps outside the measurements: warning, no constraint, the frame is processed as without sensors
    pts, _, stamp = seq.frame(sens, n)
    assert cons.AddFrame(pts, stamp, n + 1) & 256             # LSB_W_SENSOR_TIME
    # ClearSensorMeasurements: managers empty again
    cons.ClearSensorMeasurements()
    pts, _, stamp = seq.frame(sens, n + 1)
    assert cons.AddFrame(pts, stamp, n + 2) == 0

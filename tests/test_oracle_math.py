"""Oracle building blocks against independent numpy / scipy computations (CPU)."""
import numpy as np
import pytest


def _ulp_err(a, b):
    return np.abs(a - b) / np.maximum(np.spacing(np.abs(b)), 1e-300)


def test_detmath_accuracy(oracle_lib):
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-50, 50, 20000), rng.uniform(-1e-3, 1e-3, 2000), [0.0, np.pi / 4, -np.pi / 2, 1e4]])
    assert _ulp_err(oracle_lib.detmath(0, x), np.sin(x)).max() <= 2
    assert _ulp_err(oracle_lib.detmath(1, x), np.cos(x)).max() <= 2
    y, z = rng.normal(size=20000), rng.normal(size=20000)
    assert _ulp_err(oracle_lib.detmath(2, y, z), np.arctan2(y, z)).max() <= 2
    u = rng.uniform(-1, 1, 20000)
    # acos/asin via atan2(sqrt((1-x)(1+x)), x): a few ulp, absolute error tiny
    assert np.abs(oracle_lib.detmath(3, u) - np.arccos(u)).max() < 1e-15 * 8
    assert np.abs(oracle_lib.detmath(4, u) - np.arcsin(u)).max() < 1e-15 * 8
    assert oracle_lib.detmath(2, np.array([0.0]), np.array([-1.0]))[0] == np.pi


@pytest.mark.parametrize("as_float", [False, True])
def test_eigen33_against_numpy(oracle_lib, as_float):
    rng = np.random.default_rng(1)
    mats = []
    for _ in range(3000):
        p = rng.normal(size=(rng.integers(4, 12), 3)) * rng.uniform(0.01, 3, 3)
        mats.append(np.cov(p.T, bias=True))
    mats = np.array(mats)
    ev, vec = oracle_lib.eigen33(mats, as_float=as_float)
    ref = np.linalg.eigvalsh(mats)
    tol = 2e-4 if as_float else 1e-9
    scale = np.abs(ref).max(axis=1, keepdims=True)
    assert (np.abs(ev - ref) <= tol * scale).all()
    # eigenvectors: A v = lambda v for the extreme eigenvalues (what the callers use)
    for k in (0, 2):
        v = vec[:, :, k]
        res = np.einsum("nij,nj->ni", mats, v) - ev[:, k:k + 1] * v
        gap_ok = (ref[:, 2] - ref[:, 0]) > 1e-3 * scale[:, 0]
        assert (np.linalg.norm(res, axis=1)[gap_ok] <= (3e-3 if as_float else 1e-7) * scale[gap_ok, 0]).all()
        assert np.allclose(np.linalg.norm(v, axis=1), 1.0, atol=1e-4 if as_float else 1e-12)


def test_knn_against_ckdtree(oracle_lib, synth):
    from scipy.spatial import cKDTree
    seq = synth.Sequence()
    pts, _, _ = seq.frame(synth.sensor("vlp16"), 0)
    rng = np.random.default_rng(3)
    cloud = pts[rng.permutation(len(pts))[:5000]].copy()
    xyz = np.stack([cloud["x"], cloud["y"], cloud["z"]], 1)
    q = (xyz[:2000] + rng.normal(0, 0.5, (2000, 3))).astype(np.float32)
    o = oracle_lib.Oracle(nb_threads=4)
    idx, d2 = o.KnnSearch(cloud, q, 8)
    dd, ii = cKDTree(xyz.astype(np.float64)).query(q.astype(np.float64), k=8)
    # same neighbour sets wherever the 8th/9th distances are not within float rounding of each other
    d9 = cKDTree(xyz.astype(np.float64)).query(q.astype(np.float64), k=9)[0][:, 8]
    clear = (d9 - dd[:, 7]) > 1e-4
    assert clear.mean() > 0.95
    assert (np.sort(idx[clear], 1) == np.sort(ii[clear], 1)).all()
    assert np.allclose(d2[clear], dd[clear] ** 2, rtol=1e-4, atol=1e-6)
    assert (np.diff(d2, axis=1) >= 0).all()


def test_undistort_against_scipy(oracle_lib, synth):
    from scipy.spatial.transform import Rotation, Slerp
    seq = synth.Sequence()
    pts, _, _ = seq.frame(synth.sensor("tiny"), 0)
    R1 = Rotation.from_euler("zyx", [0.03, -0.01, 0.02])
    H0, H1 = np.eye(4), np.eye(4)
    H1[:3, :3] = R1.as_matrix()
    H1[:3, 3] = [0.8, 0.05, -0.02]
    o = oracle_lib.Oracle()
    out = o.Undistort(pts, H0, H1, -0.1, 0.0)
    t = (pts["time"] + 0.1) / 0.1
    rot = Slerp([0, 1], Rotation.concatenate([Rotation.identity(), R1]))(np.clip(t, 0, 1))
    xyz = np.stack([pts["x"], pts["y"], pts["z"]], 1).astype(np.float64)
    ref = rot.apply(xyz) + t[:, None] * H1[:3, 3]
    got = np.stack([out["x"], out["y"], out["z"]], 1)
    assert np.abs(got - ref).max() < 2e-5   # float32 store


def test_rolling_grid_against_dict_model(oracle_lib, synth):
    """RollingGrid::Add (MAX_INTENSITY, RollingGrid.cxx:160-318) against a dictionary model in numpy."""
    seq = synth.Sequence()
    pts, T, _ = seq.frame(synth.sensor("tiny"), 0)
    o = oracle_lib.Oracle()
    sub = pts[::3].copy()
    o.MapAdd(0, sub, False, 1.5)
    got = o.GetMap(0)
    # model: grid 50, resolution 10, leaf 0.3; origin after Roll keeps the cloud inside: recompute keys from the stored points
    leaf, res = np.float32(0.3), np.float32(10.0)
    xyz = np.stack([sub["x"], sub["y"], sub["z"]], 1)
    # the grid position is unknown here; use translation invariance of the leaf partition inside an outer voxel:
    # two points share a leaf iff they share (outer voxel, inner leaf).  Recover the origin from the oracle's own rule.
    mn, mx = xyz.min(0), xyz.max(0)
    half = np.float32(25 * 10.0)
    pos = np.zeros(3, np.float32)
    down, up = mn - (pos - half), mx - (pos + half)
    off = np.clip((up + down) / 2, np.minimum(down, 0), np.maximum(up, 0))
    pos = pos + np.round(off / res).astype(np.float32) * res
    origin = pos - half

    def rnd(v):   # round half away from zero
        return np.where(v >= 0, np.floor(v + 0.5), np.ceil(v - 0.5)).astype(np.int64)
    vout = rnd((xyz - origin) / res)
    inside = ((vout >= 0) & (vout < 50)).all(1)
    centre = vout.astype(np.float32) * res + origin
    vin = rnd((xyz - centre) / leaf)
    key_out = vout[:, 2] * 2500 + vout[:, 1] * 50 + vout[:, 0]
    key_in = vin[:, 2] * 2500 + vin[:, 1] * 50 + vin[:, 0]
    best = {}
    for i in np.flatnonzero(inside):
        k = (int(key_out[i]), int(key_in[i]))
        if k not in best or sub["intensity"][i] > sub["intensity"][best[k]]:
            best[k] = i
    order = sorted(best)
    exp = sub[[best[k] for k in order]].copy()
    exp["time"] = 1.5
    exp["label"] = 0
    assert len(got) == len(exp)
    assert got.tobytes() == exp.tobytes()


# ---------------------------------------------------------------- LM (common/lm_core.h)
def _euler_zyx(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def _lm_problem(rng, n, x_true, noise=0.0):
    """Records of the reference's three residual kinds (KeypointsMatcher.cxx:270-480): A = I - n n^T (line),
    n n^T (plane), a full SPD root (blob); X is the keypoint, P the model point."""
    R, t = _euler_zyx(*x_true[3:]), x_true[:3]
    nrm = rng.normal(size=(n, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    kind = rng.integers(0, 3, n)
    A = np.zeros((n, 3, 3))
    for i in range(n):
        nn = np.outer(nrm[i], nrm[i])
        if kind[i] == 0:
            A[i] = np.eye(3) - nn
        elif kind[i] == 1:
            A[i] = nn
        else:
            B = rng.normal(size=(3, 3))
            A[i] = B @ B.T / 3 + 0.2 * np.eye(3)
    P = rng.uniform(-30, 30, (n, 3))
    X = (P + noise * rng.normal(size=(n, 3)) - t) @ R      # R^T (P - t), row form
    w = rng.uniform(0.3, 1.0, n)
    return A, P, X, w


def _tukey(s, a2):
    v = np.clip(1 - s / a2, 0, None)
    return np.where(s <= a2, a2 / 3 * (1 - v ** 3), a2 / 3), np.where(s <= a2, v ** 2, 0.0)


def test_lm_terms_against_numpy(oracle_lib):
    """J^T J, J^T r and the cost of lm_eval_record == the reference residual r = A (R X + t - P) under
    ScaledLoss(TukeyLoss) with the Ceres corrector, Jacobian by central differences."""
    rng = np.random.default_rng(3)
    x = np.array([0.4, -0.2, 0.1, 0.03, -0.05, 0.7])
    A, P, X, w = _lm_problem(rng, 200, x + [0.05, 0.02, -0.04, 0.01, -0.01, 0.02], noise=0.02)
    w[::11] = -1.0                                          # failed matches contribute zero
    sat = 0.3
    got = oracle_lib.lm_terms(A, P, X, w, x, sat)

    def resid(xx):
        return np.einsum("nij,nj->ni", A, X @ _euler_zyx(*xx[3:]).T + xx[:3] - P)

    r = resid(x)
    J = np.zeros((len(A), 3, 6))
    for k in range(6):
        h = np.zeros(6)
        h[k] = 1e-6
        J[:, :, k] = (resid(x + h) - resid(x - h)) / 2e-6
    rho0, rho1 = _tukey((r * r).sum(1), sat * sat)
    ok = w >= 0
    scale = np.where(ok, w * rho1, 0.0)
    H = np.einsum("n,nki,nkj->nij", scale, J, J)
    g = np.einsum("n,nki,nk->ni", scale, J, r)
    iu = np.triu_indices(6)
    assert np.allclose(got[:, :21], H[:, iu[0], iu[1]], rtol=1e-6, atol=1e-9)
    assert np.allclose(got[:, 21:27], g, rtol=1e-6, atol=1e-9)
    assert np.allclose(got[:, 27], np.where(ok, 0.5 * w * rho0, 0.0), rtol=1e-12, atol=0)
    assert (got[~ok] == 0).all()


def test_lm_reduction_tree_and_thread_invariance(oracle_lib):
    """The cost after the first evaluation is the radix-32 tree sum of lm_core.h, bit for bit, and does not
    depend on the number of threads."""
    rng = np.random.default_rng(4)
    x = np.array([0.1, 0.0, -0.1, 0.0, 0.01, -0.02])
    n = 70001                                               # 69 chunks -> 3 blocks, ragged tail
    A, P, X, w = _lm_problem(rng, n, x, noise=0.05)
    terms = oracle_lib.lm_terms(A, P, X, w, x, 1.0)[:, 27]

    def fold32(v):                                          # v: (..., 32)
        for off in (16, 8, 4, 2, 1):
            v = v[..., :off] + v[..., off:2 * off]
        return v[..., 0]

    pad = np.zeros(3 * 32 * 1024)
    pad[:n] = terms
    blocks = fold32(fold32(fold32(pad.reshape(3, 32, 32, 32))))
    expect = (blocks[0] + blocks[1]) + blocks[2]
    r1 = oracle_lib.lm_solve(A, P, X, w, x, 1.0, max_iter=0, threads=1)
    r8 = oracle_lib.lm_solve(A, P, X, w, x, 1.0, max_iter=0, threads=8)
    assert r1["cost"] == expect == r8["cost"]
    s1 = oracle_lib.lm_solve(A, P, X, w, x + 0.01, 1.0, max_iter=15, threads=1)
    s8 = oracle_lib.lm_solve(A, P, X, w, x + 0.01, 1.0, max_iter=15, threads=8)
    assert np.array_equal(s1["x"], s8["x"]) and s1["iterations"] == s8["iterations"]


def test_lm_solve_against_scipy(oracle_lib):
    """Trust-region LM of lm_core.h reaches the least-squares solution scipy finds (near-quadratic loss)."""
    from scipy.optimize import least_squares
    rng = np.random.default_rng(5)
    x_true = np.array([0.3, -0.4, 0.2, 0.02, -0.03, 0.05])
    A, P, X, w = _lm_problem(rng, 3000, x_true, noise=0.01)

    def resid(xx):
        d = X @ _euler_zyx(*xx[3:]).T + xx[:3] - P
        return (np.sqrt(w)[:, None] * np.einsum("nij,nj->ni", A, d)).ravel()

    ref = least_squares(resid, np.zeros(6), method="lm", xtol=1e-14, ftol=1e-14, gtol=1e-14)
    got = oracle_lib.lm_solve(A, P, X, w, np.zeros(6), 1e3, max_iter=50)
    assert got["term"] in (1, 2, 3)                          # converged (gradient / parameter / function tolerance)
    assert np.abs(got["x"] - ref.x).max() < 2e-6
    assert np.abs(got["x"] - x_true).max() < 5e-3
    # held parameters (2-D mode): z, rx, ry stay put (LocalOptimizer.cxx:89-90)
    g2 = oracle_lib.lm_solve(A, P, X, w, np.zeros(6), 1e3, max_iter=50, two_d=True)
    assert (g2["x"][[2, 3, 4]] == 0).all() and np.abs(g2["x"][[0, 1, 5]] - x_true[[0, 1, 5]]).max() < 0.05


def test_detmath_exp(oracle_lib):
    """dm::exp (overlap estimator): <= 1 ulp in double, and the float variant equals the correctly rounded exp."""
    rng = np.random.default_rng(1)
    x = np.concatenate([np.linspace(-700, 5, 50001), -rng.uniform(0, 120, 50000)])
    assert _ulp_err(oracle_lib.detmath(5, x), np.exp(x)).max() <= 2
    xs = np.linspace(-112, 0, 50001).astype(np.float32)
    got = oracle_lib.detmath(6, xs.astype(np.float64)).astype(np.float32)
    assert np.array_equal(got, np.exp(xs.astype(np.float64)).astype(np.float32))
    assert oracle_lib.detmath(5, np.array([0.0]))[0] == 1.0 and oracle_lib.detmath(5, np.array([-800.0]))[0] == 0.0


def test_lm_sensor_terms_against_scipy(oracle_lib):
    """the wheel-odometer block r = |t - prev| - dist (CeresCostFunctions.h:267-296) and the IMU gravity
    block r = R g_cur - g_ref (:311-341), both under ScaledLoss(NULL, weight), move the solution exactly as stacking
    sqrt(weight) * r under the keypoint residuals does in scipy."""
    from scipy.optimize import least_squares
    rng = np.random.default_rng(8)
    x_true = np.array([0.3, -0.4, 0.2, 0.02, -0.03, 0.05])
    A, P, X, w = _lm_problem(rng, 400, x_true, noise=0.02)
    prev, dist, ww = np.array([-1.0, 0.5, 0.1]), 1.9, 50.0
    g_cur = np.array([0.05, -0.02, -1.0])
    g_cur /= np.linalg.norm(g_cur)
    g_ref, wg = np.array([0.0, 0.0, -1.0]), 200.0

    def resid(xx, use_w, use_g):
        R = _euler_zyx(*xx[3:])
        d = X @ R.T + xx[:3] - P
        r = [(np.sqrt(w)[:, None] * np.einsum("nij,nj->ni", A, d)).ravel()]
        if use_w:
            r.append([np.sqrt(ww) * (np.linalg.norm(xx[:3] - prev) - dist)])
        if use_g:
            r.append(np.sqrt(wg) * (R @ g_cur - g_ref))
        return np.concatenate(r)

    base = oracle_lib.lm_solve(A, P, X, w, np.zeros(6), 1e3, max_iter=50)["x"]
    for use_w, use_g in ((True, False), (False, True), (True, True)):
        ref = least_squares(resid, np.zeros(6), args=(use_w, use_g), method="lm", xtol=1e-14, ftol=1e-14, gtol=1e-14)
        got = oracle_lib.lm_solve(A, P, X, w, np.zeros(6), 1e3, max_iter=50, wheel=(prev, dist, ww) if use_w else None,
                                  gravity=(g_ref, g_cur, wg) if use_g else None)
        assert got["term"] in (1, 2, 3)
        assert np.abs(got["x"] - ref.x).max() < 1e-5, (use_w, use_g, got["x"] - ref.x)   # function tolerance 1e-6 stops in a flat valley
        assert np.abs(got["x"] - base).max() > 1e-4            # the constraint did pull the solution
        # cost of the extra blocks alone, at a fixed point (max_iter = 0: one evaluation, no step)
        c1 = oracle_lib.lm_solve(A, P, X, w, x_true, 1e3, max_iter=0, wheel=(prev, dist, ww) if use_w else None,
                                 gravity=(g_ref, g_cur, wg) if use_g else None)["cost"]
        c0 = oracle_lib.lm_solve(A, P, X, w, x_true, 1e3, max_iter=0)["cost"]
        extra = 0.5 * (resid(x_true, use_w, use_g)[3 * len(A):] ** 2).sum()
        assert abs((c1 - c0) - extra) < 1e-12 * extra
    # degenerate odometer block: |t - prev|^2 < 1e-6 counts as zero motion with zero derivative
    got = oracle_lib.lm_solve(A[:0], P[:0], X[:0], w[:0], np.array([1e-4, 0, 0, 0, 0, 0.]), 1.0, max_iter=5, wheel=(np.zeros(3), 0.5, 1.0))
    assert np.array_equal(got["x"], [1e-4, 0, 0, 0, 0, 0]) and got["cost"] == 0.5 * 0.25

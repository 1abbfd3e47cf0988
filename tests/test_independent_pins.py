"""Independent pins (CPU) for the arithmetic that the product and the oracle SHARE through
lidarslam_b200/csrc/common/*.h.  "GPU == oracle bit for bit" proves nothing about those headers, so each block is
checked here against a restatement that does not include them: numpy.linalg for the closed-form 3x3 eigen solver (every
branch driven on purpose), numpy.linalg.pinv / eigh for the registration error, a numpy trust-region loop written from
SURVEY.md Appendix A.3 (QR on the stacked Jacobian as Ceres does, explicit derivative matrices) for the LM state machine,
a pure-Python dictionary model of RollingGrid::Add for the five sampling modes, and plain-Python transcriptions of the
motion-limit check."""
import numpy as np
import pytest


# ------------------------------------------------------------------ eigen33 (common/pca.h) ----
def _perms():
    """exact signed permutation matrices (rotations by multiples of 90 degrees): conjugation is exact in floating point"""
    out = []
    for p in ((0, 1, 2), (1, 2, 0), (2, 0, 1), (0, 2, 1), (2, 1, 0), (1, 0, 2)):
        for sg in ((1, 1, 1), (-1, 1, 1), (1, -1, -1)):
            Q = np.zeros((3, 3))
            for i in range(3):
                Q[i, p[i]] = sg[i]
            out.append(Q)
    return out


def _check_eig(oracle_lib, mats, as_float, val_tol, vec_tol, label):
    mats = np.asarray(mats, dtype=np.float64)
    if as_float:
        mats = mats.astype(np.float32).astype(np.float64)
    ev, vec = oracle_lib.eigen33(mats, as_float=as_float)
    ref_val, ref_vec = np.linalg.eigh(mats)
    scale = np.maximum(np.abs(mats).reshape(-1, 9).max(1), 1e-300)
    assert (np.diff(ev, axis=1) >= 0).all(), label + ": eigenvalues not ascending"
    assert (np.abs(ev - ref_val) <= val_tol * scale[:, None]).all(), (label, np.abs(ev - ref_val).max())
    for i, (m, l, V) in enumerate(zip(mats, ev, vec)):
        # columns orthonormal
        assert np.abs(V.T @ V - np.eye(3)).max() <= vec_tol, (label, i, "not orthonormal")
        # every column lies in the eigenspace of its eigenvalue: |(m - l_k) v_k| small relative to the scale
        for k in range(3):
            assert np.linalg.norm(m @ V[:, k] - l[k] * V[:, k]) <= vec_tol * max(scale[i], 1e-300) * 4, (label, i, k)


@pytest.mark.parametrize("as_float", [False, True])
def test_eigen33_every_branch(oracle_lib, as_float):
    """pcl::eigen33 as restated in common/pca.h: the four eigenvector branches (all roots equal / two smallest equal /
    two largest equal / distinct) and both root solvers (cubic, computeRoots2 when det == 0 or the smallest root <= 0)."""
    val_tol, vec_tol = (3e-6, 2e-3) if as_float else (1e-14, 1e-7)
    Qs = _perms()
    # all eigenvalues equal (evecs = identity), including the all-zero covariance of duplicated points (scale := 1)
    _check_eig(oracle_lib, [np.eye(3) * s for s in (0.0, 1.0, 2.5e-7, 3e5)], as_float, val_tol, vec_tol, "isotropic")
    # exactly collinear neighbourhoods: rank 1, det == 0 -> computeRoots2, two SMALLEST roots equal (0, 0, l)
    rank1 = []
    for d in ([1, 0, 0], [0, 1, 0], [0, 0, 1], [1, 1, 0], [1, -1, 0], [0, 1, 1], [1, 1, 1], [2, 1, 0], [4, -2, 1]):
        d = np.asarray(d, float)
        for s in (1.0, 0.25, 64.0):
            rank1.append(s * np.outer(d, d))
    _check_eig(oracle_lib, rank1, as_float, val_tol, vec_tol, "collinear")
    # exactly coplanar: rank 2, det == 0 -> computeRoots2 with distinct non-zero roots, and with two LARGEST equal
    rank2 = []
    for Q in Qs:
        rank2.append(Q @ np.diag([0.0, 1.0, 3.0]) @ Q.T)
        rank2.append(Q @ np.diag([0.0, 0.5, 0.5]) @ Q.T)
        rank2.append(Q @ np.diag([0.0, 2.0, 2.0]) @ Q.T * 1e-3)
    # coplanar but not axis aligned: points on the plane x + y + z = 0 (integer coordinates, exact covariance)
    u, v = np.array([1., -1, 0]), np.array([1., 1, -2])
    rank2.append(np.outer(u, u) + np.outer(v, v))
    rank2.append(4 * np.outer(u, u) + np.outer(v, v))
    _check_eig(oracle_lib, rank2, as_float, val_tol, vec_tol, "coplanar")
    # full rank, two equal eigenvalues (smallest pair / largest pair), exact conjugations
    deg = []
    for Q in Qs:
        deg.append(Q @ np.diag([1.0, 1.0, 4.0]) @ Q.T)
        deg.append(Q @ np.diag([1.0, 4.0, 4.0]) @ Q.T)
        deg.append(Q @ np.diag([0.125, 0.125, 0.5]) @ Q.T)
    _check_eig(oracle_lib, deg, as_float, val_tol, vec_tol, "two equal")
    # distinct, well separated (the cubic branch), exact conjugations and generic rotations
    rng = np.random.default_rng(11)
    gen = []
    for Q in Qs:
        gen.append(Q @ np.diag([1.0, 2.0, 5.0]) @ Q.T)
    for _ in range(300):
        Q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        l = np.sort(rng.uniform(0.05, 1.0, 3)) * np.array([1.0, 2.0, 4.0])
        gen.append(Q @ np.diag(l) @ Q.T)
    _check_eig(oracle_lib, gen, as_float, max(val_tol, 1e-12), max(vec_tol, 1e-6), "distinct")


@pytest.mark.parametrize("as_float", [False, True])
def test_eigen33_near_degenerate_clouds(oracle_lib, as_float):
    """Covariances of real degenerate neighbourhoods, as the matcher meets them: noisy lines, noisy planes, a point cloud
    with duplicates.  Eigenvalues against numpy; the vectors the callers use (largest for a line, smallest for a plane)
    against numpy's up to sign wherever the spectral gap makes them well defined."""
    rng = np.random.default_rng(12)
    mats, kinds = [], []
    for _ in range(400):
        n = int(rng.integers(4, 12))
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        t = rng.uniform(-1, 1, n)
        line = np.outer(t, d) + rng.normal(0, 1e-3, (n, 3)) + rng.uniform(-30, 30, 3)
        mats.append(np.cov(line.T, bias=True)); kinds.append(0)
        Q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
        plane = (rng.uniform(-1, 1, (n, 3)) * np.array([1.0, 0.7, 1e-3])) @ Q.T + rng.uniform(-30, 30, 3)
        mats.append(np.cov(plane.T, bias=True)); kinds.append(1)
        dup = np.repeat(rng.normal(size=(2, 3)), 3, axis=0)
        mats.append(np.cov(dup.T, bias=True)); kinds.append(0)
    mats = np.array(mats)
    if as_float:
        mats = mats.astype(np.float32).astype(np.float64)
    ev, vec = oracle_lib.eigen33(mats, as_float=as_float)
    rv, rV = np.linalg.eigh(mats)
    scale = np.abs(mats).reshape(-1, 9).max(1)
    # the trigonometric closed form loses digits on close roots (two eigenvalues ~1e-7 apart here): ~1e-10 of the scale
    tol = 2e-4 if as_float else 1e-9
    assert (np.abs(ev - rv) <= tol * scale[:, None]).all()
    for kind, col in ((0, 2), (1, 0)):
        sel = np.flatnonzero(np.asarray(kinds) == kind)
        dots = np.abs(np.einsum("ni,ni->n", vec[sel, :, col], rV[sel, :, col]))
        assert (dots > 1 - (1e-3 if as_float else 1e-9)).all(), (kind, dots.min())


# ------------------------------------------------------ registration error (common/lm_covariance.h) ----
def _h21(H):
    return np.array([H[i, j] for i in range(6) for j in range(i, 6)])


@pytest.mark.parametrize("two_d", [False, True])
def test_registration_error_against_numpy(oracle_lib, two_d):
    """LocalOptimizer::EstimateRegistrationError (LocalOptimizer.cxx:112-140): ceres::Covariance with DENSE_SVD,
    min_reciprocal_condition_number 1e-14, null_space_rank -1 = the pseudo-inverse of J^T J with the singular values of
    J below 1e-7 * sigma_max dropped; then the largest eigenpair of the position and of the orientation 3x3 block."""
    rng = np.random.default_rng(13)
    free = [0, 1, 5] if two_d else list(range(6))
    cases = []
    for _ in range(60):                                    # well conditioned
        J = rng.normal(size=(40, 6)) * rng.uniform(0.1, 30, 6)
        cases.append(J.T @ J)
    for _ in range(30):                                    # rank deficient: a direction the residuals do not see at all
        J = rng.normal(size=(40, 6))
        n = rng.normal(size=6)
        if two_d:
            n[[2, 3, 4]] = 0
        n /= np.linalg.norm(n)
        J = J - np.outer(J @ n, n)
        cases.append(J.T @ J)
    for _ in range(30):                                    # nearly deficient, on both sides of the 1e-14 cut
        U, _ = np.linalg.qr(rng.normal(size=(6, 6)))
        lam = 10.0 ** rng.uniform(-1, 2, 6)
        lam[0] = lam.max() * 10.0 ** rng.choice([-12.0, -13.5, -14.5, -16.0])
        if two_d:                                          # keep the structure inside the free block
            Uf, _ = np.linalg.qr(rng.normal(size=(3, 3)))
            H = np.zeros((6, 6))
            H[np.ix_(free, free)] = Uf @ np.diag(lam[:3]) @ Uf.T
            H[2, 2] = H[3, 3] = H[4, 4] = 7.0              # held parameters: must be ignored
            cases.append(H)
        else:
            cases.append(U @ np.diag(lam) @ U.T)
    for H in cases:
        H = (H + H.T) / 2
        e = oracle_lib.registration_error(_h21(H), two_d)
        got = np.array(e.covariance).reshape(6, 6)
        Hf = H[np.ix_(free, free)]
        ref = np.zeros((6, 6))
        ref[np.ix_(free, free)] = np.linalg.pinv(Hf, rcond=1e-14, hermitian=True)
        lam = np.linalg.eigvalsh(Hf)
        ratio = lam / lam.max()
        if ((ratio > 3e-15) & (ratio < 3e-14)).any():
            continue                                       # on the cut itself: either side is legitimate
        # a kept eigenvalue l of relative size q is known to eps / q in ANY method, and 1 / l dominates the result
        rel = max(1e-9, 20 * 2.2e-16 / ratio[ratio > 3e-14].min())
        tol = np.abs(ref).max() * rel
        assert np.abs(got - ref).max() <= tol, (np.abs(got - ref).max(), tol)
        assert np.abs(got - got.T).max() <= 1e-12 * np.abs(got).max()
        for blk, val, direction, deg in ((slice(0, 3), e.position_error, e.position_error_direction, False),
                                         (slice(3, 6), e.orientation_error, e.orientation_error_direction, True)):
            lv, lV = np.linalg.eigh(ref[blk, blk])
            expect = np.sqrt(max(lv[2], 0.0))
            if deg:
                expect = expect / np.pi * 180.0
            assert abs(val - expect) <= rel * max(expect, 1e-12) + 1e-15
            d = np.array(direction)
            if lv[2] - lv[1] > 1e-6 * max(lv[2], 1e-300):
                assert abs(abs(d @ lV[:, 2]) - 1) < max(1e-8, rel)
            assert abs(np.linalg.norm(d) - 1) < 1e-9


# ------------------------------------------- trust-region LM (common/lm_core.h) vs SURVEY Appendix A.3 ----
def _rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    dRx = np.array([[0, 0, 0], [0, -sx, -cx], [0, cx, -sx]])
    dRy = np.array([[-sy, 0, cy], [0, 0, 0], [-cy, 0, -sy]])
    dRz = np.array([[-sz, -cz, 0], [cz, -sz, 0], [0, 0, 0]])
    return Rz @ Ry @ Rx, (Rz @ Ry @ dRx, Rz @ dRy @ Rx, dRz @ Ry @ Rx)


class NumpyCeres:
    """ceres::Solve as LocalOptimizer::Solve configures it (TRUST_REGION, LEVENBERG_MARQUARDT, DENSE_QR, defaults
    elsewhere), from SURVEY.md Appendix A.3.  Residual blocks r_i = A_i (R(rpy) X_i + t - P_i) under
    ScaledLoss(TukeyLoss(sat), w_i) (CeresCostFunctions.h:105-152, KeypointsMatcher.cxx:78-103)."""
    GRADIENT, PARAMETER, FUNCTION, RADIUS, MAX_ITER, INVALID = 1, 2, 3, 4, 5, 6

    def __init__(self, A, P, X, w, sat, max_iter=15, two_d=False):
        keep = w >= 0
        self.A, self.P, self.X, self.w = A[keep].reshape(-1, 3, 3), P[keep], X[keep], w[keep]
        self.a2 = sat * sat
        self.max_iter = max_iter
        self.free = [0, 1, 5] if two_d else [0, 1, 2, 3, 4, 5]

    def evaluate(self, x, jacobian):
        R, dR = _rot(*x[3:])
        d = self.X @ R.T + x[:3] - self.P
        r = np.einsum("nij,nj->ni", self.A, d)
        s = (r * r).sum(1)
        inl = s <= self.a2
        v = 1.0 - s / self.a2
        rho0 = self.w * np.where(inl, self.a2 / 3.0 * (1.0 - v ** 3), self.a2 / 3.0)
        rho1 = self.w * np.where(inl, v * v, 0.0)
        cost = 0.5 * rho0.sum()
        if not jacobian:
            return cost
        sq = np.sqrt(rho1)                                 # Corrector, rho'' <= 0 branch
        J = np.zeros((len(r), 3, 6))
        J[:, :, 0:3] = self.A
        for k in range(3):
            J[:, :, 3 + k] = np.einsum("nij,nj->ni", self.A, self.X @ dR[k].T)
        J *= sq[:, None, None]
        return cost, (sq[:, None] * r).reshape(-1), J.reshape(-1, 6)[:, self.free]

    def solve(self, x0):
        x = np.array(x0, float)
        cost, r, J = self.evaluate(x, True)
        S = 1.0 / (1.0 + np.linalg.norm(J, axis=0))        # Jacobi scaling, frozen
        Jb = J * S
        radius, dec, it, invalid, accepted = 1e4, 2.0, 0, 0, 0
        trace = []
        if np.abs(J.T @ r).max() <= 1e-10:
            return x, cost, it, accepted, self.GRADIENT, trace
        while True:
            if it >= self.max_iter:
                term = self.MAX_ITER; break
            if radius < 1e-32:
                term = self.RADIUS; break
            it += 1
            D = np.sqrt(np.clip((Jb * Jb).sum(0), 1e-6, 1e32) / radius)
            aug = np.vstack([Jb, np.diag(D)])
            rhs = np.concatenate([-r, np.zeros(len(D))])
            Q, Rq = np.linalg.qr(aug)                      # DENSE_QR on the stacked system
            delta = np.linalg.solve(Rq, Q.T @ rhs)
            mr = Jb @ delta
            mcc = -mr @ (r + mr / 2.0)
            if not (mcc > 0 and np.isfinite(delta).all()):
                invalid += 1
                if invalid >= 5:
                    term = self.INVALID; break
                radius *= 0.5
                continue
            invalid = 0
            step = np.zeros(6)
            step[self.free] = delta * S
            xc = x + step
            cost_c = self.evaluate(xc, False)
            trace.append([it, cost, cost_c, radius, 0.0])
            if np.linalg.norm(x - xc) <= 1e-8 * (np.linalg.norm(x) + 1e-8):
                term = self.PARAMETER; break
            if abs(cost - cost_c) <= 1e-6 * cost:
                term = self.FUNCTION; break
            q = (cost - cost_c) / mcc
            if q > 1e-3:
                x = xc
                cost, r, J = self.evaluate(x, True)
                Jb = J * S
                accepted += 1
                trace[-1][4] = 1.0
                radius = min(1e16, radius / max(1.0 / 3.0, 1.0 - (2.0 * q - 1.0) ** 3))
                dec = 2.0
                if np.abs(J.T @ r).max() <= 1e-10:
                    term = self.GRADIENT; break
            else:
                radius /= dec
                dec *= 2.0
        return x, cost, it, accepted, term, np.array(trace)


def _frame_records(oracle_lib, synth, sensor_name, n_frames, blobs, rng, max_per_type):
    """Residual records of a real frame: run the oracle a few frames to get keypoints and sub-maps, then build the
    point-to-line / plane / blob models in numpy (cKDTree neighbours, numpy eigh), independently of the matcher."""
    from scipy.spatial import cKDTree
    o = oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
    p = o.GetParams()
    p.use_keypoints[2] = 1 if blobs else 0
    o.SetParams(p)
    seq = synth.Sequence(scene_seed=77)
    sens = synth.sensor(sensor_name)
    for i in range(n_frames):
        pts, _, stamp = seq.frame(sens, i)
        o.AddFrame(pts, stamp, i + 1)
    T, _ = o.GetWorldTransform()
    A, P, X, w = [], [], [], []
    for k, kk in ((0, 10), (1, 5), (2, 10)):
        kp, sub = o.GetKeypoints(k), o.GetTargetSubMap(k)
        if len(kp) == 0 or len(sub) < kk:
            continue
        if len(kp) > max_per_type:
            kp = kp[np.sort(rng.permutation(len(kp))[:max_per_type])]
        xyz = np.stack([kp["x"], kp["y"], kp["z"]], 1).astype(np.float64)
        world = xyz @ T[:3, :3].T + T[:3, 3]
        cloud = np.stack([sub["x"], sub["y"], sub["z"]], 1).astype(np.float64)
        dist, idx = cKDTree(cloud).query(world, k=kk)
        for i in np.flatnonzero(dist[:, -1] < 3.0):
            nb = cloud[idx[i]]
            mean = nb.mean(0)
            lam, V = np.linalg.eigh(np.cov(nb.T, bias=True))
            if k == 0:
                if lam[0] + lam[1] > 0.04:
                    continue
                n = V[:, 2]
                Ai, wi = np.eye(3) - np.outer(n, n), 1.0 - np.sqrt(max(lam[0] + lam[1], 0)) / 0.2
            elif k == 1:
                if lam[1] < 0.04 * lam[2] or lam[0] > 0.04:
                    continue
                n = V[:, 0]
                Ai, wi = np.outer(n, n), 1.0 - np.sqrt(max(lam[0], 0)) / 0.2
            else:
                if lam[0] <= 1e-9:
                    continue
                Ai, wi = V @ np.diag(1.0 / np.sqrt(lam)) @ V.T, 1.0
            A.append(Ai.reshape(9)); P.append(mean); X.append(xyz[i]); w.append(wi)
    return np.array(A), np.array(P), np.array(X), np.array(w), T


def _xyzrpy(T):
    return np.array([T[0, 3], T[1, 3], T[2, 3], np.arctan2(T[2, 1], T[2, 2]), -np.arcsin(T[2, 0]), np.arctan2(T[1, 0], T[0, 0])])


@pytest.mark.parametrize("sensor,blobs,nmax", [("vlp16", False, 4000), ("os128", True, 12000)])
def test_trust_region_loop_against_numpy_ceres(oracle_lib, synth, sensor, blobs, nmax):
    """The LM state machine of common/lm_core.h (shared by the product and the oracle) against NumpyCeres on the residual
    records of a real frame, from a perturbed prior so that the solve has steps to take: same accept / reject sequence,
    same number of iterations, same termination, costs along the way to 1e-9 relative, final x to 1e-9."""
    rng = np.random.default_rng(21)
    A, P, X, w, T = _frame_records(oracle_lib, synth, sensor, 4, blobs, rng, nmax)
    assert len(w) > (1000 if sensor == "vlp16" else 8000)
    x_true = _xyzrpy(T)
    checked = 0
    for trial, (sat, max_iter, two_d, noise) in enumerate([(1.0, 15, False, 1.0), (0.2, 15, False, 1.0), (2.0, 15, False, 0.2),
                                                           (1.0, 6, False, 3.0), (1.0, 15, True, 1.0), (0.5, 40, False, 0.5)]):
        x0 = x_true + noise * rng.normal(size=6) * np.array([0.05, 0.05, 0.02, 0.003, 0.003, 0.01])
        if two_d:
            x0[[2, 3, 4]] = x_true[[2, 3, 4]]
        ww = w.copy()
        ww[rng.random(len(ww)) < 0.05] = -1.0              # failed matches: no residual
        got = oracle_lib.lm_solve_trace(A, P, X, ww, x0, sat, max_iter=max_iter, two_d=two_d, threads=4)
        x, cost, it, acc, term, trace = NumpyCeres(A, P, X, ww, sat, max_iter, two_d).solve(x0)
        assert got["iterations"] == it and got["accepted"] == acc and got["term"] == term, (trial, got["iterations"], it, got["accepted"], acc, got["term"], term)
        assert len(got["trace"]) == len(trace)
        if len(trace):
            assert (got["trace"][:, [0, 4]] == trace[:, [0, 4]]).all(), (trial, "accept/reject sequence")
            assert np.allclose(got["trace"][:, 1:4], trace[:, 1:4], rtol=1e-9, atol=0), trial
        assert np.abs(got["x"] - x).max() <= 1e-9, (trial, np.abs(got["x"] - x).max())
        assert abs(got["cost"] - cost) <= 1e-9 * cost
        # the normal matrix handed to the covariance estimate is J~^T J~ at the solution
        _, _, J = NumpyCeres(A, P, X, ww, sat, max_iter, False).evaluate(got["x"], True)
        H = J.T @ J
        assert np.allclose(got["H"], _h21(H), rtol=1e-9, atol=1e-9 * np.abs(H).max())
        checked += it
    assert checked > 20     # the solves did iterate


def test_trust_region_rejections_and_invalid_steps(oracle_lib):
    """Synthetic problems that force what real frames rarely produce: rejected steps (radius shrinks by 2, 4, 8 ...),
    an immediate gradient-tolerance exit, and a parameter-tolerance exit."""
    rng = np.random.default_rng(22)
    n = 600
    X = rng.uniform(-20, 20, (n, 3))
    x_true = np.array([0.3, -0.2, 0.1, 0.01, -0.02, 0.4])
    R, _ = _rot(*x_true[3:])
    normals = rng.normal(size=(n, 3))
    normals /= np.linalg.norm(normals, axis=1, keepdims=True)
    A = np.einsum("ni,nj->nij", normals, normals).reshape(n, 9)
    P = X @ R.T + x_true[:3] + rng.normal(0, 0.01, (n, 3))
    w = np.ones(n)
    seen_reject = 0
    Aid = np.tile(np.eye(3).reshape(9), (n, 1))          # point-to-point: strongly non-linear in the angles
    rng2 = np.random.default_rng(23)
    for Ak, sat in ((A, 1e3), (Aid, 3.0), (Aid, 30.0), (Aid, 1e3)):
        for _ in range(8):
            x0 = x_true + np.array([1, 1, 1, 1.2, 1.2, 2.5]) * rng2.normal(size=6)   # far outside the basin of Gauss-Newton
            got = oracle_lib.lm_solve_trace(Ak, P, X, w, x0, sat, max_iter=40, threads=1)
            x, cost, it, acc, term, trace = NumpyCeres(Ak, P, X, w, sat, 40, False).solve(x0)
            assert (got["iterations"], got["accepted"], got["term"]) == (it, acc, term)
            assert len(got["trace"]) == len(trace)
            if len(trace):
                assert (got["trace"][:, 4] == trace[:, 4]).all()
                # long non-linear paths: the radius update cubes the step quality, rounding differences grow along the way
                assert np.allclose(got["trace"][:, 1:3], trace[:, 1:3], rtol=1e-7) and np.allclose(got["trace"][:, 3], trace[:, 3], rtol=1e-5)
                seen_reject += int((trace[:, 4] == 0).sum() > 1)
            assert np.abs(got["x"] - x).max() <= 1e-8
    assert seen_reject >= 4, "too few cases exercised rejected steps"
    # gradient tolerance at iteration 0: every residual saturated (rho' = 0) -> zero gradient
    far = oracle_lib.lm_solve_trace(A, P, X, w, x_true + np.array([50., 0, 0, 0, 0, 0]), 0.01, max_iter=15)
    assert far["term"] == NumpyCeres.GRADIENT and far["iterations"] == 0 and len(far["trace"]) == 0
    # exact data, start at the solution: the first step is below the parameter tolerance (or the gradient is already flat)
    P0 = X @ R.T + x_true[:3]
    got = oracle_lib.lm_solve_trace(A, P0, X, w, x_true, 1.0, max_iter=15)
    x, cost, it, acc, term, trace = NumpyCeres(A, P0, X, w, 1.0, 15, False).solve(x_true)
    assert got["term"] in (NumpyCeres.GRADIENT, NumpyCeres.PARAMETER, NumpyCeres.FUNCTION) and got["accepted"] <= 1
    assert np.abs(got["x"] - x_true).max() < 1e-9


# ----------------------------------------------------- RollingGrid::Add, every sampling mode ----
def _rnd(v):
    """Eigen array round(): half away from zero"""
    return int(np.floor(v + np.float32(0.5))) if v >= 0 else int(np.ceil(v - np.float32(0.5)))


class PyRollingGrid:
    """RollingGrid (RollingGrid.cxx:117-318) on Python dictionaries, float32 arithmetic through numpy scalars; written
    from the reference, not from the oracle: unordered_map semantics only (no ordering assumptions)."""

    def __init__(self, grid=50, res=10.0, leaf=0.3, mode=2):
        self.G, self.leaf, self.mode = grid, leaf, mode
        self.res = int(res / leaf) * leaf                     # SetVoxelResolution (:73-80)
        self.pos = np.zeros(3, np.float32)
        self.vox = {}                                         # (idxOut, idxIn) -> [point record, count]

    def roll(self, mn, mx):
        f = np.float32
        half = f(self.G / 2 * self.res)
        res = f(self.res)
        off = np.zeros(3, np.int64)
        for a in range(3):
            down = f(mn[a] - f(self.pos[a] - half))
            up = f(mx[a] - f(self.pos[a] + half))
            o = f(f(up + down) / f(2))
            o = max(o, min(down, f(0)))
            o = min(o, max(up, f(0)))
            off[a] = _rnd(f(o / res))
        if not off.any():
            return
        G, new = self.G, {}
        for (io, ii), v in self.vox.items():
            z, rem = divmod(io, G * G)
            y, x = divmod(rem, G)
            x, y, z = x - off[0], y - off[1], z - off[2]
            if 0 <= x < G and 0 <= y < G and 0 <= z < G:
                new[(z * G * G + y * G + x, ii)] = v
        self.vox = new
        self.pos = (self.pos + off.astype(np.float32) * res).astype(np.float32)

    def add(self, pts, fixed, t):
        f = np.float32
        xyz = np.stack([pts["x"], pts["y"], pts["z"]], 1)
        self.roll(xyz.min(0), xyz.max(0))
        G, res, leaf = self.G, f(self.res), f(self.leaf)
        origin = (self.pos - f(int(G / 2) * self.res)).astype(np.float32)
        seen, mean = set(), {}
        for n in range(len(pts)):
            p = xyz[n]
            vo = [_rnd(f(f(p[a] - origin[a]) / res)) for a in range(3)]
            if not all(0 <= c < G for c in vo):
                continue
            ctr = np.array([f(f(vo[a]) * res) + origin[a] for a in range(3)], np.float32)
            vi = [_rnd(f(f(p[a] - ctr[a]) / leaf)) for a in range(3)]
            key = (vo[2] * G * G + vo[1] * G + vo[0], vi[2] * G * G + vi[1] * G + vi[0])
            if key not in self.vox:
                self.vox[key] = [pts[n].copy(), 0]
            else:
                v = self.vox[key]
                if v[0]["label"] == 1:
                    continue
                if self.mode == 1:
                    v[0] = pts[n].copy()
                elif self.mode == 2:
                    if pts[n]["intensity"] > v[0]["intensity"]:
                        v[0] = pts[n].copy()
                elif self.mode == 3:
                    c = np.array([f(f(ctr[a] - f(res / f(2))) + f(leaf * f(vi[a]))) for a in range(3)], np.float32)
                    old = np.array([v[0]["x"], v[0]["y"], v[0]["z"]], np.float32)
                    if np.sqrt(((p - c) ** 2).sum(dtype=np.float32)) < np.sqrt(((old - c) ** 2).sum(dtype=np.float32)):
                        v[0] = pts[n].copy()
                elif self.mode == 4:
                    m = mean.setdefault(key, [np.zeros(3, np.float32), 0])
                    m[0] = ((m[0] * f(m[1]) + p) / f(m[1] + 1)).astype(np.float32)
                    m[1] += 1
            if self.mode == 4:
                for mk, m in mean.items():
                    v = self.vox[mk]
                    old = np.array([v[0]["x"], v[0]["y"], v[0]["z"]], np.float32)
                    new = ((old * f(v[1]) + m[0]) / f(v[1] + 1)).astype(np.float32)
                    v[0]["x"], v[0]["y"], v[0]["z"] = new
            v = self.vox[key]
            v[0]["time"] = t
            v[0]["label"] = 1 if fixed else 0
            if key not in seen:
                seen.add(key)
                v[1] += 1

    def cloud(self):
        return np.array([self.vox[k][0] for k in sorted(self.vox, key=lambda k: (k[0], k[1]))])


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_rolling_grid_add_every_mode_against_python_model(oracle_lib, mode):
    from lidarslam_b200 import capi
    rng = np.random.default_rng(30 + mode)
    centres = rng.uniform(-12, 12, (25, 3))

    def cloud(n, spread, shift=0.0):
        p = np.zeros(n, dtype=capi.POINT_DTYPE)
        c = centres[rng.integers(0, len(centres), n)] + shift
        xyz = (c + rng.normal(0, spread, (n, 3))).astype(np.float32)
        p["x"], p["y"], p["z"], p["w"] = xyz[:, 0], xyz[:, 1], xyz[:, 2], 1.0
        p["intensity"] = rng.uniform(0, 255, n).astype(np.float32)
        p["laser_id"] = rng.integers(0, 16, n)
        return p

    o = oracle_lib.Oracle()
    o.SetVoxelGridSize(8)
    o.SetVoxelGridResolution(6.0)
    o.SetVoxelGridSamplingMode(0, mode)
    model = PyRollingGrid(grid=8, res=6.0, leaf=0.3, mode=mode)
    steps = [(500, 0.3, False, 0.0), (500, 0.3, False, 0.0), (300, 0.5, True, 0.0), (500, 0.4, False, 0.0), (400, 0.4, False, 9.0)]
    for i, (n, spread, fixed, shift) in enumerate(steps):
        c = cloud(n, spread, shift)
        o.MapAdd(0, c, fixed, 0.1 * i)
        model.add(c, fixed, 0.1 * i)
        got, exp = o.GetMap(0), model.cloud()
        assert len(got) == len(exp) and len(got) > 50, (i, len(got), len(exp))
        for fld in ("x", "y", "z", "intensity", "time", "label", "laser_id"):
            assert np.array_equal(got[fld], exp[fld]), "mode %d add %d field %s" % (mode, i, fld)


# ----------------------------------------------------------------- motion limits (common/motion_limits.h) ----
def test_motion_limits_against_python(oracle_lib, synth):
    """Slam::CheckMotionLimits (Slam.cxx:1391-1484) through the oracle on a sequence whose motion is known: constant
    velocity 8 m/s straight scene drive.  With limits above the true motion the pose complies, with limits below it
    does not; the window needs TimeWindowDuration of logged poses before it judges."""
    sens = synth.sensor("vlp16")
    for vlim, expect in ((20.0, True), (2.0, False)):
        o = oracle_lib.Oracle(nb_threads=oracle_lib.max_threads())
        p = o.GetParams()
        p.time_window_duration = 0.35
        p.logging_timeout = 2.0
        p.velocity_limits[0], p.velocity_limits[1] = vlim, 400.0
        p.acceleration_limits[0], p.acceleration_limits[1] = 1e4, 1e5
        o.SetParams(p)
        seq = synth.Sequence(scene_seed=5)
        flags = []
        for i in range(8):
            pts, _, stamp = seq.frame(sens, i)
            rc = o.AddFrame(pts, stamp, i + 1)
            flags.append((rc, o.GetComplyMotionLimits()))
        assert flags[-1][1] == expect, (vlim, flags)
        assert bool(flags[-1][0] & 128) == (not expect), flags      # LSB_W_MOTION_LIMITS
        assert flags[1][1]                                           # window not filled yet: complies by default

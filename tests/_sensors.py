"""Synthetic wheel-odometer / IMU streams for the sensor-constraint tests, derived from the ground-truth
trajectory of tools/synth: the odometer reports the chord distance to the first frame's position (what the reference's
absolute constraint compares with |t|, SensorConstraints.cxx:8-57 with an identity PreviousPose), the IMU the gravity
vector in the BASE frame."""
import numpy as np


def feed(slams, seq, sens, n_frames, noise=0.0, seed=0):
    """Adds one measurement per frame stamp (plus one 0.1 s before the first) to every handle in `slams`."""
    rng = np.random.default_rng(seed)
    gts = [seq.frame(sens, i)[1:] for i in range(n_frames)]
    p0 = gts[0][0][:3, 3]
    t0 = gts[0][1] * 1e-6
    wheel = [(t0 - 0.1, 0.0)]
    grav = [(t0 - 0.1, gts[0][0][:3, :3].T @ np.array([0, 0, -9.81]))]
    for T, stamp in gts:
        wheel.append((stamp * 1e-6, float(np.linalg.norm(T[:3, 3] - p0)) + noise * rng.normal()))
        grav.append((stamp * 1e-6, T[:3, :3].T @ np.array([0, 0, -9.81]) + noise * rng.normal(size=3)))
    for s in slams:
        for t, d in wheel:
            s.AddWheelOdomMeasurement(t, d)
        for t, a in grav:
            s.AddGravityMeasurement(t, a)
    return wheel, grav

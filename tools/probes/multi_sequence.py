"""Probe (not product): S independent OS-128 sequences on ONE GPU, one handle + host thread each ("several
handles per GPU interleave on separate streams to fill SMs left idle by the sequential phases").  Prints the aggregate
frames/s for S = 1, 2, 4, 8 and checks that every sequence ends on the pose it reaches when run alone.
    python tools/probes/multi_sequence.py [frames] [sensor]"""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
from lidarslam_b200.slam import Slam  # noqa: E402
from tools.synth import synth  # noqa: E402


def configure(s):
    p = s.GetParams()
    p.use_keypoints[2] = 1
    p.undistortion = 2
    s.SetParams(p)


def main():
    n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    sensor = sys.argv[2] if len(sys.argv) > 2 else "os128"
    warm = 4
    sens = synth.sensor(sensor)
    seqs = []
    for s in range(8):
        sq = synth.Sequence(scene_seed=1234 + s)
        seqs.append([sq.frame(sens, i) for i in range(warm + n_frames)])
    npts = len(seqs[0][0][0])
    alone = {}
    for S in (1, 2, 4, 8):
        hs = [Slam(device=0) for _ in range(S)]
        dptrs = []
        for h, frames in zip(hs, seqs):
            configure(h)
            dptrs.append([h.DevUpload(f[0]) for f in frames])
            for i in range(warm):
                h.AddFramesDev([dptrs[-1][i]], [npts], [frames[i][2]], [i + 1])
        go = threading.Barrier(S + 1)

        def work(h, frames, dp):
            go.wait()
            for i in range(warm, warm + n_frames):
                h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])

        ts = [threading.Thread(target=work, args=(h, f, d)) for h, f, d in zip(hs, seqs, dptrs)]
        for t in ts:
            t.start()
        go.wait()
        t0 = time.perf_counter()
        for t in ts:
            t.join()
        dt = time.perf_counter() - t0
        poses = [h.GetWorldTransform()[0] for h in hs]
        if S == 1:
            alone[0] = poses[0]
        same = all(np.array_equal(poses[0], alone[0]) for _ in (0,))
        print("S = %d sequences on one GPU: %7.1f frames/s aggregate (%.3f ms per frame and sequence), sequence 0 pose identical to the solo run: %s"
              % (S, S * n_frames / dt, 1e3 * dt / n_frames, same), flush=True)
        for h in hs:
            h.close()


if __name__ == "__main__":
    main()

"""Generates tests/golden/tiny_sequence.json from the CPU oracle (oracle/).

The reference ships no golden vectors (ros_wrapping/tests/README.md:3-4) and cannot be built
here, so these fixtures pin the ORACLE's own outputs (labels, keypoint counts, match counts, map
sizes, poses) on a small seeded sequence: they guard the oracle — and, in the -m gpu suite, the
CUDA path — against silent drift.  Run:  python tests/golden/make_golden.py"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

N_FRAMES = 8


def configure(s):
    p = s.GetParams()
    p.use_keypoints[2] = 1
    p.ego_motion = 3
    s.SetParams(p)


def run(slam, synth):
    configure(slam)
    seq = synth.Sequence(scene_seed=1234)
    sens = synth.sensor("tiny")
    frames = []
    for i in range(N_FRAMES):
        pts, _, stamp = seq.frame(sens, i)
        rc = slam.AddFrame(pts, stamp, i + 1)
        T, _ = slam.GetWorldTransform()
        st = slam.GetFrameStats()
        kps = [slam.GetKeypoints(k) for k in range(3)]
        frames.append({
            "rc": rc,
            "input_sha256": hashlib.sha256(pts.tobytes()).hexdigest(),
            "keypoints_sha256": [hashlib.sha256(k.tobytes()).hexdigest() for k in kps],
            "n_keypoints": [len(k) for k in kps],
            "n_matches": list(st.n_matches), "ego_n_matches": list(st.ego_n_matches),
            "loc_lm_iters": st.loc_lm_iters, "ego_lm_iters": st.ego_lm_iters, "is_keyframe": st.is_keyframe,
            "n_map": [len(slam.GetMap(k)) for k in range(3)],
            "pose": [float(v) for v in T.reshape(-1)],
        })
    return {"sensor": "tiny", "scene_seed": 1234, "frames": frames}


if __name__ == "__main__":
    from oracle import pyorc
    from tools.synth import synth
    out = run(pyorc.Oracle(nb_threads=1), synth)
    with open(os.path.join(HERE, "tiny_sequence.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", len(out["frames"]), "frames")

"""Probe (not product): queries each kNN first pass leaves to the warp pass, per search, on the bench workload (LSB_KNN_DEBUG=1)."""
import os, sys
if not os.environ.get("PROBE_QUIET"): os.environ["LSB_KNN_DEBUG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lidarslam_b200.slam import Slam
from tools.synth import synth
g = Slam(device=0)
p = g.GetParams(); p.use_keypoints[2] = 1; g.SetParams(p)
seq, sens = synth.Sequence(scene_seed=1234), synth.sensor("os128")
for i in range(12):
    pts, _, st = seq.frame(sens, i)
    if i == 11: sys.stderr.write("---- frame 11\n")
    g.AddFrame(pts, st, i + 1)

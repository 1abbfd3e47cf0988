import os, sys
sys.path.insert(0, "/root/repo")
os.environ["LSB_TRACE"] = "1"
from lidarslam_b200.slam import Slam
from tools.synth import synth
g = Slam(device=0)
p = g.GetParams(); p.use_keypoints[2] = 1; g.SetParams(p)
g.lib.lsb_set_profiling(g.h, 1)
seq, sens = synth.Sequence(scene_seed=1234), synth.sensor("os128")
fr = [seq.frame(sens, i) for i in range(16)]
dev = [g.DevUpload(f[0]) for f in fr]
for i in range(16):
    g.AddFramesDev([dev[i]], [len(fr[i][0])], [fr[i][2]], [i + 1])

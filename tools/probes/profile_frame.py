"""Probe (not product): one steady-state OS-128 frame (blobs, refined undistortion, key frame) between cudaProfilerStart /
cudaProfilerStop, for `ncu --profile-from-start off --set full ...`.   python tools/profile_frame.py [frames before] [sensor]"""
import ctypes
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from lidarslam_b200.slam import Slam
from tools.synth import synth

n_before = int(sys.argv[1]) if len(sys.argv) > 1 else 8
sensor = sys.argv[2] if len(sys.argv) > 2 else "os128"
rt = ctypes.CDLL("libcudart.so.12")
g = Slam(device=0)
p = g.GetParams()
p.use_keypoints[2] = 1
g.SetParams(p)
seq, sens = synth.Sequence(scene_seed=1234), synth.sensor(sensor)
frames = [seq.frame(sens, i) for i in range(n_before + 1)]
dev = [g.DevUpload(f[0]) for f in frames]
for i in range(n_before):
    g.AddFramesDev([dev[i]], [len(frames[i][0])], [frames[i][2]], [i + 1])
g.Synchronize()
rt.cudaProfilerStart()
g.AddFramesDev([dev[n_before]], [len(frames[n_before][0])], [frames[n_before][2]], [n_before + 1])
g.Synchronize()          # the frame's deferred map update is part of the profiled range
rt.cudaProfilerStop()
st = g.GetFrameStats()
print("profiled frame %d: keypoints %s sub-maps %s key frame %d launches %d %.3f ms" % (
    n_before, list(st.n_keypoints), list(st.n_submap), st.is_keyframe, st.kernel_launches, st.ms_total))

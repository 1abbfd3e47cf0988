"""Probe (not product): kernel timeline of steady-state OS-128 frames through CUPTI (torch.profiler sees every kernel
of the process, also those of liblidarslam_b200.so).  Writes a compact list - start, duration, stream, name - of the
kernels and memcpys of the profiled frames, for reading the idle gaps of the critical path.
    python tools/probes/frame_timeline.py [out.txt]"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402
from lidarslam_b200.slam import Slam  # noqa: E402
from tools.synth import synth  # noqa: E402


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/frame_timeline.txt"
    warm, n = 30, 3
    sens = synth.sensor("os128")
    sq = synth.Sequence(scene_seed=1234)
    frames = [sq.frame(sens, i) for i in range(warm + n)]
    npts = len(frames[0][0])
    torch.zeros(1, device="cuda")
    h = Slam(device=0)
    p = h.GetParams()
    p.use_keypoints[2] = 1
    p.undistortion = 2
    h.SetParams(p)
    dp = [h.DevUpload(f[0]) for f in frames]
    for i in range(warm):
        h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for i in range(warm, warm + n):
            h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])
        torch.cuda.synchronize()
    tmp = os.path.join(tempfile.gettempdir(), "trace.json")
    prof.export_chrome_trace(tmp)
    ev = json.load(open(tmp))["traceEvents"]
    rows = [e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset", "cuda_runtime", "cuda_driver")]
    rows.sort(key=lambda e: e["ts"])
    t0 = rows[0]["ts"] if rows else 0
    with open(out, "w") as f:
        for e in rows:
            a = e.get("args", {})
            f.write("%10.1f %8.1f %-12s s%-3s %s\n" % (e["ts"] - t0, e["dur"], e["cat"], a.get("stream", "-"), e["name"][:70]))
    print("wrote", out, len(rows), "events")


if __name__ == "__main__":
    main()

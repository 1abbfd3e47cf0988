"""Probe (not product): what the two pipelining steps buy on OS-128 - deferred map update, prefetched extraction - each alone and
together, K frames as one CUDA-event bracket (lsb_timer_begin / lsb_timer_end), no L2 flush; then a CUPTI kernel timeline of three
pipelined frames (start, duration, stream, name) for reading the critical chain map update -> sub-map -> ICP.
    python tools/probes/pipeline_probe.py [frames] [timeline.txt]"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import numpy as np  # noqa: E402
from lidarslam_b200.slam import Slam  # noqa: E402
from tools.synth import synth  # noqa: E402


def handle(defer):
    h = Slam(device=0)
    p = h.GetParams()
    p.use_keypoints[2] = 1
    p.undistortion = 2
    h.SetParams(p)
    h.SetPipelining(defer)
    return h


def run(frames, warm, K, defer, prefetch, tag):
    h = handle(defer)
    npts = len(frames[0][0])
    dp = [h.DevUpload(f[0]) for f in frames]
    for i in range(warm):
        h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])
    if prefetch:
        h.PrefetchFramesDev([dp[warm]], [npts], [frames[warm][2]], [warm + 1])
    h.TimerBegin()
    lat = []
    for i in range(warm, warm + K):
        if prefetch:
            h.PrefetchFramesDev([dp[i + 1]], [npts], [frames[i + 1][2]], [i + 2])
        t0 = time.perf_counter()
        h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])
        lat.append(1e3 * (time.perf_counter() - t0))
    ms = h.TimerEnd()
    print("%-28s %7.1f frames/s  %.3f ms/frame  call p50 %.3f p95 %.3f ms" % (tag, K / (ms * 1e-3), ms / K, np.percentile(lat, 50), np.percentile(lat, 95)), flush=True)
    h.close()


def timeline(frames, warm, n, out):
    import torch
    from torch.profiler import ProfilerActivity, profile
    torch.zeros(1, device="cuda")
    h = handle(True)
    npts = len(frames[0][0])
    dp = [h.DevUpload(f[0]) for f in frames]
    for i in range(warm):
        if i + 1 < warm:
            pass
        h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])
    h.PrefetchFramesDev([dp[warm]], [npts], [frames[warm][2]], [warm + 1])
    h.Synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for i in range(warm, warm + n):
            h.PrefetchFramesDev([dp[i + 1]], [npts], [frames[i + 1][2]], [i + 2])
            h.AddFramesDev([dp[i]], [npts], [frames[i][2]], [i + 1])
        h.Synchronize()
    tmp = os.path.join(tempfile.gettempdir(), "trace.json")
    prof.export_chrome_trace(tmp)
    ev = json.load(open(tmp))["traceEvents"]
    rows = [e for e in ev if e.get("ph") == "X" and e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    rows.sort(key=lambda e: e["ts"])
    t0 = rows[0]["ts"] if rows else 0
    with open(out, "w") as f:
        for e in rows:
            a = e.get("args", {})
            f.write("%10.1f %8.1f %-10s s%-3s %s\n" % (e["ts"] - t0, e["dur"], e["cat"], a.get("stream", "-"), e["name"][:60]))
    print("wrote", out, len(rows), "events")
    h.close()


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/pipeline_timeline.txt"
    warm = 8
    sens = synth.sensor("os128")
    sq = synth.Sequence(scene_seed=1234)
    frames = [sq.frame(sens, i) for i in range(warm + K + 1)]
    if os.environ.get("PROBE_ONLY") == "both":      # A/B runs: the pipelined loop only, several repetitions
        for rep in range(int(os.environ.get("PROBE_REPS", "4"))):
            run(frames, warm, K, True, True, "deferred + prefetch")
        return
    for rep in range(2):
        run(frames, warm, K, False, False, "synchronous")
        run(frames, warm, K, True, False, "deferred map update")
        run(frames, warm, K, False, True, "prefetch")
        run(frames, warm, K, True, True, "deferred + prefetch")
    if out != "-":
        timeline(frames, warm, 3, out)


if __name__ == "__main__":
    main()

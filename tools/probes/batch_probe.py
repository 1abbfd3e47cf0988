"""Probe (not product): aggregate frames/s of N OS-128 sequences sharing one GPU (bench.run_batch), for A/B runs of
environment knobs (CUDA_DEVICE_MAX_CONNECTIONS, LSB_LM_GRID_SHARE_MAX, LSB_HOST_LANES).
    python tools/probes/batch_probe.py [sequences] [frames]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import bench
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
k = int(sys.argv[2]) if len(sys.argv) > 2 else 20
dt, lat, _ = bench.run_batch(list(range(n)), "os128", 5, k, 0)
pooled = [v for l in lat for v in l]
print("%d sequences: %.1f frames/s  p50 %.2f ms  p95 %.2f ms   [CONN=%s SHARE_MAX=%s]" % (
    n, n * k / dt, np.percentile(pooled, 50), np.percentile(pooled, 95), os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), os.environ.get("LSB_LM_GRID_SHARE_MAX")))

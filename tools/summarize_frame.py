"""Per-kernel totals of one profiled frame from an `ncu --set full ... --page raw --csv` export, and the DRAM bytes per kNN search:
   python tools/summarize_frame.py gpurun_out/r02_frame_raw.csv profiles/r02_launch_summary.txt profiles/r02_knn_traffic.json [searches]"""
import collections
import csv
import json
import re
import sys


def main(raw, out_txt, out_json, searches=9):
    rows = list(csv.reader(open(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ni, dur = hdr.index("Kernel Name"), hdr.index("gpu__time_duration.sum")
    rd, wr = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    tscale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3}[units[dur]]
    bscale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

    def num(x):
        return float(x.replace(",", ""))
    agg = collections.OrderedDict()
    knn_bytes = knn_us = 0.0
    for r in data:
        n = re.sub(r"void |lsb::<unnamed>::|\(.*", "", r[ni])
        n = re.sub(r"cub::CUB_\w+::", "cub::", n)[:64]
        t = num(r[dur]) * tscale
        b = num(r[rd]) * bscale[units[rd]] + num(r[wr]) * bscale[units[wr]]
        a = agg.setdefault(n, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += t
        a[2] += b
        if n.startswith("k_knn"):
            knn_bytes += b
            knn_us += t
    tot = sum(a[1] for a in agg.values())
    lib = sum(a[1] for k, a in agg.items() if k.startswith("cub::"))
    nlib = sum(a[0] for k, a in agg.items() if k.startswith("cub::"))
    out = ["# per-kernel totals of ONE steady-state OS-128 frame (frame 8 of the sequence: a key frame that follows a key frame, so the sub-map selection and the map update are both in it; tools/probes/profile_frame.py)",
           "# from ncu --set full --clock-control none: cold caches, serialised, GPU at idle clocks between replays -> SHARES, not absolutes",
           "# columns: launches, total us, share of kernel time, DRAM MB (read + write), kernel",
           "# %d launches, %.1f us of kernel time; library (CUB) share %.1f %% of kernel time in %d launches" % (len(data), tot, 100 * lib / tot, nlib)]
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("%3d %9.1f us %5.1f %% %8.2f MB  %s" % (a[0], a[1], 100 * a[1] / tot, a[2] / 1e6, k))
    open(out_txt, "w").write("\n".join(out) + "\n")
    json.dump({"bytes_per_search": knn_bytes / searches, "searches": searches, "knn_kernel_us_per_frame_under_ncu": knn_us,
               "source": "dram__bytes_read.sum + dram__bytes_write.sum of every k_knn* launch of the profiled frame (3 keypoint types x 3 ICP "
                         "iterations), divided by the searches; profiles/r02_ncu_full_summary.txt holds the per-kernel values"},
              open(out_json, "w"), indent=1)
    print(out[3])


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 9)

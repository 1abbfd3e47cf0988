"""Reads a CUPTI kernel timeline written by tools/probes/pipeline_probe.py (start us, duration us, category, stream, name) and
prints, for the steady-state frame in the middle of it, the frame period, the critical chain of the blob stream from the end of
one ICP loop to the next one's first LM solve, the matching / solve phases, and the per-kernel totals of the period:
    python tools/summarize_timeline.py profiles/r02_pipeline_timeline.txt > profiles/r02_pipeline_phases.txt"""
import collections
import re
import sys


def short(name):
    name = re.sub(r"lsb::\(anonymous namespace\)::|void |cub::CUB_\w+::", "", name)
    return re.sub(r"\(.*", "", name)[:44]


def main(path):
    rows = []
    for line in open(path):
        p = line.split(None, 4)
        if len(p) == 5:
            rows.append((float(p[0]), float(p[1]), p[2], p[3], p[4].strip()))
    lm = [r for r in rows if "k_lm_solve" in r[4]]
    if len(lm) < 6:
        print("fewer than two complete frames in", path)
        return 1
    t0, t1 = lm[2][0] + lm[2][1], lm[5][0] + lm[5][1]        # end of frame 1's last solve .. end of frame 2's
    print("# %s: frame period %.1f us (end of an ICP loop to the end of the next)" % (path, t1 - t0))
    blob = [r for r in rows if "k_knn_stream<10" in r[4] and t0 < r[0] < t1]
    stream = blob[0][3] if blob else None
    ingest = [r for r in rows if "k_curvature" in r[4]]
    istream = ingest[0][3] if ingest else None
    print("# critical chain on the blob stream (%s) and the handle's stream, times relative to the end of the previous loop" % stream)
    first_lm = lm[3][0]
    for r in rows:
        if t0 - 20 <= r[0] <= first_lm + 1 and r[3] in (stream, lm[0][3]) and r[2] == "kernel":
            print("%8.1f %7.1f  %-4s %s" % (r[0] - t0, r[1], r[3], short(r[4])))
    print("# solves of the frame: start, duration")
    for r in lm[3:6]:
        print("%8.1f %7.1f" % (r[0] - t0, r[1]))
    if istream:
        ex = [r for r in rows if r[3] == istream and t0 <= r[0] < t1]
        if ex:
            print("# ingest stream (%s): extraction of the next input from %.1f to %.1f us, %.1f us of kernel / copy time" % (
                istream, ex[0][0] - t0, ex[-1][0] + ex[-1][1] - t0, sum(r[1] for r in ex)))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if t0 <= r[0] < t1:
            a = agg[short(r[4])]
            a[0] += 1
            a[1] += r[1]
    tot = sum(a[1] for a in agg.values())
    print("# per-kernel totals of the period: launches, us, share of the %.1f us of kernel + copy time (%.2f x the period)" % (tot, tot / (t1 - t0)))
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
        print("%3d %8.1f %5.1f %%  %s" % (a[0], a[1], 100 * a[1] / tot, k))
    return 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1]))

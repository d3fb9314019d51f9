#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list (B200_PROFILING.md recipe) of
`bench.py`: per-kernel device time of the LAST complete step (from one preprocess_fwd launch to the next).
Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.
    python profiles/summarize_launches.py gpurun_out/launches.csv > profiles/rNN_launches.txt"""
import collections
import csv
import re
import sys


def main(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    names = [r[4] for r in rows]
    starts = [i for i, n in enumerate(names) if "preprocess_fwd_kernel" in n]
    if len(starts) < 2:
        print("fewer than two steps in the launch list")
        return
    # the last segment that holds a whole step (forward AND backward): the bench ends with forward-only passes
    segs = [(a, b) for a, b in zip(starts[:-1], starts[1:]) if any("blend_bwd" in n for n in names[a:b])]
    if not segs:
        print("no forward+backward step in the launch list")
        return
    s, e = segs[-1]
    agg = collections.OrderedDict()
    total = 0.0
    for r in rows[s:e]:
        n = r[4]
        ours = "h3dgs::" in n or any(k in n for k in ("fill_spans_kernel", "zero_rows_kernel", "fill_zero_kernel"))
        key = re.sub(r"\(.*", "", n)
        key = key.replace("void ", "").replace("h3dgs::<unnamed>::", "h3dgs::")
        if not ours:
            key = "torch: " + key[:70]
        elif "h3dgs::" not in key:
            key = "h3dgs::" + key
        t = float(r[-1]) / 1e3
        total += t
        a = agg.setdefault(key, [0.0, 0, r[8], r[7]])
        a[0] += t
        a[1] += 1
    print(f"# launches {s}..{e - 1} of {path}: one resident step (fwd+bwd), {e - s} kernel launches, "
          f"{total:.1f} us of kernel time under ncu")
    print(f"# {'us':>9s} {'share':>6s} {'n':>3s}  {'grid':>16s} {'block':>12s}  kernel")
    ours_total = 0.0
    for k, (t, n, grid, block) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
        print(f"  {t:9.1f} {100 * t / total:5.1f}% {n:3d}  {grid:>16s} {block:>12s}  {k}")
        if k.startswith("h3dgs::"):
            ours_total += t
    print(f"# h3dgs kernels: {ours_total:.1f} us = {100 * ours_total / total:.1f}% of the step's kernel time")


if __name__ == "__main__":
    main(sys.argv[1])

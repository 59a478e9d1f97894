#!/usr/bin/env python
"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of `python bench.py ...`.

    tools/launch_list_summary.py gpurun_out/launches.csv > profiles/rNN_bench_launch_list_summary.txt

Prints (1) the launch sequence of ONE step of the bench's main loop (the 5th occurrence of
k7_match ... k7_gather: 3 warm-up + 2 timed steps precede the per-part loops) with each kernel's share
of the step, (2) all k7_* kernels of the process aggregated.  ncu times are cold-cache and serialised:
compare SHARES with the live CUDA-event numbers of bench.py, not absolutes."""
import csv
import re
import sys
from collections import OrderedDict


def short(name):
    name = re.sub(r"^void ", "", name)
    name = re.sub(r"\(.*$", "", name)
    name = name.replace("(bool)", "").replace("(int)", "")
    return name[:70]


def main():
    rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
    hdr = rows[0]
    ik, iv, ig = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    seq = [(short(r[ik]), float(r[iv]) / 1e3, r[ig]) for r in rows[1:]]
    starts = [i for i, (k, _, _) in enumerate(seq) if k.startswith("k7_score_pipe")]
    ends = [i for i, (k, _, _) in enumerate(seq) if k.startswith("k7_scale_multi")]
    print(f"{len(seq)} launches captured; score-kernel launches: {len(starts)}")
    if len(starts) >= 5 and len(ends) >= 5:
        # one step of the main loop as the profiler serialises it: from a score kernel (the step's first launch, on the
        # second stream) to the backward's scale kernel (its last)
        s = starts[4]
        e = next(i for i in ends if i > s)
        step = seq[s:e + 1]
        tot = sum(t for _, t, _ in step)
        print(f"\none step of the main loop (launches {s}..{e}): {len(step)} launches, {tot:.1f} us serialised")
        print(f"{'kernel':72s} {'grid':>16s} {'us':>9s} {'share':>7s}")
        for k, t, g in step:
            print(f"{k:72s} {g:>16s} {t:9.1f} {100 * t / tot:6.1f}%")
    agg = OrderedDict()
    for k, t, _ in seq:
        if not k.startswith("k7_"):
            continue
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += t
    tot = sum(v[1] for v in agg.values())
    print(f"\nall k7_* kernels of the process ({tot:.1f} us)")
    print(f"{'kernel':72s} {'launches':>8s} {'total us':>10s} {'avg us':>9s} {'share':>7s}")
    for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:72s} {n:8d} {t:10.1f} {t / n:9.1f} {100 * t / tot:6.1f}%")


if __name__ == "__main__":
    main()

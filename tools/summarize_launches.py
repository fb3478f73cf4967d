#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name."""
import collections
import csv
import sys


def main(path):
    with open(path) as f:
        lines = [l for l in f if l.startswith('"')]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        a = agg.setdefault(row["Kernel Name"], [0, 0.0])
        a[0] += 1
        a[1] += float(row["Metric Value"].replace(",", ""))
    tot = sum(a[1] for a in agg.values())
    # launches that are not part of a frame: bench.py's fp64 peak probe, torch fills of the L2 flush, the pair-count
    # diagnostic behind the roofline, and the untimed mesh producer (GPU Delaunay + trim: 24-NN, stars, split, duplicates)
    NOT_FRAME = ("k_dfma_probe", "k_kde_groups_count", "k_delaunay", "void k_delaunay", "void k_knn_thread<24>", "k_split", "k_dup")
    def is_aside(k):
        return "at::" in k or any(k.startswith(p) for p in NOT_FRAME)
    aside = sum(a[1] for k, a in agg.items() if is_aside(k))
    print(f"# {path}: {sum(a[0] for a in agg.values())} launches, {tot / 1e6:.3f} ms device time (cold-cache, serialised); "
          f"{(tot - aside) / 1e6:.3f} ms without the probe / L2 flush / diagnostic / mesh-producer launches (= 'share_of_frame' denominator)")
    print(f"{'kernel':70s} {'launches':>8s} {'total_ms':>12s} {'avg_us':>10s} {'share':>8s} {'share_of_frame':>15s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        frame = "" if is_aside(k) else f"{a[1] / (tot - aside):15.4f}"
        print(f"{k[:70]:70s} {a[0]:8d} {a[1] / 1e6:12.3f} {a[1] / a[0] / 1e3:10.1f} {a[1] / tot:8.4f} {frame}")


if __name__ == "__main__":
    main(sys.argv[1])

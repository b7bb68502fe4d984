#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel time of one
training step (the span between the last two input-layout kernels)."""
import collections
import csv
import re
import sys


def load(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    r = csv.reader(lines)
    hdr = next(r)
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    rows = []
    for row in r:
        if len(row) <= vi:
            continue
        v = float(row[vi].replace(",", ""))
        if row[ui] == "ns":
            v /= 1e3
        elif row[ui] == "ms":
            v *= 1e3
        rows.append((row[ki], v))
    return rows


def main():
    path = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/launches.csv"
    marker = sys.argv[2] if len(sys.argv) > 2 else "nchw_to_nhwc"
    rows = load(path)
    idx = [i for i, (k, _) in enumerate(rows) if marker in k]
    a, b = idx[-2], idx[-1]
    step = rows[a:b]
    tot = sum(v for _, v in step)
    print("# one step: %d kernels, %.1f us total (serialised, cold-cache; compare shares)" % (len(step), tot))
    for i, (k, v) in enumerate(step):
        name = re.sub(r"\(.*", "", k)
        name = re.sub(r"^void ", "", name)[:64]
        print("%3d %-66s %9.1f  %5.1f%%" % (i, name, v, 100 * v / tot))
    agg = collections.defaultdict(float)
    for k, v in step:
        name = re.sub(r"\(.*", "", k)
        agg[re.sub(r"^void ", "", name)[:64]] += v
    print("# by kernel")
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1]):
        print("    %-66s %9.1f  %5.1f%%" % (k, v, 100 * v / tot))


if __name__ == "__main__":
    main()

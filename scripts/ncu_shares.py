#!/usr/bin/env python
"""Per-kernel share of one bench step from an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`).
Per-launch times under ncu are cold-cache and serialised: the SHARES are what is comparable with the CUDA-event step time."""
import collections
import csv
import sys


def main(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 8]
    hi = [i for i, r in enumerate(rows) if "Kernel Name" in r][0]
    h = rows[hi]
    kn, mn, mv, mu = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
    scale = {"ns": 1e-6, "nsecond": 1e-6, "us": 1e-3, "usecond": 1e-3, "ms": 1.0, "msecond": 1.0, "s": 1e3, "second": 1e3}
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[hi + 1:]:
        if r[mn] == "gpu__time_duration.sum":
            name = r[kn].split("(")[0][:70]
            agg[name][0] += 1
            agg[name][1] += float(r[mv].replace(",", "")) * scale.get(r[mu], 1.0)
    tot = sum(v[1] for v in agg.values())
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{v[1]:10.3f} ms {100 * v[1] / tot:5.1f}% x{v[0]:4d}  {k}")
    print(f"{tot:10.3f} ms total under ncu (counted pass + timed step + probes)")


if __name__ == "__main__":
    main(sys.argv[1])

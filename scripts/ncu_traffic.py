#!/usr/bin/env python
"""Sum DRAM traffic of the traversal kernels of ONE timed bench step from an ncu csv
(`ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv` on `bench.py --steps 1 --warmup 0 --mode M`)
and record it under profiles/traffic.json, keyed "workload/pipeline" — bench.py reports it as roofline.traffic."""
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main(path, key):
    rows = [r for r in csv.reader(open(path)) if len(r) > 8]
    h = rows[0]
    ki, mi, vi, ui, idi = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit"), h.index("ID")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9}
    per = {}
    for r in rows[1:]:
        d = per.setdefault(r[idi], {"name": r[ki]})
        d[r[mi]] = float(r[vi].replace(",", "")) * scale.get(r[ui], 1)
    trav = re.compile(r"(wf_trace|wf_shadow|wf_primary|render_mega|trace_rays)")
    # the timed step launches the COUNT=false instantiations (template argument 3 of <LAYOUT, ORDER, COUNT[, SHADOW0]>, 2 of the octet
    # kernels' <ORDER, COUNT>); the counted pass the others
    timed = [d for d in per.values() if trav.search(d["name"]) and re.search(r"<(\d+, \d+, (0|false)(, \w+)?|\d+, (0|false))>\(", d["name"].replace("(bool)", "").replace("(int)", ""))]
    out = {"dram_bytes_per_step": sum(d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0) for d in timed),
           "traversal_seconds_under_ncu": sum(d.get("gpu__time_duration.sum", 0) for d in timed), "launches": len(timed),
           "kernels": sorted({d["name"].split("(")[0] for d in timed}), "source": os.path.basename(path)}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    allv = json.load(open(tp)) if os.path.exists(tp) else {}
    if len(sys.argv) > 3:  # captured at a reduced spp: bench.py scales linearly
        out["spp"] = int(sys.argv[3])
        out["note"] = f"captured at --spp {sys.argv[3]} (one pass); scales linearly with spp"
    allv[key] = out
    json.dump(allv, open(tp, "w"), indent=1, sort_keys=True)
    print(key, out)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])

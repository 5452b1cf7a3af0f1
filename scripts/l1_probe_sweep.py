"""Sector rate of an incoherent per-lane gather of node-sized records as a function of the table size: L1-resident tables
show the L1 pipe's own roof, tables above the 126 MB L2 the DRAM-served floor.  python scripts/l1_probe_sweep.py  (GPU box)"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import minipath_b200 as mp  # noqa: E402

ctx = mp.Context(0)
rows = []
for sectors in (5, 6):
    for n in (512, 1024, 4096, 16384, 65536, 262144, 1 << 20, 1 << 22, 1 << 23):
        g = ctx.probe_l1_gather(n, sectors)
        rows.append({"records": n, "sectors_per_record": sectors, "table_bytes": n * sectors * 32, "gsectors_per_s": g, "tb_per_s": g * 32e-3})
        print(f"{sectors} sectors x {n:8d} records = {n * sectors * 32 / 2**20:9.2f} MiB: {g:8.1f} Gsectors/s = {g * 32e-3:6.2f} TB/s", flush=True)
print(json.dumps(rows))

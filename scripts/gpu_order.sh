#!/bin/bash
# reference vs near-first visiting order on the two headline workloads (same pipeline)
set -u
mkdir -p gpurun_out
for O in reference near_first; do
  A="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool-s0 --order $O"
  timeout 300 python bench.py $A > gpurun_out/order_default_$O.json 2>gpurun_out/order_default_$O.err
  timeout 300 python bench.py $A --workload atrium-1080p --spp 32 > gpurun_out/order_atrium_$O.json 2>gpurun_out/order_atrium_$O.err
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/order_*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); r=d.get("roofline") or {}
        print(f, round(d["value"]), round(d["ms_per_step"],3), r.get("frac"), r.get("per_ray"))
    except Exception as e: print(f,"ERR",e)
PY

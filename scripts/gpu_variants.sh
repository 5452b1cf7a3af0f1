#!/bin/bash
# A/B of library builds (build/libminipath_b200_*.so) on the stand-in, pool pipelines only.
set -u
mkdir -p gpurun_out
for v in default "$@"; do
  if [ $v = default ]; then unset MINIPATH_B200_LIB; else export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so; fi
  for m in wavefront-pool wavefront-hybrid-pool; do
    timeout 600 python bench.py --workload atrium-1080p --spp 8 --no-cpu-baseline --no-e2e --steps 3 --mode $m > gpurun_out/var_${v}_$m.json 2> gpurun_out/var_${v}_$m.err
  done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/var_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],2))
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-300:])
PY

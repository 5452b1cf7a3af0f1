#!/bin/bash
# A/B of library builds (build/libminipath_b200_*.so) on the coherent-ray kernels: default workload (fused depth-0 kernel) and teapot-bench (megakernel).
set -u
mkdir -p gpurun_out
rm -f gpurun_out/pv*
for rep in 1; do
for v in default "$@"; do
  if [ $v = default ]; then unset MINIPATH_B200_LIB; else export MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$v.so; fi
  timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 5 --mode wavefront-hybrid-pool > gpurun_out/pv${rep}_${v}_default.json 2> gpurun_out/pv${rep}_${v}_default.err
  timeout 600 python bench.py --workload teapot-bench --no-cpu-baseline --no-e2e --steps 5 --mode megakernel > gpurun_out/pv${rep}_${v}_tb.json 2> gpurun_out/pv${rep}_${v}_tb.err
  timeout 600 python bench.py --workload cube-1080p --no-cpu-baseline --no-e2e --steps 5 --mode megakernel > gpurun_out/pv${rep}_${v}_cube.json 2> gpurun_out/pv${rep}_${v}_cube.err
done
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/pv*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],3))
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-300:])
PY

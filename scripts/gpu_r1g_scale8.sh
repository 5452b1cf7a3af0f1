#!/bin/bash
# 8-GPU lines of the session-g build: default workload (spp split, weak) and the stand-in with tile sharding (strong), plus their N=1 partners.
set -u
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 300 $T --master-port 29541 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/scale_default_n8.json 2> gpurun_out/scale_default_n8.err; echo rc=$?
timeout 300 $T --master-port 29542 bench.py --gpus 8 --steps 3 --warmup 3 --workload atrium-1080p --spp 32 --shard tiles --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool-s0 > gpurun_out/scale_atrium_tiles_n8.json 2> gpurun_out/scale_atrium_tiles_n8.err; echo rc=$?
timeout 300 python bench.py --steps 3 --warmup 3 --workload atrium-1080p --spp 32 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool-s0 > gpurun_out/scale_atrium_n1.json 2> gpurun_out/scale_atrium_n1.err; echo rc=$?
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/scale_*n[18].json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["n_gpus"], round(d["value"]), round(d["ms_per_step"],3), d["scaling"], d["config"].get("sharding"))
    except Exception as e: print(f,"ERR",e)
PY

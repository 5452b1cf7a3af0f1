#!/bin/bash
set -u
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 python bench.py --gpus 1 --workload atrium-1080p --spp 32 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/scale2_atrium_n1.json 2> gpurun_out/scale2_atrium_n1.err
timeout 300 $TR --master-port 29532 bench.py --gpus $N --workload atrium-1080p --spp 32 --steps 2 --warmup 3 --shard tiles --no-e2e > gpurun_out/scale2_atrium_tiles_n$N.json 2> gpurun_out/scale2_atrium_tiles_n$N.err
timeout 300 $TR --master-port 29533 bench.py --gpus $N --workload atrium-1080p --spp 32 --steps 2 --warmup 3 > gpurun_out/scale2_atrium_spp_n$N.json 2> gpurun_out/scale2_atrium_spp_n$N.err
timeout 300 $TR --master-port 29531 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/scale2_default_n$N.json 2> gpurun_out/scale2_default_n$N.err
timeout 300 $TR --master-port 29534 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/scale2_ref_n$N.json 2> gpurun_out/scale2_ref_n$N.err
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/scale2_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["n_gpus"], d["scaling"], round(d["value"]), round(d["ms_per_step"],2), d["config"].get("pipeline"), d["e2e"] and round(d["e2e"]["value"]))
    except Exception as e: print(f,"ERR",open(f.replace(".json",".err")).read()[-500:])
PY

#!/bin/bash
# After the occupancy change of the coherent-ray kernels: parity suite, smoke, bench lines of the affected workloads, ncu captures of both kernels.
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 900 python bench.py --workload teapot-bench > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
timeout 900 python bench.py --workload cube-1080p --no-cpu-baseline > gpurun_out/bench_cube.json 2> gpurun_out/bench_cube.err
timeout 900 python bench.py --workload atrium-1080p --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_atrium.json 2> gpurun_out/bench_atrium.err
A="--steps 1 --warmup 0 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"wf_primary" -s 2 -c 1 -o gpurun_out/prof_primary_after python bench.py $A --mode wavefront-hybrid-pool > gpurun_out/ncu_primary.log 2>&1; echo "rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"render_mega" -s 2 -c 1 -o gpurun_out/prof_mega_after python bench.py $A --workload teapot-bench --mode megakernel > gpurun_out/ncu_mega.log 2>&1; echo "rc=$?"
tail -2 gpurun_out/pytest_gpu.log; tail -1 gpurun_out/smoke.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/bench_*.json")):
    try: d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d["value"]), round(d["ms_per_step"],3), d.get("e2e",{}).get("value"), d["roofline"]["frac"], d["config"].get("pipeline"))
    except Exception as e: print(f,"ERR",e)
PY

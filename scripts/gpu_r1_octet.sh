#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 600 python bench.py --no-cpu-baseline --order near_first --no-e2e > gpurun_out/bench_near.json 2> gpurun_out/bench_near.err
timeout 600 python bench.py --workload teapot-bench --no-cpu-baseline > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
timeout 900 python bench.py --workload atrium-1080p --spp 16 --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_atrium16.json 2> gpurun_out/bench_atrium16.err
A="--steps 1 --warmup 0 --spp 2 --no-cpu-baseline --no-e2e --mode wavefront-pool"
python bench.py $A > gpurun_out/plain_wfo.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:wf_trace_pool -s 12 -c 3 -o gpurun_out/prof_wftrace_pool python bench.py $A > gpurun_out/ncu_wfo.log 2>&1
ls -la gpurun_out

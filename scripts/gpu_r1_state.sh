#!/bin/bash
# Re-establish the measured state of the tree: parity tests, smoke, default bench, stand-in bench, launch list + full capture of the pool kernel.
set -u
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 900 python bench.py --workload atrium-1080p --spp 16 --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_atrium16.json 2> gpurun_out/bench_atrium16.err
timeout 600 python bench.py --workload teapot-bench --no-cpu-baseline > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
A="--workload atrium-1080p --steps 1 --warmup 0 --spp 2 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool"
timeout 300 python bench.py $A > gpurun_out/plain_pool.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_atrium_hybrid_pool.csv python bench.py $A > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_trace_pool -s 6 -c 3 -o gpurun_out/prof_atrium_pool python bench.py $A > gpurun_out/ncu_pool.log 2>&1
ls -la gpurun_out

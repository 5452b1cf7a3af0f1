#!/bin/bash
# Round-end evidence: parity suite, smoke, default bench (both arms), the other workloads, launch list + full capture of the default command.
set -u
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
timeout 900 python bench.py --workload teapot-bench > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
timeout 900 python bench.py --workload cube-1080p --no-cpu-baseline > gpurun_out/bench_cube.json 2> gpurun_out/bench_cube.err
timeout 900 python bench.py --workload atrium-1080p --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_atrium.json 2> gpurun_out/bench_atrium.err
timeout 900 python bench.py --workload atrium-4k --spp 16 --steps 2 --no-cpu-baseline --no-e2e > gpurun_out/bench_atrium4k_spp16.json 2> gpurun_out/bench_atrium4k_spp16.err
A="--steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_default.csv python bench.py $A > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wf_(trace|shadow)_pool|wf_trace_kernel" -s 60 -c 3 -o gpurun_out/prof_default python bench.py $A > gpurun_out/ncu_default.log 2>&1
ls -la gpurun_out

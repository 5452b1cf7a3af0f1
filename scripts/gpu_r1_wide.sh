#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python scripts/soup_sweep.py --sizes 1e6,1e7 > gpurun_out/soup_sweep.jsonl 2> gpurun_out/soup_sweep.err
timeout 600 python bench.py --workload atrium-4k --spp 4 --no-cpu-baseline --no-e2e --steps 2 > gpurun_out/bench_atrium4k_spp4.json 2> gpurun_out/bench_atrium4k_spp4.err
timeout 600 python bench.py --workload cube-1080p --no-cpu-baseline --steps 3 > gpurun_out/bench_cube.json 2> gpurun_out/bench_cube.err
tail -n 3 gpurun_out/pytest_gpu.log; cut -c1-330 gpurun_out/soup_sweep.jsonl; tail -n 3 gpurun_out/soup_sweep.err

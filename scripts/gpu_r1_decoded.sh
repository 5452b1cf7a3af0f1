#!/bin/bash
# Decoded node layout: parity, A/B against the quantised records, L1 access-pattern microbenchmark.
set -u
mkdir -p gpurun_out
timeout 120 scripts/ubench/l1_patterns > gpurun_out/ubench_l1_patterns.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for L in decoded quantised; do
timeout 600 python bench.py --layout $L --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_default_$L.json 2> gpurun_out/bench_default_$L.err
timeout 900 python bench.py --layout $L --workload atrium-1080p --spp 16 --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_atrium16_$L.json 2> gpurun_out/bench_atrium16_$L.err
timeout 600 python bench.py --layout $L --workload teapot-bench --no-cpu-baseline --no-e2e > gpurun_out/bench_teapot_bench_$L.json 2> gpurun_out/bench_teapot_bench_$L.err
done
ls -la gpurun_out

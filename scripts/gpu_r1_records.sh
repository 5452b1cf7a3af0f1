#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for L in auto decoded quantised; do
timeout 900 python bench.py --layout $L --workload atrium-1080p --spp 16 --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_atrium16_$L.json 2> gpurun_out/bench_atrium16_$L.err
done
timeout 600 python bench.py --no-cpu-baseline --no-e2e --steps 3 > gpurun_out/bench_default_auto.json 2> gpurun_out/bench_default_auto.err
ls -la gpurun_out

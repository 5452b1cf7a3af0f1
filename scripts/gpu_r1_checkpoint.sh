#!/bin/bash
# Checkpoint: full parity suite, smoke, launch list and full capture of the pool kernels on the stand-in.
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
bash scripts/gpu_launchlist.sh atrium-1080p wavefront-hybrid-pool
A="--workload atrium-1080p --steps 1 --warmup 0 --spp 2 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wf_(trace|shadow)_pool" -s 20 -c 2 -o gpurun_out/prof_atrium_pool python bench.py $A > gpurun_out/ncu_pool.log 2>&1
ls -la gpurun_out

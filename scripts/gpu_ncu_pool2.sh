#!/bin/bash
set -u
mkdir -p gpurun_out
A="--workload atrium-1080p --steps 1 --warmup 0 --spp 8 --no-cpu-baseline --no-e2e --mode wavefront-hybrid-pool"
timeout 300 python bench.py $A > gpurun_out/plain_pool.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"wf_(trace|shadow)_pool" -s 10 -c 2 -o gpurun_out/prof_atrium_pool_big python bench.py $A > gpurun_out/ncu_pool.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_atrium_big.csv python bench.py $A > gpurun_out/ncu_list.log 2>&1
ls -la gpurun_out | tail -5

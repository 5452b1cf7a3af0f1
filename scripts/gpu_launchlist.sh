#!/bin/bash
set -u
mkdir -p gpurun_out
W=${1:-atrium-1080p}; M=${2:-wavefront-hybrid-pool}
A="--workload $W --steps 1 --warmup 0 --spp 2 --no-cpu-baseline --no-e2e --mode $M"
timeout 300 python bench.py $A > gpurun_out/plain_list.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_${W}_${M}.csv python bench.py $A > gpurun_out/ncu_list.log 2>&1
tail -c 300 gpurun_out/ncu_list.log

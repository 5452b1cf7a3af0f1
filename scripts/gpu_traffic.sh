#!/bin/bash
set -u
mkdir -p gpurun_out
run() { # workload mode spp
  A="--workload $1 --steps 1 --warmup 0 --no-cpu-baseline --no-e2e --mode $2"
  [ -n "${3:-}" ] && A="$A --spp $3"
  timeout 300 python bench.py $A > gpurun_out/plain_$1_$2.log 2>&1 && \
  timeout 900 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/traffic_$1_$2.csv python bench.py $A > gpurun_out/ncu_traffic_$1_$2.log 2>&1
}
run teapot-1080p wavefront-hybrid-pool
run teapot-1080p megakernel
run teapot-bench megakernel
run atrium-1080p wavefront-hybrid-pool 16
run cube-1080p megakernel
ls -la gpurun_out | grep traffic

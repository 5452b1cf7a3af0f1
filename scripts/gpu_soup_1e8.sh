#!/bin/bash
# 1e8-triangle soup point of the traversal sweep (BASELINE configs[4]); the fork-join host builder makes the build feasible.
set -u
mkdir -p gpurun_out
nproc > gpurun_out/nproc.txt; free -g > gpurun_out/free.txt
( time timeout 1500 python scripts/soup_sweep.py --sizes 1e8 --reps 2 ) > gpurun_out/soup_1e8.jsonl 2> gpurun_out/soup_1e8.err; echo "rc=$?" >> gpurun_out/soup_1e8.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_subset.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_subset.log
tail -5 gpurun_out/soup_1e8.err; tail -3 gpurun_out/pytest_gpu_subset.log

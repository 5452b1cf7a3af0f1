#!/bin/bash
# lane-owned pool slots (POOL_LANE_OWNED=1 build): parity of every context that uses the pool kernels, then A/B against the default build
set -u
mkdir -p gpurun_out
V=${1:-owned}
MINIPATH_B200_LIB=$PWD/build/libminipath_b200_$V.so timeout 600 python -m pytest tests -m gpu -x -q -k "pool" > gpurun_out/pytest_$V.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_$V.log
tail -5 gpurun_out/pytest_$V.log
bash scripts/gpu_variants5.sh "$@"

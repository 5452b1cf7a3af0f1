#!/bin/bash
# block-compacted depth-0 sun rays (WF_PRIMARY_S0_BLOCK=1 build): parity of the S0 contexts, then A/B against the default build
set -u
mkdir -p gpurun_out
MINIPATH_B200_LIB=$PWD/build/libminipath_b200_s0blk.so timeout 600 python -m pytest tests -m gpu -x -q -k "hybrid_pool_s0 or auto_pipeline" > gpurun_out/pytest_s0blk.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_s0blk.log
tail -3 gpurun_out/pytest_s0blk.log
bash scripts/gpu_variants5.sh s0blk

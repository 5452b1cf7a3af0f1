#!/bin/bash
# Full ncu capture of the coherent-ray kernels: the fused depth-0 wavefront kernel (default workload) and the megakernel (teapot-bench).
set -u
mkdir -p gpurun_out
A="--steps 1 --warmup 0 --no-cpu-baseline --no-e2e"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"wf_primary" -s 1 -c 1 -o gpurun_out/prof_primary python bench.py $A --mode wavefront-hybrid-pool > gpurun_out/ncu_primary.log 2>&1; echo "rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"render_mega" -s 1 -c 1 -o gpurun_out/prof_mega python bench.py $A --workload teapot-bench --mode megakernel > gpurun_out/ncu_mega.log 2>&1; echo "rc=$?"
ls -la gpurun_out/*.ncu-rep

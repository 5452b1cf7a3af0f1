#!/bin/bash
# Round-1 first GPU pass: parity tests, smoke, bench (A/B of pipelines and orders), ncu launch list + full captures.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "rc=$?" >> gpurun_out/bench_default.err
python bench.py --impl reference > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
for m in megakernel wavefront; do for o in reference near_first; do
  python bench.py --mode $m --order $o --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/bench_${m}_${o}.json 2> gpurun_out/bench_${m}_${o}.err
done; done
python bench.py --workload teapot-bench --steps 5 --warmup 3 > gpurun_out/bench_teapot_bench.json 2> gpurun_out/bench_teapot_bench.err
# ncu: launch list of a short wavefront run, then full captures of the two traversal kernels
A="--steps 1 --warmup 0 --spp 2 --no-cpu-baseline --no-e2e"
python bench.py $A --mode wavefront > gpurun_out/plain_wf.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_wavefront.csv python bench.py $A --mode wavefront > gpurun_out/ncu_wf_list.log 2>&1
python bench.py $A --mode megakernel > gpurun_out/plain_mega.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_mega -s 1 -c 1 -o gpurun_out/prof_mega python bench.py $A --mode megakernel > gpurun_out/ncu_mega.log 2>&1
python bench.py $A --mode wavefront > gpurun_out/plain_wf2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:wf_trace -s 12 -c 3 -o gpurun_out/prof_wftrace python bench.py $A --mode wavefront > gpurun_out/ncu_wftrace.log 2>&1
ls -la gpurun_out

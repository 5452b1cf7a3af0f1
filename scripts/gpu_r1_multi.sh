#!/bin/bash
# N-GPU bench exactly as the driver launches it (torchrun, one rank per GPU, NCCL film reduce).
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_L.txt
for n in 1 $N; do
  if [ $n -eq 1 ]; then
    python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $n --steps 3 --warmup 3 --shard tiles --no-e2e > gpurun_out/scale_tiles_n$n.json 2> gpurun_out/scale_tiles_n$n.err
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus $n --steps 2 --warmup 1 > gpurun_out/scale_ref_n$n.json 2> gpurun_out/scale_ref_n$n.err
  fi
done
tail -c 600 gpurun_out/scale_n$N.err
ls -la gpurun_out | grep scale

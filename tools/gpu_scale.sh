#!/bin/bash
# Multi-GPU bench of the contract command on N GPUs of one box:
#   gpurun --gpus N --timeout 900 -- 'bash tools/gpu_scale.sh N TAG'
n=${1:-2}; tag=${2:-run}
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29531 \
  bench.py --gpus $n --steps 2 --warmup 3 > gpurun_out/${tag}_bench_n${n}.json 2> gpurun_out/${tag}_bench_n${n}.err
tail -c 1500 gpurun_out/${tag}_bench_n${n}.json; tail -5 gpurun_out/${tag}_bench_n${n}.err

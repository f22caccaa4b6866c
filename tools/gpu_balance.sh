#!/bin/bash
# GPU-box call (N GPUs): balance probe (identical poses + per-rank times), then the contract bench on N GPUs.
n=${1:-2}; tag=${2:-bal}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 tools/balance_probe.py > gpurun_out/${tag}_probe_n${n}.txt 2> gpurun_out/${tag}_probe_n${n}.err
grep -h "rank" gpurun_out/${tag}_probe_n${n}.txt | sort | tail -40; tail -3 gpurun_out/${tag}_probe_n${n}.err
bash tools/gpu_scale.sh $n $tag

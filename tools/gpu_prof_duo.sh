#!/bin/bash
# ncu --set full capture of the two-row K3 kernel (256 rows = 128 CTAs, 10 iterations) + phase shares
tag=${1:-r2b}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pose_duo -s 1 -c 1 -f -o gpurun_out/${tag}_k3duo python tools/ncu_k3.py 4 10 > gpurun_out/${tag}_ncu_k3duo.log 2>&1
python tools/ncu_phases.py gpurun_out/${tag}_k3duo.ncu-rep > gpurun_out/${tag}_k3duo_phases.txt 2>&1; cat gpurun_out/${tag}_k3duo_phases.txt

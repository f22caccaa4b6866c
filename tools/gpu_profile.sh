#!/bin/bash
# ncu evidence of the contract bench: launch list of `python bench.py --no-cpu-baseline`, one --set full capture
# of K3 (tools/ncu_k3.py 2 10) with its phase shares.
tag=${1:-prof}
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --no-cpu-baseline > gpurun_out/${tag}_ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pose_sym -s 1 -c 1 -f -o gpurun_out/${tag}_k3 python tools/ncu_k3.py 2 10 > gpurun_out/${tag}_ncu_k3.log 2>&1
python tools/ncu_phases.py gpurun_out/${tag}_k3.ncu-rep > gpurun_out/${tag}_k3_phases.txt 2>&1; cat gpurun_out/${tag}_k3_phases.txt

#!/bin/bash
# Round-2 ncu evidence: launch list of the contract bench (python bench.py --no-cpu-baseline), one --set full
# capture of the two-row K3 kernel (256 rows = 128 CTAs x 10 iterations) with phase shares and raw metrics.
tag=${1:-r2p}
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --no-cpu-baseline > gpurun_out/${tag}_ncu_bench.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pose_duo -s 1 -c 1 -f -o gpurun_out/${tag}_k3duo python tools/ncu_k3.py 4 10 > gpurun_out/${tag}_ncu_k3duo.log 2>&1
python tools/ncu_phases.py gpurun_out/${tag}_k3duo.ncu-rep > gpurun_out/${tag}_k3duo_phases.txt 2>&1; cat gpurun_out/${tag}_k3duo_phases.txt | head -30
ncu -i gpurun_out/${tag}_k3duo.ncu-rep --page raw --csv > gpurun_out/${tag}_k3duo_raw.csv 2>/dev/null
wc -l gpurun_out/${tag}_launches.csv

#!/bin/bash
# One GPU-box call for a K3 change: the K3 tests, variant timings (16 = old recovery, 17, 12 = default),
# the contract bench, its ncu launch list, one ncu --set full capture of K3 + phase shares.
#   gpurun --timeout 1500 -- 'bash tools/gpu_k3_recovery.sh TAG'
tag=${1:-run}
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_pose_gpu.py tests/test_train_utils_gpu.py tests/test_register_gpu.py -m gpu -x -q) > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
python tools/microbench.py pose 37 40 16,17,12,16,12 > gpurun_out/${tag}_variants.jsonl 2> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_variants.jsonl
python tools/microbench.py solve 37 40 > gpurun_out/${tag}_solve.jsonl 2>> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_solve.jsonl
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 300 gpurun_out/${tag}_bench.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv python bench.py --no-cpu-baseline > gpurun_out/${tag}_ncu_bench.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pose_sym -s 1 -c 1 -f -o gpurun_out/${tag}_k3 python tools/ncu_k3.py 2 10 > gpurun_out/${tag}_ncu_k3.log 2>&1
python tools/ncu_phases.py gpurun_out/${tag}_k3.ncu-rep > gpurun_out/${tag}_k3_phases.txt 2>&1; cat gpurun_out/${tag}_k3_phases.txt

#!/bin/bash
# One GPU-box call for the one-metric K3 kernel (train_utils.solve shape): tests + timings.
tag=${1:-run}
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_pose_gpu.py tests/test_train_utils_gpu.py tests/test_register_gpu.py tests/test_capi_status_gpu.py -m gpu -x -q) > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
python tools/microbench.py solve 37 40 > gpurun_out/${tag}_solve.jsonl 2> gpurun_out/${tag}_solve.err; cat gpurun_out/${tag}_solve.jsonl
python tools/microbench.py solve 74 40 >> gpurun_out/${tag}_solve.jsonl 2>> gpurun_out/${tag}_solve.err; tail -2 gpurun_out/${tag}_solve.jsonl
python tools/microbench.py pose 37 40 16,12 > gpurun_out/${tag}_variants.jsonl 2>> gpurun_out/${tag}_solve.err; cat gpurun_out/${tag}_variants.jsonl

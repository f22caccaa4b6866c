#!/bin/bash
# Quick GPU-box call after a K3 change: the K3 tests and the contract bench.
tag=${1:-quick}
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_pose_gpu.py tests/test_train_utils_gpu.py tests/test_register_gpu.py -m gpu -x -q) > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 300 gpurun_out/${tag}_bench.json

#!/bin/bash
tag=${1:-r2c}
mkdir -p gpurun_out
(time timeout 900 python -m pytest tests/test_small_parity_gpu.py tests/test_full_parity_gpu.py tests/test_distribution_parity_gpu.py tests/test_chamfer_gpu.py -m gpu -q -s -k "not test_edge_cases") > gpurun_out/${tag}_tests.log 2>&1; tail -12 gpurun_out/${tag}_tests.log
timeout 300 python tools/hard_probe.py chair 0 150 > gpurun_out/${tag}_hard.jsonl 2> gpurun_out/${tag}_hard.err
timeout 300 python tools/hard_probe.py ellipsoids 0 150 >> gpurun_out/${tag}_hard.jsonl 2>> gpurun_out/${tag}_hard.err
cat gpurun_out/${tag}_hard.jsonl; tail -3 gpurun_out/${tag}_hard.err

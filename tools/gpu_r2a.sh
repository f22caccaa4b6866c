#!/bin/bash
# r2 first GPU call: the two-row K3 kernel against the one-row kernel (bits), the full-size parity
# reports, and timings of both kernels side by side.
tag=${1:-r2a}
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_pose_gpu.py -m gpu -x -q -k "duo") > gpurun_out/${tag}_duo.log 2>&1; tail -5 gpurun_out/${tag}_duo.log
(time timeout 900 python -m pytest tests/test_full_parity_gpu.py tests/test_distribution_parity_gpu.py -m gpu -q -s) > gpurun_out/${tag}_parity.log 2>&1; tail -15 gpurun_out/${tag}_parity.log
timeout 600 python tools/microbench.py pose 37 40 13,20,13,20 > gpurun_out/${tag}_variants.jsonl 2> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_variants.jsonl
HOUV_POSE_VARIANT=20 timeout 600 python tools/microbench.py pose 74 200 > gpurun_out/${tag}_pose200.jsonl 2>> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_pose200.jsonl

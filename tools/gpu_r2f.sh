#!/bin/bash
tag=${1:-r2f}
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_pose_gpu.py -m gpu -x -q -k "duo") > gpurun_out/${tag}_duo.log 2>&1; tail -5 gpurun_out/${tag}_duo.log
timeout 600 python tools/microbench.py pose 37 40 13,20,21,20,21 > gpurun_out/${tag}_variants.jsonl 2> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_variants.jsonl; tail -3 gpurun_out/${tag}_variants.err

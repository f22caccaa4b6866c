#!/bin/bash
# One GPU-box call: GPU parity tests, both bench arms, smoke().  Usage (from the repo root):
#   gpurun --timeout 1700 -- 'bash tools/gpu_check.sh TAG'
tag=${1:-run}
mkdir -p gpurun_out
(time timeout 1000 python -m pytest tests -m gpu -x -q) > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
(time python bench.py) > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 600 gpurun_out/${tag}_bench.json
(time python bench.py --impl reference) > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; tail -c 400 gpurun_out/${tag}_bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log

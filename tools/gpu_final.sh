#!/bin/bash
# Round-end check on one GPU box: the whole GPU suite, both bench arms, smoke().
# (compute-sanitizer is closed on this pool; tools/sanitize_workload.py is the workload for it elsewhere.)
tag=${1:-final}
mkdir -p gpurun_out
(time timeout 1000 python -m pytest tests -m gpu -x -q) > gpurun_out/${tag}_pytest.log 2>&1; tail -3 gpurun_out/${tag}_pytest.log
python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; tail -c 300 gpurun_out/${tag}_bench.json
python bench.py --impl reference > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; tail -c 200 gpurun_out/${tag}_bench_ref.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; tail -2 gpurun_out/${tag}_smoke.log

#!/bin/bash
# r2: K1/K2 tests, then the artefacts kept under profiles/ (full-size parity table, distributional parity,
# K3 variant timings, K2 variant timings, contract bench N=1 + reference arm)
tag=${1:-r2l}
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_chamfer_gpu.py tests/test_reference_ext_gpu.py tests/test_calc_cd_gpu.py -m gpu -x -q) > gpurun_out/${tag}_k12.log 2>&1; tail -3 gpurun_out/${tag}_k12.log
timeout 600 python tools/full_parity.py > gpurun_out/${tag}_full_parity.json 2> gpurun_out/${tag}_full_parity.err; tail -2 gpurun_out/${tag}_full_parity.err
timeout 600 python -c "
import json,sys
sys.path.insert(0,'.')
from tests.test_distribution_parity_gpu import distribution_report
print(json.dumps(distribution_report(), indent=1))" > gpurun_out/${tag}_distribution_parity.json 2> gpurun_out/${tag}_dist.err; tail -2 gpurun_out/${tag}_dist.err
timeout 600 python tools/microbench.py pose 37 40 13,20,21,20,21 > gpurun_out/${tag}_k3_variants.jsonl 2> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_k3_variants.jsonl
for v in 0 1 2; do HOUV_K2_VARIANT=$v timeout 200 python tools/k2_time.py; done > gpurun_out/${tag}_k2_variants.jsonl 2>> gpurun_out/${tag}_variants.err; cat gpurun_out/${tag}_k2_variants.jsonl
(time timeout 900 python bench.py) > gpurun_out/${tag}_bench_n1.json 2> gpurun_out/${tag}_bench.err; tail -c 400 gpurun_out/${tag}_bench_n1.json; tail -4 gpurun_out/${tag}_bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; tail -c 300 gpurun_out/${tag}_bench_ref.json

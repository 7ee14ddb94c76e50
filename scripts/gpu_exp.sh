#!/bin/bash
# one-off: bulk L2 prefetch of the SE trace rows, prefetch distance 2 / 3 / 4 pairs against the line-wise prefetch (base)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/exp_tests.txt 2>&1; tail -2 gpurun_out/exp_tests.txt
bash scripts/gpu_variants.sh --workload trace --no-cpu-baseline --no-e2e
python scripts/show_bench.py gpurun_out/bench_var_*.json

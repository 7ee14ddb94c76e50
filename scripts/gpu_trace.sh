#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --workload trace > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 10 -c 1 -f -o gpurun_out/k_ue_env_trace \
    python bench.py --workload trace --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_trace.log 2>&1

#!/bin/bash
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:k_env_step -s 516 -c 1 -f -o gpurun_out/k_env_step_cfg2 \
    python bench.py --workload cfg2 --no-graph --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_cfg2.log 2>&1
tail -3 gpurun_out/ncu_cfg2.log

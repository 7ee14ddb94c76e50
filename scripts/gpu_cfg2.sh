#!/bin/bash
mkdir -p gpurun_out
for m in 0 1 2; do
  IRS_STEP_MODE=$m python bench.py --workload cfg2 --no-cpu-baseline --no-e2e --steps 200 > gpurun_out/cfg2_mode$m.json 2> gpurun_out/cfg2_mode$m.err
  IRS_STEP_MODE=$m python bench.py --workload mix --envs 8192 --no-cpu-baseline --no-e2e --steps 200 > gpurun_out/e8k_mode$m.json 2> gpurun_out/e8k_mode$m.err
done

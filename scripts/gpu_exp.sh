#!/bin/bash
# one-off: final pass - bench part of the round plus the lane8 ncu capture
bash scripts/gpu_round.sh bench > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_action_8|k_env_tail_8" -s 1030 -c 2 -f -o gpurun_out/k_per_env_mix \
    python bench.py --workload mix --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_perenv.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log
python scripts/show_bench.py gpurun_out/bench.json gpurun_out/bench_mix.json gpurun_out/bench_trace.json gpurun_out/bench_cfg2.json gpurun_out/bench_d401.json gpurun_out/bench_d4096.json | cut -c1-200

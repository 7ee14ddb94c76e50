#!/bin/bash
# ncu passes of the default sweep and the mult_slice mix: launch list + one full capture of k_ue_env and of the per-env kernels.
# The buffers are first filled to their steady state (--fill = what bench.py's plateau detection settles on); unprofiled
# launches are skipped with -s (counts matching kernels: 4 launches per step at 65,536 envs).
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -s 11790 -c 16 --csv --log-file gpurun_out/launches_sweep.csv \
    python bench.py --fill 2944 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 2950 -c 1 -f -o gpurun_out/k_ue_env_sweep \
    python bench.py --fill 2944 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_sweep.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 518 -c 1 -f -o gpurun_out/k_ue_env_mix \
    python bench.py --workload mix --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_mix.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_action_t|k_env_tail_t" -s 1030 -c 2 -f -o gpurun_out/k_per_env_mix \
    python bench.py --workload mix --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_perenv.log 2>&1
ls -la gpurun_out | tail -12

#!/bin/bash
# One GPU-box pass: parity tests, smoke, both bench arms, the other workloads (part "bench"); ncu launch list + full
# captures (part "ncu").  usage: gpu_round.sh [all|bench|ncu]   Run "ncu" first when kernels changed: bench.py's issue
# roofline divides the instruction count of the committed capture (profiles/ue_env_ncu.json) by the live kernel time.
PART=${1:-all}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
if [ $PART = all ] || [ $PART = bench ]; then
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --d2h-probe > gpurun_out/d2h_n1.json 2>&1
python bench.py --workload mix > gpurun_out/bench_mix.json 2> gpurun_out/bench_mix.err
python bench.py --workload trace > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err
python bench.py --workload cfg2 --no-cpu-baseline > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err
python bench.py --depth 401 --steps 100 --no-cpu-baseline > gpurun_out/bench_d401.json 2> gpurun_out/bench_d401.err
python bench.py --depth 4096 --envs 16384 --steps 50 --no-cpu-baseline > gpurun_out/bench_d4096.json 2> gpurun_out/bench_d4096.err
fi
if [ $PART = all ] || [ $PART = ncu ]; then
# ncu passes: the buffers are first filled to their steady state (the fill lengths are what bench.py's plateau detection
# settles on, see config.fill_steps of the bench lines); unprofiled launches are skipped with -s (counts matching kernels:
# 4 launches per step at 65,536 envs)
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:^k_ -s 11790 -c 16 --csv --log-file gpurun_out/launches_sweep.csv \
    python bench.py --fill 2944 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 2950 -c 1 -f -o gpurun_out/k_ue_env_sweep \
    python bench.py --fill 2944 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_sweep.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 518 -c 1 -f -o gpurun_out/k_ue_env_mix \
    python bench.py --workload mix --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_mix.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 518 -c 1 -f -o gpurun_out/k_ue_env_trace \
    python bench.py --workload trace --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_trace.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_ue_env -s 1670 -c 1 -f -o gpurun_out/k_ue_env_d401 \
    python bench.py --depth 401 --fill 1664 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_d401.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_action_8|k_env_tail_8" -s 1030 -c 2 -f -o gpurun_out/k_per_env_mix \
    python bench.py --workload mix --fill 512 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_perenv.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_env_step -s 516 -c 1 -f -o gpurun_out/k_env_step_cfg2 \
    python bench.py --workload cfg2 --no-graph --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu_full_cfg2.log 2>&1
fi
ls -la gpurun_out | tail -30

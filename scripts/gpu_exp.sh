#!/bin/bash
# one-off: GPU suite, A/B of the table-driven small divisions, and the one-launch / three-launch crossover
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu > gpurun_out/exp_tests.txt 2>&1; tail -3 gpurun_out/exp_tests.txt
for w in sweep mix; do
  a=""; [ $w != sweep ] && a="--workload $w"
  for v in new base; do
    l=""; [ $v = base ] && l="$PWD/build_variants/lib_base.so"
    IRS_LIB_PATH=$l python bench.py $a --no-cpu-baseline --no-e2e > gpurun_out/bench_${w}_$v.json 2> gpurun_out/bench_${w}_$v.err
    python scripts/show_bench.py gpurun_out/bench_${w}_$v.json
  done
done
for E in 1024 2368 4096 8192 16384; do
  for m in 1000000 0; do
    IRS_FUSED_MAX_ENVS=$m python bench.py --workload mix --envs $E --no-cpu-baseline --no-e2e > gpurun_out/bench_x_${E}_$m.json 2> gpurun_out/bench_x_${E}_$m.err
    echo "E=$E fused_max=$m"; python scripts/show_bench.py gpurun_out/bench_x_${E}_$m.json
  done
done

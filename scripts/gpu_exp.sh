#!/bin/bash
# one-off: row kernels specialised by SE source at compile time: parity, then A/B
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/exp_tests.txt 2>&1; tail -2 gpurun_out/exp_tests.txt
for w in sweep mix trace cfg2; do
  a=""; [ $w != sweep ] && a="--workload $w"
  for v in new base new base; do
    l=""; [ $v = base ] && l="$PWD/build_variants/lib_base.so"
    IRS_LIB_PATH=$l python bench.py $a --no-cpu-baseline --no-e2e > gpurun_out/bench_${w}_$v.json 2> gpurun_out/bench_${w}_$v.err
    python scripts/show_bench.py gpurun_out/bench_${w}_$v.json | cut -c1-200
  done
done

#!/bin/bash
# a short bench per library variant under build_variants/ (IRS_LIB_PATH) plus the default build; WL="sweep mix" picks workloads
mkdir -p gpurun_out
for w in ${WL:-sweep}; do
  python bench.py --workload $w --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/var_default_$w.json 2> gpurun_out/var_default_$w.err
  for f in build_variants/*.so; do
    [ -f "$f" ] || continue
    n=$(basename $f .so)
    IRS_LIB_PATH=$PWD/$f python bench.py --workload $w --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/var_${n}_$w.json 2> gpurun_out/var_${n}_$w.err
  done
done

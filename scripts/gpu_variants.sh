#!/bin/bash
# a short bench per library variant under build_variants/ (IRS_LIB_PATH) plus the default build
mkdir -p gpurun_out
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/var_default.json 2> gpurun_out/var_default.err
for f in build_variants/*.so; do
  [ -f "$f" ] || continue
  n=$(basename $f .so)
  IRS_LIB_PATH=$PWD/$f python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/var_$n.json 2> gpurun_out/var_$n.err
done

#!/bin/bash
# build variants (build_variants/lib_<tag>.so) against the default library: default sweep, device-timed
mkdir -p gpurun_out
python bench.py --steps 60 --no-cpu-baseline --no-e2e > gpurun_out/bench_var_default.json 2> gpurun_out/bench_var_default.err
for f in build_variants/lib_*.so; do
  tag=$(basename $f .so); tag=${tag#lib_}
  IRS_LIB_PATH=$PWD/$f python bench.py --steps 60 --no-cpu-baseline --no-e2e > gpurun_out/bench_var_$tag.json 2> gpurun_out/bench_var_$tag.err
done

#!/bin/bash
# build variants (build_variants/lib_<tag>.so) against the default library; args = bench flags (default: config 2)
mkdir -p gpurun_out
ARGS="${@:---workload cfg2 --no-cpu-baseline --no-e2e}"
python bench.py $ARGS > gpurun_out/bench_var_default.json 2> gpurun_out/bench_var_default.err
for f in build_variants/lib_*.so; do
  tag=$(basename $f .so); tag=${tag#lib_}
  IRS_LIB_PATH=$PWD/$f python bench.py $ARGS > gpurun_out/bench_var_$tag.json 2> gpurun_out/bench_var_$tag.err
done

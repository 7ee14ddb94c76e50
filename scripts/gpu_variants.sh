#!/bin/bash
# tests on the default build, then a short bench per library variant under build_variants/
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py --steps 100 --warmup 10 --no-cpu-baseline > gpurun_out/var_default.json 2> gpurun_out/var_default.err
for f in build_variants/*.so; do
  n=$(basename $f .so)
  IRS_LIB_PATH=$PWD/$f python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-e2e > gpurun_out/var_$n.json 2> gpurun_out/var_$n.err
done

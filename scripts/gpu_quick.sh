#!/bin/bash
# usage: gpu_quick.sh [bench args...]  -> tests + one bench line
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py "$@" > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err

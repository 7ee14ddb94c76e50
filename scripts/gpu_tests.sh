#!/bin/bash
# usage: gpu_tests.sh [pytest args...]  -> GPU test suite only
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -30 gpurun_out/pytest_gpu.log

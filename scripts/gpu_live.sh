#!/bin/bash
# GPU suite + the main bench workloads (no CPU arm)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for w in sweep mix trace cfg2; do
  python bench.py --workload $w --no-cpu-baseline > gpurun_out/live_$w.json 2> gpurun_out/live_$w.err
done

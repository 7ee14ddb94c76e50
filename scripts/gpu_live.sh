#!/bin/bash
# the bench workloads at their steady state (live-extent rows)
mkdir -p gpurun_out
for w in sweep mix trace cfg2; do
  python bench.py --workload $w > gpurun_out/live_$w.json 2> gpurun_out/live_$w.err
done
python bench.py --workload sweep --depth 401 --no-cpu-baseline > gpurun_out/live_sweep401.json 2> gpurun_out/live_sweep401.err
python bench.py --workload sweep --depth 4096 --envs 16384 --no-cpu-baseline > gpurun_out/live_sweep4096.json 2> gpurun_out/live_sweep4096.err

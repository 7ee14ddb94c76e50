#!/bin/bash
mkdir -p gpurun_out
for d in 401 1024 4096; do
  e=65536; [ $d = 4096 ] && e=16384
  python bench.py --depth $d --envs $e --steps 50 --warmup 10 --no-cpu-baseline > gpurun_out/depth_$d.json 2> gpurun_out/depth_$d.err
done
python bench.py --workload trace --no-cpu-baseline --steps 100 > gpurun_out/bench_trace2.json 2> gpurun_out/bench_trace2.err

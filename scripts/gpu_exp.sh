#!/bin/bash
mkdir -p gpurun_out
for f in 2368 4096 6144 8192; do IRS_HOST_PLAN=geom IRS_HOST_FIRST=$f python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_geom_$f.json 2> gpurun_out/bench_geom_$f.err; done

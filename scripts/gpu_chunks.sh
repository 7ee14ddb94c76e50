#!/bin/bash
# e2e (irs_step_host) against the chunk plan, default sweep
mkdir -p gpurun_out
for c in 3 4 5 6 8 12; do
  IRS_HOST_CHUNKS=$c python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_chunks_$c.json 2> gpurun_out/bench_chunks_$c.err
done
IRS_HOST_PLAN=geom python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_chunks_geom.json 2> gpurun_out/bench_chunks_geom.err
IRS_HOST_PLAN=geom IRS_HOST_STREAMS=2 python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_chunks_geom2.json 2> gpurun_out/bench_chunks_geom2.err
IRS_HOST_CHUNKS=8 IRS_HOST_STREAMS=2 python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_chunks_8s2.json 2> gpurun_out/bench_chunks_8s2.err

#!/bin/bash
# e2e (irs_step_host) against the number of pipeline chunks, default sweep
mkdir -p gpurun_out
for c in 0 4 6 8 12 16; do
  IRS_HOST_CHUNKS=$c python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_chunks_$c.json 2> gpurun_out/bench_chunks_$c.err
done

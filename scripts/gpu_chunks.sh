#!/bin/bash
mkdir -p gpurun_out
for c in 1 2 3 4 6 8 12; do
  IRS_HOST_CHUNKS=$c python bench.py --no-cpu-baseline --steps 30 --warmup 5 > gpurun_out/chunks_$c.json 2> gpurun_out/chunks_$c.err
done

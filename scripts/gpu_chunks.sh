#!/bin/bash
# e2e (irs_step_host) against IRS_HOST_CHUNKS (0 = the library's wave-aligned plan), workloads in WL
mkdir -p gpurun_out
for w in ${WL:-sweep}; do
  for c in ${CH:-0 2 3 4}; do
    if [ $c = 0 ]; then unset IRS_HOST_CHUNKS; else export IRS_HOST_CHUNKS=$c; fi
    python bench.py --workload $w --no-cpu-baseline --steps 30 --warmup 5 > gpurun_out/chunks_${w}_c$c.json 2> gpurun_out/chunks_${w}_c$c.err
  done
done

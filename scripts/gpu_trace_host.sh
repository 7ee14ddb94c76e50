#!/bin/bash
# timelines of irs_step_host (IRS_HOST_TRACE=1) and e2e for chunk plans / stream counts: gpu_trace_host.sh "4 1" "0 2" ...
mkdir -p gpurun_out
for cfg in "$@"; do
  set -- $cfg
  tag="c$1_s$2"
  IRS_HOST_TRACE=1 IRS_HOST_CHUNKS=$1 IRS_HOST_STREAMS=$2 python bench.py --steps 30 --no-cpu-baseline > gpurun_out/bench_host_$tag.json 2> gpurun_out/bench_host_$tag.err
  grep "irs_step_host trace" gpurun_out/bench_host_$tag.err | tail -3 > gpurun_out/host_trace_$tag.txt
done

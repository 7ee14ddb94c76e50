#!/bin/bash
# one-off: record-block prefetch at the entry of the lane8 stages (sweep; trace)
mkdir -p gpurun_out
bash scripts/gpu_variants.sh --no-cpu-baseline --no-e2e --steps 100
for t in default pf; do mv gpurun_out/bench_var_$t.json gpurun_out/bench_varsweep_$t.json; done
bash scripts/gpu_variants.sh --workload trace --no-cpu-baseline --no-e2e --steps 100
for t in default pf; do mv gpurun_out/bench_var_$t.json gpurun_out/bench_vartrace_$t.json; done
python scripts/show_bench.py gpurun_out/bench_varsweep_*.json gpurun_out/bench_vartrace_*.json | cut -c1-190

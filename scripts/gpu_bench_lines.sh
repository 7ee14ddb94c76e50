#!/bin/bash
# both bench arms + the other workloads (no profiler)
mkdir -p gpurun_out
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --workload mix > gpurun_out/bench_mix.json 2> gpurun_out/bench_mix.err
python bench.py --workload trace > gpurun_out/bench_trace.json 2> gpurun_out/bench_trace.err
python bench.py --workload cfg2 --no-cpu-baseline > gpurun_out/bench_cfg2.json 2> gpurun_out/bench_cfg2.err
python bench.py --depth 401 --steps 100 --no-cpu-baseline > gpurun_out/bench_d401.json 2> gpurun_out/bench_d401.err
python bench.py --depth 4096 --envs 16384 --steps 50 --no-cpu-baseline > gpurun_out/bench_d4096.json 2> gpurun_out/bench_d4096.err

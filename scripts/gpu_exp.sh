#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_host_path.py -x -q > gpurun_out/pytest_host.log 2>&1; tail -3 gpurun_out/pytest_host.log
python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_small1.json 2> gpurun_out/bench_small1.err
for c in 5 6 8; do IRS_HOST_CHUNKS=$c python bench.py --steps 40 --no-cpu-baseline > gpurun_out/bench_small1_c$c.json 2> gpurun_out/bench_small1_c$c.err; done

#!/bin/bash
# one-off experiment: parity of the warp-per-env cores, then the phase stamps of the one-launch step (debug build)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -x -q -m gpu -k "one_launch or warp_per_env or not (lane8 or thread_per_env)" > gpurun_out/exp_tests.txt 2>&1
tail -5 gpurun_out/exp_tests.txt
IRS_LIB_PATH=$PWD/build_variants/lib_phaseclock.so python scripts/phase_clock.py 1024 > gpurun_out/phase_clock_cold.txt 2>&1
cat gpurun_out/phase_clock_cold.txt
python bench.py --workload cfg2 --no-cpu-baseline --no-e2e > gpurun_out/bench_cfg2_exp.json 2> gpurun_out/bench_cfg2_exp.err
python scripts/show_bench.py gpurun_out/bench_cfg2_exp.json 2>/dev/null || tail -c 600 gpurun_out/bench_cfg2_exp.json

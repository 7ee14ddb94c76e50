#!/bin/bash
# k_ue_env warps-per-env split on small batches: GPU suite, then cfg2 / 2,048-env benches with the split forced off and automatic
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for s in 0 1 2 auto; do
  if [ $s = auto ]; then unset IRS_UE_SPLIT; else export IRS_UE_SPLIT=$s; fi
  python bench.py --workload cfg2 --no-cpu-baseline --steps 300 > gpurun_out/split_cfg2_$s.json 2> gpurun_out/split_cfg2_$s.err
  python bench.py --workload mix --envs 2048 --no-cpu-baseline --steps 300 > gpurun_out/split_e2048_$s.json 2> gpurun_out/split_e2048_$s.err
  python bench.py --workload mix --envs 4096 --no-cpu-baseline --steps 300 > gpurun_out/split_e4096_$s.json 2> gpurun_out/split_e4096_$s.err
done

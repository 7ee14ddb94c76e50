#!/bin/bash
# e2e (irs_step_host) against the chunking knobs: IRS_HOST_PRIO x IRS_HOST_CHUNKS, workloads in WL
mkdir -p gpurun_out
IRS_HOST_PRIO=1 python -m pytest tests/test_gpu_host_path.py -m gpu -x -q > gpurun_out/pytest_prio.log 2>&1
for w in ${WL:-sweep}; do
  for p in 0 1; do
    for c in 0 8 12; do
      if [ $c = 0 ]; then unset IRS_HOST_CHUNKS; else export IRS_HOST_CHUNKS=$c; fi
      IRS_HOST_PRIO=$p python bench.py --workload $w --no-cpu-baseline --steps 30 --warmup 5 > gpurun_out/chunks_${w}_p${p}_c$c.json 2> gpurun_out/chunks_${w}_p${p}_c$c.err
    done
  done
done

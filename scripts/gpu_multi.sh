#!/bin/bash
# usage: gpu_multi.sh N  -> bench (both arms) + D2H probe on N GPUs of one box
N=$1
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_${N}gpu.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
$TR bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
$TR bench.py --gpus $N --d2h-probe > gpurun_out/d2h_n$N.json 2> gpurun_out/d2h_n$N.err
$TR bench.py --gpus $N --impl reference --steps 20 --warmup 3 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
tail -2 gpurun_out/bench_n$N.err

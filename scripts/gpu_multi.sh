#!/bin/bash
# multi-GPU visit (gpurun --gpus N): C-path and torch.distributed assembly tests, then the bench at N ranks
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/multi.txt
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -x -q >> gpurun_out/multi.txt 2>&1
tail -3 gpurun_out/multi.txt
time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench exit $?"; tail -1 gpurun_out/bench_n$N.json | cut -c1-300; grep -v "^W\|Warning\|warn" gpurun_out/bench_n$N.err | tail -5

#!/bin/bash
# 2-GPU visit: NCCL assembly parity test, then the bench at N=2 (torchrun, one rank per GPU)
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/multi.txt
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q >> gpurun_out/multi.txt 2>&1
tail -3 gpurun_out/multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "bench exit $?"; tail -1 gpurun_out/bench_n2.json | cut -c1-600; tail -3 gpurun_out/bench_n2.err

#!/bin/bash
# multi-GPU visit (gpurun --gpus N): NCCL assembly parity test, then the bench at N ranks (torchrun)
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/multi.txt
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -x -q >> gpurun_out/multi.txt 2>&1
tail -3 gpurun_out/multi.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench exit $?"; tail -1 gpurun_out/bench_n$N.json | cut -c1-400; tail -3 gpurun_out/bench_n$N.err

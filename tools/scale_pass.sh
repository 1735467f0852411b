#!/bin/bash
# bench.py at N GPUs of one box exactly as the driver launches it (torchrun, NCCL), bounded by `timeout`.
# usage: tools/scale_pass.sh N [extra bench flags]    -> gpurun_out/r02_bench_c2_n${N}.json
set -u
N=${1:-2}; shift || true
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 10 --warmup 3 "$@" > gpurun_out/r02_bench_c2_n${N}.json 2> gpurun_out/r02_bench_c2_n${N}.err
echo "rc=$?"; tail -c 400 gpurun_out/r02_bench_c2_n${N}.json

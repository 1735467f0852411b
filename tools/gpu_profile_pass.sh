#!/bin/bash
# One profiling pass on the GPU box: bench (plain), launch list of the same command under ncu, and one
# `ncu --set full` capture of the hot kernels from tools/ncu_targets.py.  Tag = $1.
set -u
T=${1:-r01x}
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_ncu_bench.log 2>&1; echo "launch list rc=$?"
python tools/ncu_targets.py > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:tc_gemm|tc_linear|tc_mlp_chain|nll_head|attn_|setconv_|gauss_nll|add_layernorm' -c 80 \
    -o gpurun_out/${T}_full -f python tools/ncu_targets.py > gpurun_out/${T}_ncu_full.log 2>&1; echo "full rc=$?"
ls -la gpurun_out/
tail -c 1500 gpurun_out/${T}_bench.json

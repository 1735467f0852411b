#!/bin/bash
# One profiling pass on the GPU box: bench (plain), launch list of the same command under ncu, and one
# `ncu --set full` capture of the hot kernels from tools/ncu_targets.py.  Tag = $1.
# The .ncu-rep stays on the box (gpurun copies back at most 64 MiB): only its raw-page CSV comes home.
set -u
T=${1:-r02}
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_fast.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/${T}_ncu_bench.log 2>&1; echo "launch list rc=$?"
python tools/ncu_targets.py > /dev/null 2>&1 && \
ncu --set full --clock-control none -k 'regex:tc_gemm|tc_linear|tc_mlp_chain|tc_chain_bwd|nll_head|attn_|setconv_|gauss_nll|add_layernorm|skinny_' -c 100 \
    -o /tmp/${T}_full -f python tools/ncu_targets.py > gpurun_out/${T}_ncu_full.log 2>&1; echo "full rc=$?"
ncu -i /tmp/${T}_full.ncu-rep --page raw --csv > gpurun_out/${T}_full_raw.csv 2>/dev/null; echo "raw csv rc=$?"
ls -la gpurun_out/ /tmp/${T}_full.ncu-rep
tail -c 600 gpurun_out/${T}_bench_fast.json

#!/bin/bash
# Light evidence pass (no `ncu --set full`): tests, bench line, launch list of the same command, launch lists of the
# REPLAYED C2 / C3 steps.  Tag = $1.
set -u
T=${1:-r02}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/${T}_gputest.log 2>&1; echo "tests rc=$?"; tail -1 gpurun_out/${T}_gputest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_fast.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/${T}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/${T}_ncu_bench.log 2>&1; echo "launch list rc=$?"
for spec in "AttnCNP 256" "ConvCNP_unet 32"; do
  set -- $spec
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \
      --log-file gpurun_out/${T}_graph_${1}.csv python tools/graph_launch_list.py $1 $2 > gpurun_out/${T}_graph_${1}.log 2>&1; echo "graph list $1 rc=$?"
done
ls -la gpurun_out | head -30

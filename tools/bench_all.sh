#!/bin/bash
# every single-GPU workload of bench.py at HEAD -> gpurun_out/${T}_bench_<workload>.json   (T = $1, default r02)
T=${1:-r02}
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/${T}_bench_fast.json 2> gpurun_out/${T}_bench_fast.err
for wl in convcnp_c3 attnlnp_c4 cnp_c1 attncnp_c5_infer convcnp_c5_infer; do
  python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_${wl}.json 2> gpurun_out/${T}_bench_${wl}.err
done
python bench.py --workload attnlnp_c4 --scaling strong --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${T}_bench_attnlnp_c4_strong.json 2>/dev/null
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2>/dev/null
ls gpurun_out | grep ${T}_bench

#!/bin/bash
# GPU box: option sort_rays (LBVH walkers take their rays bin by bin) across scenes and primitive counts.
mkdir -p gpurun_out
T=${TAG:-sr}
run() { # name scene nx ny spp extra
  timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err
  echo "$1: exit $? $(python -c "import json; d=json.load(open('gpurun_out/${T}_$1.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms', {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
S="--opt sort_rays=2 --opt sort_obits=5 --opt sort_dbits=1"
run s2_default 2 1200 600 64 ""
run s2_walker 2 1200 600 64 "--opt walker=1"
run s2_walker_sort 2 1200 600 64 "--opt walker=1 $S"
run s1_default 1 1920 1080 32 ""
run s1_pool_walker 1 1920 1080 32 "--opt stream=0 --opt walker=1"
run s1_pool_walker_sort 1 1920 1080 32 "--opt stream=0 --opt walker=1 $S"
for n in 16384 131072; do
  run s5_${n}_default 5 1920 1080 8 "--stress-prims $n"
  run s5_${n}_sort 5 1920 1080 8 "--stress-prims $n $S"
done
run s4_sort_o5d1_r1 4 800 800 64 "--opt sort_rays=1 --opt sort_obits=5 --opt sort_dbits=1"
run s4_sort_o5d1_r2 4 800 800 64 "$S"
run s4_default 4 800 800 64 ""

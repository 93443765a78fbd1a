#!/bin/bash
# GPU box: leaf size 1 everywhere — streaming / flat kernels against the walkers.
mkdir -p gpurun_out
run() { # name scene nx ny spp extra
  timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/ac3_$1.json 2> gpurun_out/ac3_$1.err
  echo "$1: $(python -c "import json; d=json.load(open('gpurun_out/ac3_$1.json')); print(round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
for n in 131072 1048576; do
  run s5_${n}_walker 5 1920 1080 8 "--stress-prims $n"
  run s5_${n}_stream_leaf1 5 1920 1080 8 "--stress-prims $n --opt walker=0 --opt stream=1 --opt leaf_size=1"
done
run s2_auto 2 1200 600 64 ""
run s2_flat_leaf1 2 1200 600 64 "--opt walker=0 --opt leaf_size=1"
run s2_flat_leaf2 2 1200 600 64 "--opt walker=0 --opt leaf_size=2"
run s4_auto 4 800 800 64 ""
run s4_flat_leaf1 4 800 800 64 "--opt walker=0 --opt leaf_size=1"
run s1_pool_flat_leaf1 1 1920 1080 32 "--opt stream=0 --opt walker=0 --opt leaf_size=1"

#!/bin/bash
# GPU box: the no-NEE side of the auto rules after the walker defaults changed (leaf 1, min_active 12, step_nodes 4).
mkdir -p gpurun_out
run() { # name scene nx ny spp extra
  timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/ac2_$1.json 2> gpurun_out/ac2_$1.err
  echo "$1: $(python -c "import json; d=json.load(open('gpurun_out/ac2_$1.json')); print(round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
run s1_auto 1 1920 1080 32 ""
run s1_walker 1 1920 1080 32 "--opt stream=0 --opt walker=1"
run s1_stream_leaf1 1 1920 1080 32 "--opt leaf_size=1"
run s1_stream_leaf2 1 1920 1080 32 "--opt leaf_size=2"
run s1_stream_leaf8 1 1920 1080 32 "--opt leaf_size=8"
for n in 512 2048 16384; do
  run s5_${n}_auto 5 1920 1080 8 "--stress-prims $n"
  run s5_${n}_walker 5 1920 1080 8 "--stress-prims $n --opt walker=1 --opt stream=0"
  run s5_${n}_stream_leaf1 5 1920 1080 8 "--stress-prims $n --opt leaf_size=1"
done

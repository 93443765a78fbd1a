#!/bin/bash
# GPU box: second round of the walker sweep with binned rays: leaf size, refill threshold, step.
mkdir -p gpurun_out
run() { # name scene nx ny spp extra
  timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/sw4_$1.json 2> gpurun_out/sw4_$1.err
  echo "$1: $(python -c "import json; d=json.load(open('gpurun_out/sw4_$1.json')); print(round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
for o in "leaf_size=2,min_active=16" "leaf_size=2,min_active=12" "leaf_size=2,min_active=8" "leaf_size=1,min_active=16" "leaf_size=2,min_active=16,step_nodes=4" "leaf_size=2,min_active=16,step_prims=4" "leaf_size=3,min_active=16" "min_active=12" "min_active=8" "leaf_size=2,min_active=20"; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  tag=$(echo $o | tr ',=' '__')
  run s4_$tag 4 800 800 64 "$args"
  run s5_$tag 5 1920 1080 8 "$args"
  run s2_$tag 2 1200 600 64 "$args"
done

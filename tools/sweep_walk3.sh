#!/bin/bash
# GPU box: walker settings again, now that the rays arrive bin by bin (round 2).
mkdir -p gpurun_out
run() { # name scene nx ny spp extra
  timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/sw3_$1.json 2> gpurun_out/sw3_$1.err
  echo "$1: $(python -c "import json; d=json.load(open('gpurun_out/sw3_$1.json')); print(round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
for o in "step_nodes=8" "step_nodes=4" "step_nodes=16" "step_nodes=32,step_prims=16" "min_active=16" "min_active=28" "min_active=30" "step_prims=4" "step_prims=16" "leaf_size=2" "leaf_size=8" "step_nodes=16,min_active=28" "step_nodes=4,min_active=28"; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  tag=$(echo $o | tr ',=' '__')
  run s4_$tag 4 800 800 64 "$args"
done
for o in "step_nodes=8" "step_nodes=16" "min_active=28" "leaf_size=8" "leaf_size=2" "step_nodes=16,min_active=28"; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  tag=$(echo $o | tr ',=' '__')
  run s5_$tag 5 1920 1080 8 "$args"
done

#!/bin/bash
# Runs on the GPU box: streaming wavefront (LBVH walk inside the fused kernels) vs the slot pool on the larger scenes.
mkdir -p gpurun_out
run() {  # scene nx ny spp opts...
  local s=$1 nx=$2 ny=$3 spp=$4; shift 4
  local args=""; for kv in "$@"; do args="$args --opt $kv"; done
  timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --scene $s --nx $nx --ny $ny --spp $spp $args > gpurun_out/tmp.json 2> gpurun_out/tmp.err
  echo "scene $s $* : exit $? $(python -c "import json; d=json.load(open('gpurun_out/tmp.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms', {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
for sc in "1 1920 1080 32" "2 1200 600 64" "4 800 800 64" "3 600 600 128"; do
  run $sc stream=0
  run $sc stream=1 shade_ctas=2
  run $sc stream=1 shade_ctas=3
done

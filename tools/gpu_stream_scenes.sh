#!/bin/bash
# Runs on the GPU box: option sets on the larger scenes (streaming wavefront vs slot pool, waves).
mkdir -p gpurun_out
run() {  # scene nx ny spp opts...
  local s=$1 nx=$2 ny=$3 spp=$4; shift 4
  local args=""; for kv in "$@"; do args="$args --opt $kv"; done
  timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --scene $s --nx $nx --ny $ny --spp $spp $args > gpurun_out/tmp.json 2> gpurun_out/tmp.err
  echo "scene $s $* : exit $? $(python -c "import json; d=json.load(open('gpurun_out/tmp.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms')" 2>&1 | tail -1)"
}
for sc in "1 1920 1080 32" "2 1200 600 64" "4 800 800 64" "3 600 600 128"; do
  for o in ${OPTSETS:-waves=1 waves=2 waves=3}; do run $sc $(echo $o | tr ',' ' '); done
done

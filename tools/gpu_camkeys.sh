#!/bin/bash
# GPU box: camera rays binned by direction (option sort_camera) on the walker scenes.
mkdir -p gpurun_out
T=${TAG:-ck}
run() { # name scene nx ny spp extra
  timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err
  echo "$1: exit $? $(python -c "import json; d=json.load(open('gpurun_out/${T}_$1.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms', {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
run s4_auto 4 800 800 64 ""
run s4_nocam 4 800 800 64 "--opt sort_camera=0"
run s4_off 4 800 800 64 "--opt sort_rays=0"
for n in 16384 131072 1048576; do
  run s5_${n}_off 5 1920 1080 8 "--stress-prims $n --opt sort_rays=0"
  run s5_${n}_cam 5 1920 1080 8 "--stress-prims $n --opt sort_rays=1"
  run s5_${n}_nocam 5 1920 1080 8 "--stress-prims $n --opt sort_rays=1 --opt sort_camera=0"
done
run s2_walker_cam 2 1200 600 64 "--opt walker=1 --opt sort_rays=1"
run s2_default 2 1200 600 64 ""
run s1_pool_walker_cam 1 1920 1080 32 "--opt stream=0 --opt walker=1 --opt sort_rays=1"
run s1_default 1 1920 1080 32 ""

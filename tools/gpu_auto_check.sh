#!/bin/bash
# GPU box: the auto rules of round 2 (walkers + ray binning) against their alternatives.
mkdir -p gpurun_out
T=${TAG:-auto}
run() { # name scene nx ny spp extra
  timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err
  echo "$1: exit $? $(python -c "import json; d=json.load(open('gpurun_out/${T}_$1.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms', {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
run s4_auto 4 800 800 64 ""
run s4_sort2 4 800 800 64 "--opt sort_rays=2"
run s2_auto 2 1200 600 64 ""
run s2_sort2 2 1200 600 64 "--opt sort_rays=2"
run s2_flat 2 1200 600 64 "--opt walker=0"
run s1_auto 1 1920 1080 32 ""
for n in 2048 16384; do
  run s5_${n}_auto 5 1920 1080 8 "--stress-prims $n"
  run s5_${n}_stream 5 1920 1080 8 "--stress-prims $n --opt walker=0 --opt stream=1"
done
run s5_512_auto 5 1920 1080 8 "--stress-prims 512"
run s5_512_walker 5 1920 1080 8 "--stress-prims 512 --opt walker=1 --opt stream=0"

#!/bin/bash
# GPU box: A/B of a tuning build (raytracing-weekend_b200/librtw_$V.so, `make variant`) against librtw.so on the LBVH scenes.
mkdir -p gpurun_out
V=${V:-f1}
run() { # name scene nx ny spp extra
  timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/ab_$1.json 2> gpurun_out/ab_$1.err
  echo "$1: $(python -c "import json; d=json.load(open('gpurun_out/ab_$1.json')); print(round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
for rep in 1 2; do
for lib in base $V; do
  if [ "$lib" = "base" ]; then unset RTW_LIB; else export RTW_LIB=$PWD/raytracing-weekend_b200/librtw_$lib.so; fi
  run s4_${lib}_$rep 4 800 800 64 ""
  run s1_${lib}_$rep 1 1920 1080 32 ""
  run s5_${lib}_$rep 5 1920 1080 8 ""
  run s2_${lib}_$rep 2 1200 600 64 ""
done
done

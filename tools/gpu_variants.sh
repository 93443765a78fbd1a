#!/bin/bash
# Runs on the GPU box: tuning builds of librtw (RTW_LIB=...) x option sets, one bench line each.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -q -x -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
for lib in ${LIBS}; do
  for o in ${OPTSETS}; do
    args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
    RTW_LIB=$PWD/raytracing-weekend_b200/$lib timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline $args > gpurun_out/tmp.json 2> gpurun_out/tmp.err
    echo "$lib $o: exit $? $(python -c "import json; d=json.load(open('gpurun_out/tmp.json')); print(round(d['value'],1), round(d['ms_per_step'],2), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
  done
done

#!/bin/bash
# Runs on the GPU box: GPU tests, then the bench with the streaming wavefront against the slot pool, shade_ctas sweep.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -q -x -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.log
for o in "stream=0" "stream=1 shade_ctas=2" "stream=1 shade_ctas=3" "stream=1 shade_ctas=4" ${EXTRA}; do
  args=""; for kv in $o; do args="$args --opt $kv"; done
  tag=$(echo $o | tr ' =' '__')
  timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline $args > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
  echo "$o: exit $? $(python -c "import json,sys; d=json.load(open('gpurun_out/bench_$tag.json')); print(d['value'], d['ms_per_step'], d['mrays_per_s'], {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
done

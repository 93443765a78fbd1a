#!/bin/bash
# Runs on the GPU box: option sweep on one scene (default scene 4, 800x800, 64 spp).
mkdir -p gpurun_out
T=${TAG:-sweep}
for o in ${OPTSETS}; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene ${SCENE:-4} --nx ${NX:-800} --ny ${NY:-800} --spp ${SPP:-64} $args > gpurun_out/${T}_tmp.json 2> gpurun_out/${T}_tmp.err
  echo "$o: exit $? $(python -c "import json; d=json.load(open('gpurun_out/${T}_tmp.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms', {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
done

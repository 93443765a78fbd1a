#!/bin/bash
# Runs on the GPU box: LBVH-scene variants.  One short bench line per (scene, option set).
mkdir -p gpurun_out
T=${TAG:-r02d}
if [ -z "$NOTEST" ]; then
echo "== pytest"; timeout 1500 python -m pytest ${TESTS:-tests/test_gpu_tier1.py tests/test_gpu_render.py tests/test_gpu_ref_pin.py tests/test_gpu_edges.py tests/test_golden.py} -q -x -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/${T}_pytest.log
fi
run() { # name scene nx ny spp extra
  timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-relmse --scene $2 --nx $3 --ny $4 --spp $5 $6 > gpurun_out/${T}_$1.json 2> gpurun_out/${T}_$1.err
  echo "$1: exit $? $(python -c "import json; d=json.load(open('gpurun_out/${T}_$1.json')); print(round(d['value'],1), 'Mpaths/s', round(d['mrays_per_s'],0), 'Mrays/s', round(d['ms_per_step'],2), 'ms', {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
for o in ${OPTSETS:-wide=1 wide=0}; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  tag=$(echo $o | tr ',=' '__')
  run s4_$tag 4 800 800 ${SPP4:-64} "$args"
  run s1_$tag 1 1920 1080 ${SPP1:-32} "$args"
  run s2_$tag 2 1200 600 ${SPP2:-64} "$args"
  if [ -n "$C5" ]; then run s5_$tag 5 1920 1080 4 "$args"; fi
done

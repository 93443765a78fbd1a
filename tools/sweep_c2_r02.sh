#!/bin/bash
# GPU box: C2 scheduling options once more, after the hit record moved into the slot record (round 2).
mkdir -p gpurun_out
for o in "" "waves=3" "shade_ctas=4" "shade_ctas=2" "pool_paths=1572864" "pool_paths=2621440" "pool_paths=3145728" "lag=2" "lag=5" "tail_paths=65536" "tail_paths=262144"; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-relmse $args > gpurun_out/c2s_tmp.json 2>/dev/null
  echo "[$o] $(python -c "import json; d=json.load(open('gpurun_out/c2s_tmp.json')); print(round(d['value'],1), round(d['e2e']['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})")"
done

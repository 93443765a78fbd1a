#!/bin/bash
# GPU box: slots per wave on C2 (two waves), finer than the powers of two.
mkdir -p gpurun_out
for o in "" "pool_paths=2359296" "pool_paths=2621440" "pool_paths=2883584" "pool_paths=3145728" "pool_paths=4194304" "" "pool_paths=2621440" "pool_paths=2883584"; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-relmse $args > gpurun_out/c2p_tmp.json 2>/dev/null
  echo "[$o] $(python -c "import json; d=json.load(open('gpurun_out/c2p_tmp.json')); print(round(d['value'],1), round(d['e2e']['value'],1), round(d['ms_per_step'],2))")"
done
# the share of one GPU of eight (128 spp) with the same pools
for o in "" "pool_paths=2621440"; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  python bench.py --steps 5 --warmup 3 --spp 128 --no-cpu-baseline --no-relmse $args > gpurun_out/c2p_tmp.json 2>/dev/null
  echo "128 spp [$o] $(python -c "import json; d=json.load(open('gpurun_out/c2p_tmp.json')); print(round(d['value'],1), round(d['ms_per_step'],3))")"
done

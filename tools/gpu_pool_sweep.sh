#!/bin/bash
# GPU box: C2 with small pools (does an L2-resident pool beat the HBM-streaming one?) and the run-to-run spread of the default.
mkdir -p gpurun_out
T=${TAG:-pool}
for cfg in "0 0" "262144 2" "262144 4" "524288 2" "524288 4" "1048576 2" "1048576 4" "131072 4"; do
  set -- $cfg
  args="--opt pool_paths=$1 --opt waves=$2"; [ "$1" = "0" ] && args=""
  timeout 300 python bench.py --steps 3 --warmup 3 --spp ${SPP:-256} --no-cpu-baseline --no-relmse $args > gpurun_out/${T}_$1_$2.json 2> gpurun_out/${T}_$1_$2.err
  echo "pool $1 waves $2: $(python -c "import json; d=json.load(open('gpurun_out/${T}_$1_$2.json')); print(round(d['value'],1), 'Mpaths/s', round(d['ms_per_step'],2), 'ms', d['gpu_launches'], {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
done
if [ -n "$SPREAD" ]; then
  for i in 1 2 3 4 5; do
    timeout 300 python bench.py --no-cpu-baseline --no-relmse > gpurun_out/${T}_spread_$i.json 2> gpurun_out/${T}_spread_$i.err
    echo "default run $i: $(python -c "import json; d=json.load(open('gpurun_out/${T}_spread_$i.json')); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['ms_per_step'],2))" 2>&1 | tail -1)"
  done
fi

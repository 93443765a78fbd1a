#!/bin/bash
# Runs on the GPU box: BASELINE.json's configurations at their FULL settings, one bench line each under
# gpurun_out/r02_configs/ (copied to profiles/r02_configs/).  GPUS = the GPU counts to run (default "1").
#   CONFIGS="C3 C4"   which configurations;  C5 needs --no-relmse (its converged reference alone is a minute of GPU time)
mkdir -p gpurun_out/r02_configs
for cfg in ${CONFIGS:-C3 C4}; do
  for n in ${GPUS:-1}; do
    extra="--steps ${STEPS:-1} --warmup ${WARMUP:-1} --no-cpu-baseline"
    if [ "$cfg" = "C5" ] || [ "$n" != "1" ]; then extra="$extra --no-relmse"; fi
    out=gpurun_out/r02_configs/${cfg}_n${n}.json
    if [ "$n" = "1" ]; then
      timeout ${TMO:-1500} python bench.py --config $cfg --gpus 1 $extra > $out 2> gpurun_out/r02_configs/${cfg}_n${n}.err
    else
      timeout ${TMO:-1500} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --config $cfg --gpus $n $extra > $out 2> gpurun_out/r02_configs/${cfg}_n${n}.err
    fi
    echo "$cfg n=$n: exit $? $(python -c "import json; d=json.load(open('$out')); print(round(d['value'],1), 'Mpaths/s', round(d['mrays_per_s'],0), 'Mrays/s', round(d['ms_per_step'],1), 'ms/frame build', round(d['build']['lbvh_build_ms'],2), 'ms', 'dominant', d['roofline']['kernel'], round(d['roofline']['frac'],3), 'l2', d['roofline'].get('l2',{}).get('frac'))" 2>&1 | tail -1)"
  done
done

#!/bin/bash
# On an 8-GPU box: bench.py at N = 2, 4, 8 launched the way the driver does (torchrun, NCCL, 127.0.0.1).
mkdir -p gpurun_out
for n in ${NS:-2 4 8}; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
    bench.py --gpus $n --steps ${STEPS:-3} --warmup ${WARMUP:-3} > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "N=$n exit $?"; python -c "
import json,sys
d=json.loads(open('gpurun_out/bench_n$n.json').read().strip().splitlines()[-1])
print('N=%d value %.1f Mpaths/s e2e %.1f ms/step %.2f clocks %s' % (d['n_gpus'], d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks']))"
done

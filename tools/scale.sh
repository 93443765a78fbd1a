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
[ -n "$SKIP_REF" ] && exit 0
# the reference arm under the same launcher: rank 0 prints, the others exit 0
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29555 \
  bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err
echo "reference arm N=2 exit $?"; cut -c1-300 gpurun_out/bench_ref_n2.json

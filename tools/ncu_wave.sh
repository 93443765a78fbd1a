#!/bin/bash
# Runs on the GPU box: one `ncu --set full` capture of a steady-state wavefront iteration (the 5 kernels
# of iteration ~40 of a C2 frame), after the same command exited 0 without ncu.  TAG names the report.
TAG=${TAG:-r01x}
mkdir -p gpurun_out
timeout 600 python bench.py --steps 1 --warmup 0 --spp ${SPP:-256} --no-cpu-baseline ${BENCH_OPTS} > gpurun_out/ncu_plain_run.log 2>&1; echo "plain exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on --launch-skip ${SKIP:-230} --launch-count ${COUNT:-5} \
  -o gpurun_out/${TAG}_wave -f python bench.py --steps 1 --warmup 0 --spp ${SPP:-256} --no-cpu-baseline ${BENCH_OPTS} > gpurun_out/ncu_full.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out/${TAG}_wave.ncu-rep

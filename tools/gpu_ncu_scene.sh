#!/bin/bash
# Runs on the GPU box: `ncu --set full` of the traversal (and shading) kernels of one steady-state iteration of an
# LBVH scene, after the same command exited 0 without ncu.  TAG names the report; SCENE/NX/NY/SPP pick the frame.
T=${TAG:-r02e}
mkdir -p gpurun_out
ARGS="--steps 1 --warmup 0 --no-cpu-baseline --no-relmse --scene ${SCENE:-4} --nx ${NX:-800} --ny ${NY:-800} --spp ${SPP:-16} ${BENCH_OPTS}"
timeout 600 python bench.py $ARGS > gpurun_out/${T}_plain.json 2> gpurun_out/${T}_plain.err; echo "plain exit $?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"${KERNELS:-k_traverse|k_shadow|k_shade_all}" --launch-skip ${SKIP:-24} --launch-count ${COUNT:-3} \
  -o gpurun_out/${T}_scene${SCENE:-4} -f python bench.py $ARGS > gpurun_out/${T}_ncu.log 2>&1; echo "ncu exit $?"; tail -2 gpurun_out/${T}_ncu.log
ls -la gpurun_out/${T}_scene${SCENE:-4}.ncu-rep

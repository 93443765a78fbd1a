#!/bin/bash
# Runs on the GPU box (via gpurun): GPU tests, smoke, a short bench; logs under gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.csv 2>&1
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_gpu.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -3 gpurun_out/smoke.log
echo "== bench"; timeout 900 python bench.py --steps ${STEPS:-2} --warmup ${WARMUP:-1} > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
echo "== cli"; timeout 300 ./raytracing-weekend_b200/rtw_cli -s 0 -dx 600 -dy 600 -ns 64 -v -j -o gpurun_out/cornell_64spp.ppm 2> gpurun_out/cli.log; echo "cli exit $?"; cat gpurun_out/cli.log

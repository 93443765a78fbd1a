#!/bin/bash
# Runs on the GPU box: the render-parity tests, then bench.py once per k_shade_all variant (option shade_ctas).
mkdir -p gpurun_out
echo "== pytest (render + tier1)"; timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_tier1.py -q -m gpu -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/pytest_gpu.log
for c in ${CTAS:-2 3 4}; do
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --opt shade_ctas=$c > gpurun_out/bench_ctas$c.json 2> gpurun_out/bench_ctas$c.err; echo "ctas=$c exit $?"
  python - <<P
import json
d=json.load(open("gpurun_out/bench_ctas$c.json"))
print("shade_ctas=$c", round(d["value"],1), "Mpaths/s", {k:round(v["ms_per_iteration"],4) for k,v in d["roofline"]["stages"].items()})
P
done

#!/bin/bash
# Runs on the GPU box (gpurun): round-2 check of a kernel change.
#   TESTS="..."     pytest targets (default: the render / reference-pin tests)
#   OPTSETS="a=1,b=2 c=3"   one short bench line per option set (besides the default)
#   SANITIZE=1      compute-sanitizer memcheck / racecheck / initcheck over smoke() and one LBVH frame
#   NCU=TAG         launch list + one --set full capture of a steady-state iteration -> gpurun_out/TAG_*
mkdir -p gpurun_out
T=${TAG:-r02}
echo "== pytest"; timeout 1500 python -m pytest ${TESTS:-tests/test_gpu_render.py tests/test_gpu_ref_pin.py tests/test_gpu_edges.py tests/test_golden.py} -q -x -m gpu > gpurun_out/${T}_pytest.log 2>&1; echo "pytest exit $?"; tail -4 gpurun_out/${T}_pytest.log
echo "== smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
run_bench() { # name, extra args
  timeout 600 python bench.py --steps ${STEPS:-4} --warmup 3 --no-cpu-baseline --no-relmse $2 > gpurun_out/${T}_bench_$1.json 2> gpurun_out/${T}_bench_$1.err
  echo "bench $1: exit $? $(python -c "import json; d=json.load(open('gpurun_out/${T}_bench_$1.json')); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['ms_per_step'],2), 'frac', round(d['roofline']['frac'],3), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})" 2>&1 | tail -1)"
}
run_bench default ""
for o in ${OPTSETS}; do
  args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
  run_bench $(echo $o | tr ',=' '__') "$args"
done
if [ -n "$SANITIZE" ]; then
  for tool in memcheck racecheck initcheck; do
    echo "== compute-sanitizer $tool (smoke)"
    timeout 900 compute-sanitizer --tool $tool --log-file gpurun_out/${T}_sanitizer_${tool}_smoke.log python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_sanitizer_${tool}_smoke.out 2>&1; echo "exit $?"; tail -3 gpurun_out/${T}_sanitizer_${tool}_smoke.log
    echo "== compute-sanitizer $tool (scene 4, two waves, LBVH walkers)"
    timeout 1200 compute-sanitizer --tool $tool --log-file gpurun_out/${T}_sanitizer_${tool}_scene4.log python tools/sanitizer_frame.py > gpurun_out/${T}_sanitizer_${tool}_scene4.out 2>&1; echo "exit $?"; tail -3 gpurun_out/${T}_sanitizer_${tool}_scene4.log
  done
fi
if [ -n "$NCU" ]; then
  echo "== ncu launch list"; timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1400 --csv --log-file gpurun_out/${NCU}_launches_c2_128spp.csv python bench.py --steps 1 --warmup 0 --spp 128 --no-cpu-baseline --no-relmse > gpurun_out/ncu_list.log 2>&1; echo "ncu exit $?"
  timeout 1200 ncu --set full --clock-control none --import-source on --launch-skip ${SKIP:-230} --launch-count ${COUNT:-6} \
    -o gpurun_out/${NCU}_wave -f python bench.py --steps 1 --warmup 0 --spp 256 --no-cpu-baseline --no-relmse > gpurun_out/ncu_full.log 2>&1; echo "ncu full exit $?"; tail -2 gpurun_out/ncu_full.log
  ls -la gpurun_out/${NCU}_wave.ncu-rep
fi

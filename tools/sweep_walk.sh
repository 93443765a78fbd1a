# tuning sweep of the traversal variants (run on the GPU box): flat kernels vs persistent walkers
for sc in "1 960 540" "2 1200 600" "4 800 800"; do set -- $sc
for o in "walker=0" "walker=1 step_nodes=16 step_prims=8 min_active=22" "walker=1 step_nodes=64 step_prims=32 min_active=12"; do
args=""; for kv in $o; do args="$args --opt $kv"; done
python bench.py --steps 2 --warmup 1 --scene $1 --nx $2 --ny $3 --spp 64 --no-cpu-baseline $args 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('scene $1 [$o]', round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})"
done; done

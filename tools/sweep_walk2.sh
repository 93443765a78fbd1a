# Runs on the GPU box: step sizes / refill threshold of the persistent walkers on scenes 4 and 2 (lane utilisation).
for sc in "4 800 800 32" "2 1200 600 32"; do set -- $sc
for o in ${SETS:-"walker=1,step_nodes=16,step_prims=8,min_active=22" "walker=1,step_nodes=4,step_prims=4,min_active=22" "walker=1,step_nodes=2,step_prims=2,min_active=22" "walker=1,step_nodes=1,step_prims=1,min_active=22" "walker=1,step_nodes=4,step_prims=2,min_active=28" "walker=1,step_nodes=2,step_prims=1,min_active=28" "walker=1,step_nodes=4,step_prims=4,min_active=16" "walker=1,step_nodes=8,step_prims=2,min_active=24" "walker=0"}; do
args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
python bench.py --steps 1 --warmup 1 --scene $1 --nx $2 --ny $3 --spp $4 --no-cpu-baseline $args 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('scene $1 [$o]', round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})"
done; done

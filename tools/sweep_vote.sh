# Runs on the GPU box: warp-vote thresholds of the walkers' inner loops (Walker::step_voted) on scenes 4 and 2.
for sc in "4 800 800 32" "2 1200 600 32"; do set -- $sc
for o in ${SETS:-"walker=1" "walker=1,node_min=4,prim_min=0" "walker=1,node_min=8,prim_min=0" "walker=1,node_min=12,prim_min=0" "walker=1,node_min=16,prim_min=0" "walker=1,node_min=8,prim_min=4" "walker=1,node_min=8,prim_min=8" "walker=1,node_min=12,prim_min=8" "walker=1,node_min=8,prim_min=4,step_nodes=32,step_prims=16" "walker=1,node_min=12,prim_min=6,step_nodes=64,step_prims=32,min_active=26"}; do
args=""; for kv in $(echo $o | tr ',' ' '); do args="$args --opt $kv"; done
python bench.py --steps 1 --warmup 1 --scene $1 --nx $2 --ny $3 --spp $4 --no-cpu-baseline $args 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('scene $1 [$o]', round(d['value'],1), {k:round(v['ms_per_iteration'],4) for k,v in d['roofline']['stages'].items()})"
done; done

for spec in "base:" "m1:GSB_SORT_MODE=1" "m2:GSB_SORT_MODE=2" "m2s5:GSB_SORT_MODE=2,GSB_SORT=5" "tp1:GSB_TREELET_PASSES=1" "tp0:GSB_TREELET_PASSES=0"; do
  tag=${spec%%:*}; envs=${spec#*:}
  ( IFS=','; for kv in $envs; do [ -n "$kv" ] && export "$kv"; done; unset IFS
    python bench.py --scene field --steps 3 --spp-per-step 16 --no-e2e --no-cpu-baseline --no-extra 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$tag', round(d['value'],1), {k: round(v,1) for k,v in d['stage_ms_per_step'].items()}, 'nodes', round(d['roofline']['fetch_per_ray']['nodes_visited'],2), 'build', round(d['bvh']['build_ms'],1))" )
done

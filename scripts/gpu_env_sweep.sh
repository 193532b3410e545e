#!/bin/bash
# Cornell bench line per environment setting: each argument is "ENV=VAL[,ENV2=VAL2]" ("" = defaults)
for spec in "$@"; do
  ( IFS=','; for kv in $spec; do [ -n "$kv" ] && export "$kv"; done; unset IFS
    timeout 300 python bench.py --scene ${SCENE:-cornell} --steps 3 --warmup 2 --spp-per-step ${SPP:-128} --no-e2e --no-cpu-baseline --no-instrumented --no-extra 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print('[$spec]', round(d['value'],1), {k: round(x,1) for k,x in d['stage_ms_per_step'].items()})
" )
done

#!/bin/bash
# diagnostic: do the small shadow-ray count differences against the oracle come from FMA contraction / fast division in k_shade?
mkdir -p gpurun_out
cd gpuspectral_b200
for variant in "fmad_off:-prec-div=false -prec-sqrt=false -fmad=false" "ieee:-fmad=false"; do
  tag=${variant%%:*}; flags=${variant#*:}
  touch csrc/wavefront.cu
  make -j8 WF_FLAGS="$flags" > /dev/null 2>&1 || { echo build failed; exit 1; }
  (cd .. && rm -f gpurun_out/parity_full_size.jsonl && timeout 600 python -m pytest tests/test_gpu_parity.py -q -x -k "1080p or 55m_same" 2>&1 | tail -2; python - <<PY
import json
for l in open('gpurun_out/parity_full_size.jsonl'):
    d=json.loads(l)
    if 'ext_rays_gpu' in d: print('$tag', d['test'], 'ext', d['ext_rays_gpu']-d['ext_rays_oracle'], 'shadow', d['shadow_rays_gpu']-d['shadow_rays_oracle'], 'of', d['shadow_rays_oracle'], 'diverged px', d['pixels_diverged'], 'bit-equal', round(d['pixels_bit_equal'],4))
PY
  )
done

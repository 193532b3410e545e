"""Renders a fixed small workload (host-driven loop, so every k_shade launch is visible to ncu) and prints the number of
vertices shaded; scripts/gpu_r2_shade_ops.sh runs it under ncu and divides the FP32 operation counts by that number."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpuspectral_b200 as g  # noqa: E402

scene = sys.argv[1]
with g.Context(0) as ctx:
    ctx.upload(g.Scene.builtin(scene))
    ctx.set_params(max_depth=8)
    ctx.set_profiling(True)          # host loop: plain launches
    ctx.render(640, 360, 0, 0)
    ctx.clear()
    ctx.reset_counters()
    ctx.render(640, 360, 0, 16)
    c = ctx.counters()
    print(json.dumps({"scene": scene, "shaded_vertices": c["shaded_vertices"], "paths": c["paths"], "extension_rays": c["extension_rays"]}))

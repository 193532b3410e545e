"""Times the pieces of one e2e step on the box: per-frame upload (TLAS rebuild), render, read-back."""
import time, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import gpuspectral_b200 as g
import bench

scene = g.Scene.builtin("cornell")
ctx = g.Context(0)
ctx.upload(scene)
print("bvh", ctx.bvh_stats()["build_ms"])
ctx.set_params(max_depth=8)
W, H = 1920, 1080
ctx.render(W, H, 0, 0); ctx.clear()
mesh_map = list(ctx._mesh_map)
host = np.empty((H, W, 4), np.float32)
for spp in (1, 16, 128):
    for it in range(3):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        bench.per_frame_upload(ctx, scene, mesh_map)
        ctx.synchronize(); t1 = time.perf_counter()
        ctx.render(W, H, it * spp, spp)
        t2 = time.perf_counter()
        ctx.synchronize(); t3 = time.perf_counter()
        ctx.read_hdr(out=host); t4 = time.perf_counter()
        print(f"spp {spp}: upload {1e3*(t1-t0):.2f} ms, render call {1e3*(t2-t1):.2f} ms, render wait {1e3*(t3-t2):.2f} ms, read_hdr {1e3*(t4-t3):.2f} ms, bvh build {ctx.bvh_stats()['build_ms']:.2f}")

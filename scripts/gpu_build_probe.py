"""Uploads the instanced field scene twice and prints the BVH build statistics (second build: warm allocator pool)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpuspectral_b200 as g
scene = g.Scene.builtin("field")
ctx = g.Context(0)
for k in range(3):
    t0 = time.perf_counter(); ctx.upload(scene); t1 = time.perf_counter()
    s = ctx.bvh_stats()
    print(f"build {k}: device {s['build_ms']:.1f} ms, wall of upload() {1e3*(t1-t0):.0f} ms, {s['unique_triangles']/s['build_ms']/1e3:.1f} Mtris/s, nodes {s['blas_nodes']}")

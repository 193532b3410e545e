import sys, time
sys.path.insert(0, '.')
import gpuspectral_b200 as g
s = g.Scene.builtin("field")
for k in range(3):
    t = time.perf_counter()
    with g.Context(0) as c:
        c.upload(s)
        st = c.bvh_stats()
        print(k, 'build_ms', round(st['build_ms'], 1), 'wall upload s', round(time.perf_counter() - t, 3))

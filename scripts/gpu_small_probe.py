"""Small levels: what the candidate builds find (GSB_BUILD_TRACE=1 prints every candidate), and the rate with the winner."""
import sys, time
sys.path.insert(0, ".")
import gpuspectral_b200 as g
for name in sys.argv[1:] or ["cornell"]:
    s = g.Scene.builtin(name)
    with g.Context(0) as c:
        c.upload(s); c.upload(s)
        st = c.bvh_stats()
        print(name, "build_ms %.2f" % st["build_ms"], "world nodes", st["world_nodes"], flush=True)
        c.set_params(max_depth=8); c.render(1920, 1080, 0, 2); c.synchronize(); c.clear()
        best = 1e9
        for k in range(3):
            t = time.perf_counter(); c.render(1920, 1080, k * 64, 64); c.synchronize(); best = min(best, time.perf_counter() - t)
        print(name, "Mpaths/s %.1f" % (1920 * 1080 * 64 / best / 1e6), flush=True)

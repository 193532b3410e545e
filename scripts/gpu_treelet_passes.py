"""What each treelet pass buys on the instanced field (and the gallery): SAH cost, build time, paths per second."""
import os, sys, time
sys.path.insert(0, '.')
import gpuspectral_b200 as g
for name, spp in (("field", 16), ("gallery", 32)):
    s = g.Scene.builtin(name)
    for passes in (0, 1, 2):
        os.environ["GSB_TREELET_PASSES"] = str(passes)
        with g.Context(0) as c:
            c.upload(s); c.upload(s)
            st = c.bvh_stats()
            c.set_params(max_depth=8)
            c.render(1920, 1080, 0, 2); c.synchronize(); c.clear()
            best = 1e9
            for k in range(3):
                t = time.perf_counter()
                c.render(1920, 1080, k * spp, spp); c.synchronize()
                best = min(best, time.perf_counter() - t)
            print(name, "passes", passes, "build_ms %.1f" % st["build_ms"], "sah %.1f -> %.1f" % (st["sah_cost_before"], st["sah_cost_after"]),
                  "Mpaths/s %.1f" % (1920 * 1080 * spp / best / 1e6), flush=True)

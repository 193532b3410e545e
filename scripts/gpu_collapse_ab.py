"""Greedy (surface area) against SAH-table cut of the 8-wide collapse: node counts, fetches per ray, paths per second."""
import os, sys, time
sys.path.insert(0, '.')
import gpuspectral_b200 as g
CONFIGS = [("greedy", {"GSB_COLLAPSE_SAH": "0", "GSB_COLLAPSE_SAH_TLAS": "0"}), ("sah blas, greedy tlas", {"GSB_COLLAPSE_SAH": "1", "GSB_COLLAPSE_SAH_TLAS": "0"}),
           ("sah both", {"GSB_COLLAPSE_SAH": "1", "GSB_COLLAPSE_SAH_TLAS": "1"})]
if len(sys.argv) > 1 and sys.argv[1] == "gamma":
    CONFIGS = [("gamma %d" % g_, {"GSB_TREELET_GAMMA": str(g_)}) for g_ in (7, 10, 14, 24)]
if len(sys.argv) > 1 and sys.argv[1] == "small":
    CONFIGS = [("gamma %d passes %d leaf %d" % (g_, p_, l_), {"GSB_TREELET_GAMMA": str(g_), "GSB_TREELET_PASSES": str(p_), "GSB_MAX_LEAF": str(l_)})
               for l_ in (3, 2) for g_, p_ in ((7, 1), (14, 1), (7, 0), (7, 2), (14, 2), (24, 1))]
if len(sys.argv) > 1 and sys.argv[1] == "inst":
    CONFIGS = [("inst cost %.1f" % c, {"GSB_SAH_COST_INST": str(c)}) for c in (0.5, 1.0, 2.0, 4.0, 8.0)] + [("tlas leaf 2", {"GSB_MAX_LEAF_TLAS": "2"})]
if len(sys.argv) > 1 and sys.argv[1] == "costs":
    CONFIGS = [("greedy", {"GSB_COLLAPSE_SAH": "0"})] + [("sah prim %.2f" % c, {"GSB_COLLAPSE_SAH": "1", "GSB_SAH_COST_PRIM": str(c)}) for c in (0.5, 0.75, 1.0, 1.5, 2.5)]
for name, spp in ((("cornell", 64),) if len(sys.argv) > 1 and sys.argv[1] == "small" else (("field", 16),) if len(sys.argv) > 1 and sys.argv[1] == "inst" else (("cornell", 64), ("gallery", 32), ("field", 16))):
    s = g.Scene.builtin(name)
    for label, env in CONFIGS:
        os.environ.update(env)
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
            c.set_instrumented(3); c.reset_counters()
            c.render(1920, 1080, 100, 2); c.synchronize()
            k = c.counters()
            rays = max(1, k.get("rays_counted", 0))
            print(name, label, "build_ms %.1f" % st["build_ms"], "nodes", st["blas_nodes"] + st["world_nodes"], "bytes %.1f MB" % (st["bvh_bytes"] / 1e6),
                  "per ray: nodes %.2f tris %.2f inst %.2f" % (k.get("nodes_visited", 0) / rays, k.get("tris_tested", 0) / rays, k.get("inst_entered", 0) / rays),
                  "sah %.0f -> %.0f" % (st["sah_cost_before"], st["sah_cost_after"]), "Mpaths/s %.1f" % (1920 * 1080 * spp / best / 1e6), flush=True)

"""Freezes outputs of THE REFERENCE'S OWN SHADERS (oracle/_ref/libgsref.so, built from /root/reference by oracle/Makefile)
as golden vectors: tests/golden/ref_vectors.npz.  Run in the build container, where /root/reference exists.

The inputs are regenerated from fixed seeds by tests/refpin.py; only the reference's outputs are stored.  The oracle is
checked against this file bit for bit (tests/test_oracle_cpu.py), so the pin travels to machines without the reference.
Whole-shader runs (render / probes) use the oracle's intersector as traceRayEXT — the ray/triangle search has no source
in the reference — so they pin everything in raygen.rgen / rayhit.rchit / miss.rmiss / shadowmiss.rmiss around it.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

N = 1024


def main():
    import gpuspectral_b200 as g
    import refpin
    from oracle import oracle as orc
    from oracle import ref
    orc.build()
    ref.build(force=True)
    out = {}
    idx, ts = refpin.seeds(4096, 1), refpin.seeds(4096, 2)
    idx[:3], ts[:3] = [0, 1, 2073599], [0, 0, 1023]
    out["rng"] = ref.rng(idx, ts, 8)

    r = ref.RefScene()
    refpin.fill_bsdf_tables(r)
    for t, recs in refpin.BSDF_RECORDS.items():
        for k in range(len(recs)):
            h = (t << 16) | k
            wo, wi, sd = refpin.unit_dirs(N, 10 * t + k), refpin.unit_dirs(N, 1000 + 10 * t + k), refpin.seeds(N, 2000 + 10 * t + k)
            for name, a in zip(("wi", "f", "pdf", "delta", "seed"), r.bsdf_sample(h, wo, sd)):
                out[f"sample_{t}_{k}_{name}"] = a
            for name, a in zip(("f", "pdf", "delta"), r.bsdf_eval(h, wo, wi)):
                out[f"eval_{t}_{k}_{name}"] = a
    frame, a, b = ref.onb(refpin.unit_dirs(N, 77), refpin.unit_dirs(N, 78))
    out["onb_frame"], out["onb_local"], out["onb_world"] = frame, a, b

    cornell = g.Scene.builtin("cornell")
    o, r = orc.OracleScene.from_host_scene(cornell), ref.RefScene.from_host_scene(cornell)
    pos = (np.random.default_rng(5).random((N, 3)) * np.array([2, 2, 2]) - np.array([1, 0, 1])).astype(np.float32)
    for name, a in zip(("lpos", "emission", "pdf", "seed"), r.sample_light(pos, refpin.seeds(N, 6))):
        out[f"light_{name}"] = a
    W, H = 64, 48
    px, py = np.arange(W * H) % W, np.arange(W * H) // W
    out["camera_rays"] = r.camera_rays(W, H, px, py).view(np.float32).reshape(-1, 8)
    img, ext, sh = r.render(o, W, H, 0, 4)
    out["cornell_image"], out["cornell_rays"] = img, np.array([ext, sh], np.uint64)
    probes = []
    for (x, y) in refpin.PROBE_PIXELS:
        for t in (0, 3):
            _, b = r.path_probe(o, W, H, x, y, t, 64)
            probes.append(b)
    out["cornell_probe_lengths"] = np.array([len(b) for b in probes], np.uint32)
    out["cornell_probes"] = np.concatenate(probes).view(np.uint32).reshape(-1, 12)

    o, r = orc.OracleScene(), ref.RefScene()
    refpin.all_types_scene(g, o, r)
    img, ext, sh = r.render(o, W, H, 0, 2)
    out["alltypes_image"], out["alltypes_rays"] = img, np.array([ext, sh], np.uint64)

    path = os.path.join(HERE, "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()

"""Generates the committed fixtures of tests/golden/ (run in the build container, where /root/reference exists).

tungsten_cornell_64.npy   64x64x3 float32 box-filtered copy of the reference repo's
                          src/GPUSpectral/assets/scenes/cornell-box/TungstenRender.exr (1024x1024, CC0 Bitterli scene,
                          rendered by Tungsten - an independent unbiased renderer, NOT by GPUSpectral): loose sanity only.
tungsten_cornell_256_f16.npy   the same image box-filtered to 256x256, float16 (converged-render comparison on the GPU box).
oracle_cornell_48x48_4spp_d8.npy   the oracle's own image for a fixed input: pins the oracle against accidental edits.
oracle_paths.json         per-bounce records of a few fixed paths (instance, primitive, t bits, seed) from the oracle.
rng_kat.json              RNG known answers derived from the GLSL text in SURVEY.md 8(c) (tea -> pcgHash -> draws).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

REF_SCENE = "/root/reference/src/GPUSpectral/assets/scenes/cornell-box"


def tungsten():
    os.environ["OPENCV_IO_ENABLE_OPENEXR"] = "1"
    import cv2
    img = cv2.imread(os.path.join(REF_SCENE, "TungstenRender.exr"), cv2.IMREAD_UNCHANGED)[..., :3][..., ::-1]  # BGR -> RGB
    assert img.shape[:2] == (1024, 1024)
    small = img.astype(np.float64).reshape(64, 16, 64, 16, 3).mean(axis=(1, 3)).astype(np.float32)
    np.save(os.path.join(HERE, "tungsten_cornell_64.npy"), small)
    # 256 x 256 (4 x 4 box filter), float16: the yardstick of the converged-image figure (tests/test_gpu_parity.py)
    quarter = img.astype(np.float64).reshape(256, 4, 256, 4, 3).mean(axis=(1, 3)).astype(np.float16)
    np.save(os.path.join(HERE, "tungsten_cornell_256_f16.npy"), quarter)


def oracle_fixtures():
    import gpuspectral_b200 as g
    from oracle import oracle as orc
    orc.build()
    s = g.Scene.load_xml(os.path.join(REF_SCENE, "scene.xml"))
    o = orc.OracleScene.from_host_scene(s)
    o.set_params(max_depth=8)
    img, c, _ = o.render(48, 48, 0, 4)
    np.save(os.path.join(HERE, "oracle_cornell_48x48_4spp_d8.npy"), img)
    paths = []
    for (x, y, ts) in [(24, 24, 0), (10, 40, 3), (47, 0, 1), (0, 47, 2), (30, 12, 5)]:
        rgb, b = o.path_probe(48, 48, x, y, ts)
        paths.append({"x": x, "y": y, "timestamp": ts, "rgb_bits": [int(v) for v in rgb.view(np.uint32)],
                      "bounces": [{"inst": int(r["inst"]), "prim": int(r["prim"]), "t_bits": int(np.float32(r["t"]).view(np.uint32)),
                                   "seed": int(r["seed"]), "shadow_cast": int(r["shadow_cast"]), "shadow_visible": int(r["shadow_visible"])} for r in b]})
    json.dump({"scene": "cornell-box/scene.xml", "width": 48, "height": 48, "max_depth": 8, "counters": c, "paths": paths},
              open(os.path.join(HERE, "oracle_paths.json"), "w"), indent=1)


def rng_kat():
    # (pixel index, timestamp) -> tea, seed = pcgHash(tea), first draws of randPcg()
    kat = [
        {"idx": 0, "ts": 0, "tea": 0x5df5f2bf, "seed": 0x66f8bb0c, "draws": [0xe4737e43, 0x8a69cf8e, 0xab953492]},
        {"idx": 1, "ts": 0, "tea": 0xc09848f2, "seed": 0x3a910142, "draws": [0x223461c8, 0x0739263e, 0x4f8dd381]},
        {"idx": 0, "ts": 1, "tea": 0x5d8714d7, "seed": 0xc9a3ba20, "draws": [0xb168c2ce, 0xf2630107, 0xafac66a3]},
        {"idx": 32896, "ts": 0, "tea": 0x6645fac4, "seed": 0x7b81d2a0, "draws": [0x63096906]},
        {"idx": 1037760, "ts": 7, "tea": 0x8c3d8ccd, "seed": 0xc509fb3d, "draws": [0x827ae01d]},
        {"idx": 2073599, "ts": 1023, "tea": 0x7fa03ab2, "seed": 0xd8b9587c, "draws": [0xb2036de8]},
    ]
    json.dump(kat, open(os.path.join(HERE, "rng_kat.json"), "w"), indent=1)


if __name__ == "__main__":
    tungsten()
    oracle_fixtures()
    rng_kat()
    print("fixtures written to", HERE)

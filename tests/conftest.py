import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this environment")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def gsb():
    import gpuspectral_b200
    return gpuspectral_b200


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.build()
    return oracle


SCENES = {}


def host_scene(name):
    """Session-cached host scenes (built by the C++ host layer)."""
    import gpuspectral_b200 as g
    if name not in SCENES:
        if name == "cornell":
            SCENES[name] = g.Scene.builtin("cornell")
        elif name == "gallery":
            SCENES[name] = g.Scene.builtin("gallery")
        elif name == "field_small":
            SCENES[name] = g.Scene.builtin("field", 1234, 3, 3, 3)   # 3 meshes x 1280 tris, 9 instances
        elif name == "field_medium":
            SCENES[name] = g.Scene.builtin("field", 77, 4, 5, 5)     # 4 meshes x 20480 tris, 25 instances
        else:
            raise KeyError(name)
    return SCENES[name]


@pytest.fixture(scope="session")
def scenes():
    return host_scene


def random_rays(scene, n, seed, spread=1.5):
    """Rays with origins in / around the scene bounds and random directions, plus rays aimed at vertices and edges."""
    rng = np.random.default_rng(seed)
    pts = []
    inst, ni = scene.instances
    for i in range(ni):
        T = np.array(inst[i].transform, np.float64).reshape(4, 4).T
        p, _ = scene.mesh(inst[i].mesh)
        if len(p) > 3000:
            p = p[rng.choice(len(p), 3000, replace=False)]
        pts.append((np.c_[p.astype(np.float64), np.ones(len(p))] @ T.T)[:, :3])
    pts = np.concatenate(pts)
    lo, hi = pts.min(0), pts.max(0)
    c, e = 0.5 * (lo + hi), 0.5 * (hi - lo) * spread
    o = (c + (rng.random((n, 3)) * 2 - 1) * e)
    d = rng.normal(size=(n, 3))
    # a third of the rays are aimed exactly at scene vertices (shared edges / corners)
    k = n // 3
    tgt = pts[rng.integers(0, len(pts), k)]
    d[:k] = tgt - o[:k]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    import gpuspectral_b200 as g
    return g.make_rays(o.astype(np.float32), d.astype(np.float32), 0.0, 1e10)

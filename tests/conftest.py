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


class ArchiveScene:
    """A scene archive (oracle.OracleScene.dump_host_scene -> tests/golden/scene_*.npz) with the host Scene's read interface
    (mesh_count, mesh, bsdfs, lights, instances, camera), so that Context.upload / OracleScene.from_host_scene /
    random_rays take it like a Scene.  Lets scenes of the reference checkout travel to the GPU box as flat records."""

    def __init__(self, path):
        import ctypes as C

        import gpuspectral_b200._native as N
        z = np.load(path)
        self._z = z
        self.mesh_count = int(z["mesh_count"][0])
        ni, nl = int(z["instances_count"][0]), int(z["lights_count"][0])
        self._inst = (N.Instance * max(1, ni))()
        C.memmove(self._inst, z["instances"].tobytes(), ni * C.sizeof(N.Instance))
        self._ni = ni
        self._lights = (N.TriangleLight * max(1, nl))()
        C.memmove(self._lights, z["lights"].tobytes(), nl * C.sizeof(N.TriangleLight))
        self._nl = nl
        from oracle import oracle as orc
        self._tex, self._nt, self._tex_keep = orc.archive_textures(z)

    def mesh(self, i):
        return self._z[f"mesh{i}_p"], self._z[f"mesh{i}_n"]

    def mesh_uv(self, i):
        return self._z[f"mesh{i}_uv"] if f"mesh{i}_uv" in self._z else None

    @property
    def textures(self):
        import ctypes as C

        import gpuspectral_b200._native as N
        return C.cast(self._tex, C.POINTER(N.Texture)), self._nt

    def bsdfs(self, t):
        return self._z[f"bsdf{t}"].tobytes(), int(self._z[f"bsdf{t}_count"][0])

    @property
    def lights(self):
        return self._lights, self._nl

    @property
    def instances(self):
        import ctypes as C
        copy = type(self._inst)()
        C.memmove(copy, self._inst, C.sizeof(copy))
        return copy, self._ni

    @property
    def camera(self):
        return self._z["eye"], self._z["to_world"], float(self._z["fov"][0])

    @property
    def instanced_triangles(self):
        return sum(len(self._z[f"mesh{self._inst[i].mesh}_p"]) // 3 for i in range(self._ni))

"""Cross-check of the product's own Mitsuba-XML reader + loader (gpuspectral_b200/host/xml.cpp, loader.cpp) against the
REFERENCE'S OWN parser: external/TinyParser-Mitsuba (+ its tinyxml2), the library engine/Loader.cpp:253-349 reads scenes
with, compiled from /root/reference into oracle/_ref/libtpm.so (oracle/Makefile, oracle/ref/tpm_glue.cpp).

On every scene.xml the reference ships (7 files, 123 shapes): per shape, in file order, the instance matrix equals the
parser's `to_world`; the BSDF the loader attached carries the parameters the parser read, under the conventions of
engine/Loader.cpp:145-234 (restated here: alpha * sqrt 2, R0 from int_ior, defaults); area emitters' radiance; the
camera's fov and matrix (Loader.cpp:323-336, Camera.cpp:20-45).  Most of those files reference meshes the reference never
shipped: the loader's placeholder mode (GSB_LOADER_MISSING_MESH) keeps the shape list complete.
Needs the reference checkout (build container only): skipped elsewhere."""
import ctypes as C
import glob
import json
import math
import os
import struct

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = sorted(glob.glob("/root/reference/src/GPUSpectral/assets/scenes/*/scene.xml"))
LIB = os.path.join(ROOT, "oracle", "_ref", "libtpm.so")

pytestmark = pytest.mark.skipif(not SCENES, reason="the reference checkout is not on this machine")

OT_BSDF, OT_EMITTER, OT_SENSOR, OT_SHAPE = 1, 2, 9, 10


def _dump(path):
    if not os.path.exists(LIB):
        import subprocess
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "_ref/libtpm.so"], check=True, capture_output=True)
    L = C.CDLL(LIB)
    L.tpm_dump_scene.restype = C.c_long
    L.tpm_dump_scene.argtypes = [C.c_char_p, C.c_char_p, C.c_long]
    buf = C.create_string_buffer(1 << 23)
    n = L.tpm_dump_scene(path.encode(), buf, len(buf))
    assert n > 0, buf.value
    return json.loads(buf.value)


def _leaf_bsdf(obj):
    """The innermost bsdf of a shape's material chain (twosided / bumpmap wrappers hold it as a child) + twosided flag."""
    two = False
    cur = obj
    while True:
        if cur["plugin"] == "twosided":
            two = True
        inner = [c for c in cur["children"] + list(cur["named"].values()) if c["type"] == OT_BSDF]
        if cur["plugin"] in ("twosided", "bumpmap", "mask") and inner:
            cur = inner[0]
            continue
        return cur, two


def _num(obj, name, default):
    p = obj["props"].get(name)
    return float(p["v"]) if p and p["t"] in ("num", "int") else default


def _rgb(obj, name, default):
    p = obj["props"].get(name)
    if p and p["t"] == "rgb":
        return [float(x) for x in p["v"]]
    if p and p["t"] in ("num", "int"):
        return [float(p["v"])] * 3
    return default


RECORD = {0: "3fi", 1: "2f", 2: "2f", 4: "10fi", 3: "6f", 7: "7fi", 8: "3fi"}


def _expected_record(bsdf):
    """engine/Loader.cpp:145-234 (and the rough-dielectric extension): plugin -> (BSDF type, parameter tuple)."""
    t = bsdf["plugin"]
    f32 = lambda x: struct.unpack("f", struct.pack("f", x))[0]
    r0 = lambda ior: f32(f32(f32(ior - 1.0) / f32(ior + 1.0)) ** 2)
    if t == "diffuse":
        return 0, _rgb(bsdf, "reflectance", [0.5, 0.5, 0.5]) + [0]
    if t == "dielectric":
        return 1, [_num(bsdf, "int_ior", 1.5046), _num(bsdf, "ext_ior", 1.000277)]
    if t == "conductor":
        return 2, [_num(bsdf, "eta", 0.0), 1.0]
    if t == "plastic":
        ior = f32(_num(bsdf, "int_ior", 1.3))
        return 3, _rgb(bsdf, "diffuse_reflectance", [0.5] * 3) + [ior, 1.0, r0(ior)]
    if t == "roughconductor":
        return 4, _rgb(bsdf, "eta", [0.0] * 3) + _rgb(bsdf, "k", [0.0] * 3) + _rgb(bsdf, "specular_reflectance", [1.0] * 3) + \
            [f32(f32(math.sqrt(2)) * f32(_num(bsdf, "alpha", 0.1))), 0]
    if t == "roughplastic":
        ior = f32(_num(bsdf, "int_ior", 1.3))
        return 7, _rgb(bsdf, "diffuse_reflectance", [0.5] * 3) + [ior, 1.0, r0(ior), f32(f32(math.sqrt(2.0)) * f32(_num(bsdf, "alpha", 0.1))), 0]
    if t == "roughdielectric":
        return 8, [_num(bsdf, "int_ior", 1.5046), _num(bsdf, "ext_ior", 1.000277), _num(bsdf, "alpha", 0.1), 0]
    return None, None


@pytest.mark.parametrize("path", SCENES, ids=[p.split("/")[-2] for p in SCENES])
def test_loader_equals_the_reference_parser(gsb, path, monkeypatch):
    monkeypatch.setenv("GSB_LOADER_MISSING_MESH", "placeholder")
    ref = _dump(path)
    ours = gsb.Scene.load_xml(path)
    shapes = [c for c in ref["children"] if c["type"] == OT_SHAPE and c["plugin"] in ("obj", "rectangle", "cube", "disk")]
    inst, ni = ours.instances
    assert ni == len(shapes), "shape list differs"
    checked = {"matrices": 0, "bsdfs": 0, "emitters": 0}
    for i, sh in enumerate(shapes):
        # instance matrix: the parser stores row-major, gsb_instance column-major
        tw = sh["props"].get("to_world")
        M = np.array(tw["v"], np.float64).reshape(4, 4) if tw else np.eye(4)
        got = np.array(list(inst[i].transform), np.float64).reshape(4, 4).T
        assert np.allclose(got, M, rtol=1e-5, atol=1e-5 * max(1.0, np.abs(M).max())), (i, got, M)
        checked["matrices"] += 1
        # material chain
        mats = [c for c in sh["children"] + list(sh["named"].values()) if c["type"] == OT_BSDF]
        if mats:
            leaf, two = _leaf_bsdf(mats[0])
            btype, want = _expected_record(leaf)
            handle = inst[i].bsdf
            if btype is not None:
                assert handle >> 16 == btype, (i, leaf["plugin"], handle >> 16)
                raw, n = ours.bsdfs(btype)
                stride = gsb.BSDF_STRIDES[btype]
                rec = struct.unpack(RECORD[btype], raw[(handle & 0xffff) * stride:(handle & 0xffff) * stride + stride])
                textured = any(c["type"] == 12 for c in leaf["children"] + list(leaf["named"].values()))
                if not textured:
                    assert np.allclose(rec, want, rtol=2e-6, atol=1e-7), (i, leaf["plugin"], rec, want)
                checked["bsdfs"] += 1
            assert bool(inst[i].twofaced) == two, (i, two)
        # area emitter
        ems = [c for c in sh["children"] + list(sh["named"].values()) if c["type"] == OT_EMITTER]
        want_em = _rgb(ems[0], "radiance", [1.0, 1.0, 1.0]) if ems else [0.0, 0.0, 0.0]
        assert np.allclose(list(inst[i].emission), want_em, rtol=1e-6), (i, list(inst[i].emission), want_em)
        checked["emitters"] += len(ems)
    # camera: fov in degrees -> radians (Loader.cpp:323-336), matrix row-major -> column-major
    sensor = [c for c in ref["children"] if c["type"] == OT_SENSOR][0]
    eye, to_world, fov = ours.camera
    assert fov == pytest.approx(math.radians(_num(sensor, "fov", 0.0)), rel=1e-6)
    Mc = np.array(sensor["props"]["to_world"]["v"], np.float64).reshape(4, 4)
    assert np.allclose(eye, Mc[:3, 3], rtol=1e-5, atol=1e-5)
    assert checked["matrices"] == len(shapes) and checked["bsdfs"] > 0

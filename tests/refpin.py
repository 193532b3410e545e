"""Shared helpers of the tests that pin the oracle to the reference's own shader text (oracle/_ref) and of
tests/golden/make_ref_vectors.py, which freezes outputs of that library as golden vectors."""
import ctypes as C
import struct

import numpy as np

# One parameter table per reference BSDF type (rayhit.rchit:22-70 / Scene.h:29-82 layouts), several records each.
BSDF_RECORDS = {
    0: [struct.pack("3fi", 0.63, 0.065, 0.05, 0), struct.pack("3fi", 0.725, 0.71, 0.68, 0)],
    1: [struct.pack("2f", 1.5, 1.0), struct.pack("2f", 1.3, 1.0)],
    2: [struct.pack("2f", 0.0, 1.0), struct.pack("2f", 1.8, 1.0)],
    3: [struct.pack("6f", 1.0, 0.38, 0.013, 1.3, 1.0, 0.017), struct.pack("6f", 0.2, 0.4, 0.9, 1.5, 1.0, 0.04)],
    4: [struct.pack("10fi", 1.657, 0.880, 0.521, 9.224, 6.270, 4.837, 1, 1, 1, 0.1414, 0),
        struct.pack("10fi", 0.143, 0.375, 1.442, 3.983, 2.386, 1.603, 0.9, 0.9, 0.9, 0.4243, 0)],
    5: [struct.pack("4f", 0.5, 0.5, 0.5, 0.04), struct.pack("4f", 0.8, 0.3, 0.1, 0.08)],
    6: [struct.pack("5f", 0.5, 0.3, 0.2, 0.04, 0.2), struct.pack("5f", 0.2, 0.6, 0.7, 0.06, 0.07)],
    7: [struct.pack("7fi", 1.0, 0.58, 0.13, 1.3, 1.0, 0.017, 0.0707, 0), struct.pack("7fi", 0.43, 0.58, 1.0, 1.5, 1.0, 0.04, 0.4243, 0)],
}


def unit_dirs(n, seed):
    rng = np.random.default_rng(seed)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d = d.astype(np.float32)
    # the exact poles and the horizon are the branches of onbCreate / the grazing limits of the rough lobes
    d[:6] = [[0, 0, 1], [0, 0, -1], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0.6, 0.8, 0]]
    return d


def seeds(n, seed):
    return np.random.default_rng(seed).integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)


def bits_equal(a, b):
    """Bit-for-bit equality of two arrays; float NaNs compare equal to NaNs (payload bits are not part of the contract)."""
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind == "f":
        return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))
    return a == b


def fill_bsdf_tables(*scenes):
    for s in scenes:
        for t, recs in BSDF_RECORDS.items():
            s.set_bsdfs(t, b"".join(recs), len(recs))


def all_types_scene(gsb, *targets):
    """The BSDF gallery with every one of the reference's eight BSDF types in use: the gallery's rough-dielectric ball
    (an extension with no reference source) becomes a SmoothFloor ball, the aluminium ball a RoughFloor ball.  Feeds the
    same flat records to every target (OracleScene / RefScene / Context-like objects with the flat-record setters)."""
    sc = gsb.Scene.builtin("gallery")
    inst, ni = sc.instances
    tables = {t: list(sc.bsdfs(t)) for t in range(8)}
    tables[5] = [b"".join(BSDF_RECORDS[5]), 2]
    tables[6] = [b"".join(BSDF_RECORDS[6]), 2]
    swapped = 0
    for i in range(ni):
        t = inst[i].bsdf >> 16
        if t == 8:
            inst[i].bsdf = (5 << 16) | 0
            inst[i].twofaced = 1
            swapped += 1
        elif t == 4 and (inst[i].bsdf & 0xffff) == 1:
            inst[i].bsdf = (6 << 16) | 1
            swapped += 1
    assert swapped == 2, "the gallery no longer has the two balls this helper rewrites"
    lights, nl = sc.lights
    for tgt in targets:
        for m in range(sc.mesh_count):
            tgt.add_mesh(*sc.mesh(m))
        for t in range(8):
            tgt.set_bsdfs(t, tables[t][0], tables[t][1])
        tgt.set_lights(lights, nl)
        tgt.set_instances(inst, ni)
        tgt.set_camera(*sc.camera)
    return sc


PROBE_PIXELS = [(5, 7), (31, 20), (48, 40), (60, 11), (17, 33), (40, 25), (2, 46), (63, 47)]

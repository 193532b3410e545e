"""glTF 2.0 input (host/gltf.cpp) — CPU only: a hand-built scene written as .gltf (external .bin + data URI) and as .glb
must come out as the same host Scene: meshes un-indexed, node hierarchy composed, materials mapped, emissive triangles
turned into lights, the camera upright in the reference's ray-generation formula (checked through the oracle)."""
import base64
import json
import struct

import numpy as np
import pytest


def _build(tmp_path, glb):
    # buffer 0: floor quad (4 vertices + normals, u16 indices); buffer 1: light triangle (positions only) + tetrahedron (u32 indices)
    quad_p = np.float32([[-2, 0, -2], [2, 0, -2], [2, 0, 2], [-2, 0, 2]])
    quad_n = np.float32([[0, 1, 0]] * 4)
    quad_i = np.uint16([0, 2, 1, 0, 3, 2])
    b0 = quad_p.tobytes() + quad_n.tobytes() + quad_i.tobytes()
    tri_p = np.float32([[-0.5, 0, -0.5], [0, 0, 0.5], [0.5, 0, -0.5]])        # normal +y; the node turns it to face down
    tet_p = np.float32([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]])
    tet_i = np.uint32([0, 2, 1, 0, 1, 3, 0, 3, 2, 1, 2, 3])
    b1 = tri_p.tobytes() + tet_p.tobytes() + tet_i.tobytes()
    views = [
        {"buffer": 0, "byteOffset": 0, "byteLength": 48}, {"buffer": 0, "byteOffset": 48, "byteLength": 48},
        {"buffer": 0, "byteOffset": 96, "byteLength": 12},
        {"buffer": 1, "byteOffset": 0, "byteLength": 36}, {"buffer": 1, "byteOffset": 36, "byteLength": 48},
        {"buffer": 1, "byteOffset": 84, "byteLength": 48},
    ]
    acc = [
        {"bufferView": 0, "componentType": 5126, "count": 4, "type": "VEC3"},
        {"bufferView": 1, "componentType": 5126, "count": 4, "type": "VEC3"},
        {"bufferView": 2, "componentType": 5123, "count": 6, "type": "SCALAR"},
        {"bufferView": 3, "componentType": 5126, "count": 3, "type": "VEC3"},
        {"bufferView": 4, "componentType": 5126, "count": 4, "type": "VEC3"},
        {"bufferView": 5, "componentType": 5125, "count": 12, "type": "SCALAR"},
    ]
    doc = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 1, 4]}],
        "nodes": [
            {"name": "floor", "mesh": 0},
            {"name": "rig", "translation": [0, 3, 0], "children": [2, 3]},
            {"name": "light", "mesh": 1, "rotation": [1, 0, 0, 0]},                         # 180 deg about x: faces down
            {"name": "glass", "mesh": 2, "translation": [1, -2.5, 0], "scale": [0.5, 0.5, 0.5]},
            {"name": "cam", "camera": 0, "translation": [0, 1, 5]},
        ],
        "cameras": [{"type": "perspective", "perspective": {"yfov": 0.6, "aspectRatio": 16 / 9, "znear": 0.1}}],
        "meshes": [
            {"name": "quad", "primitives": [{"attributes": {"POSITION": 0, "NORMAL": 1}, "indices": 2, "material": 0}]},
            {"name": "tri", "primitives": [{"attributes": {"POSITION": 3}, "material": 1}]},
            {"name": "tet", "primitives": [{"attributes": {"POSITION": 4}, "indices": 5, "material": 2},
                                           {"attributes": {"POSITION": 4}, "indices": 5, "material": 3, "mode": 1}]},
        ],
        "materials": [
            {"pbrMetallicRoughness": {"baseColorFactor": [0.8, 0.2, 0.2, 1], "metallicFactor": 0, "roughnessFactor": 1}},
            {"emissiveFactor": [1, 1, 0.5], "extensions": {"KHR_materials_emissive_strength": {"emissiveStrength": 10}},
             "pbrMetallicRoughness": {"metallicFactor": 0}},
            {"pbrMetallicRoughness": {"metallicFactor": 0, "roughnessFactor": 0.5},
             "extensions": {"KHR_materials_transmission": {"transmissionFactor": 1}, "KHR_materials_ior": {"ior": 1.33}}},
            {"pbrMetallicRoughness": {"roughnessFactor": 0.2}},
        ],
        "bufferViews": views, "accessors": acc,
    }
    if glb:
        for v in views:
            if v["buffer"] == 1:
                v["byteOffset"] += len(b0)
            v["buffer"] = 0
        blob = b0 + b1
        doc["buffers"] = [{"byteLength": len(blob)}]
        js = json.dumps(doc).encode()
        js += b" " * (-len(js) % 4)
        blob += b"\0" * (-len(blob) % 4)
        path = tmp_path / "scene.glb"
        body = struct.pack("<II", len(js), 0x4E4F534A) + js + struct.pack("<II", len(blob), 0x004E4942) + blob
        path.write_bytes(b"glTF" + struct.pack("<II", 2, 12 + len(body)) + body)
    else:
        (tmp_path / "floor.bin").write_bytes(b0)
        doc["buffers"] = [{"uri": "floor.bin", "byteLength": len(b0)},
                          {"uri": "data:application/octet-stream;base64," + base64.b64encode(b1).decode(), "byteLength": len(b1)}]
        path = tmp_path / "scene.gltf"
        path.write_text(json.dumps(doc, indent=1))
    return path


@pytest.mark.parametrize("glb", [False, True])
def test_gltf_scene_maps_onto_the_host_scene(gsb, orc, tmp_path, glb):
    s = gsb.Scene.load(_build(tmp_path, glb))
    assert s.mesh_count == 3
    pos, nrm = s.mesh(0)
    assert pos.shape == (6, 3) and np.allclose(nrm, [0, 1, 0])                    # indexed quad -> 2 un-indexed triangles
    assert np.array_equal(pos[1], [2, 0, 2])                                       # index order kept
    pos, nrm = s.mesh(1)
    assert pos.shape == (3, 3) and np.allclose(np.abs(nrm), [0, 1, 0])             # no NORMAL: face normal
    assert s.mesh(2)[0].shape == (12, 3)
    inst, n = s.instances
    assert n == 3                                                                  # the mode-1 primitive is skipped
    assert any("mode 1" in w for w in s.warnings)
    T = [np.array(inst[i].transform, np.float32).reshape(4, 4).T for i in range(n)]
    assert np.allclose(T[0], np.eye(4))
    assert np.allclose(T[1], [[1, 0, 0, 0], [0, -1, 0, 3], [0, 0, -1, 0], [0, 0, 0, 1]], atol=1e-6)      # parent T * child R
    assert np.allclose(T[2], [[0.5, 0, 0, 1], [0, 0.5, 0, 0.5], [0, 0, 0.5, 0], [0, 0, 0, 1]], atol=1e-6)
    # materials: diffuse floor (+ the light's black-ish default), rough dielectric for the transmissive tetrahedron
    assert [inst[i].bsdf >> 16 for i in range(n)] == [gsb.BSDF_DIFFUSE, gsb.BSDF_DIFFUSE, 8]
    raw, cnt = s.bsdfs(8)
    assert cnt == 1 and np.allclose(struct.unpack("3fi", bytes(raw)[:16])[:3], [1.33, 1.0, 0.25])
    raw, cnt = s.bsdfs(gsb.BSDF_DIFFUSE)
    assert np.allclose(struct.unpack("3f", bytes(raw)[:12]), [0.8, 0.2, 0.2])
    assert list(inst[1].emission) == [10.0, 10.0, 5.0] and inst[2].twofaced == 0
    lights, nl = s.lights
    assert nl == 1
    lp = np.array(lights[0].positions, np.float32)[:, :3]
    assert np.allclose(lp[:, 1], 3.0) and np.allclose(sorted(lp[:, 0]), [-0.5, 0, 0.5])
    # camera: at (0, 1, 5) looking down -Z, upright, fov on the larger axis
    w, h = s.film
    assert (w, h) == (1920, 1080)
    o = orc.OracleScene.from_host_scene(s)
    r = o.camera_rays(w, h, np.uint32([w // 2, w // 2, w - 1]), np.uint32([h // 2, 0, h // 2]))
    assert np.allclose(r["origin"], [0, 1, 5])
    assert np.allclose(r["dir"][0], [0, 0, -1], atol=2e-3)
    assert r["dir"][1][1] > 0.25 and abs(r["dir"][1][0]) < 2e-3                    # top row looks up
    assert r["dir"][2][0] > 0.4                                                    # right column looks right
    assert np.isclose(np.arctan2(r["dir"][1][1], -r["dir"][1][2]), 0.3, atol=2e-3)  # yfov / 2
    # and it renders: the floor is lit by the triangle light
    o.set_params(max_depth=4)
    img, counters, _ = o.render(64, 36, 0, 4)
    assert np.isfinite(img).all() and img[..., :3].mean() > 1e-3 and counters["shadow_rays"] > 0


def test_gltf_errors_are_reported(gsb, tmp_path):
    p = tmp_path / "broken.gltf"
    p.write_text('{"asset": {"version": "2.0"}, "scenes": [{"nodes": [0]}], "nodes": [{"mesh": 3}]}')
    with pytest.raises(gsb.GsbError):
        gsb.Scene.load(p)
    with pytest.raises(gsb.GsbError):
        gsb.Scene.load(tmp_path / "missing.glb")


@pytest.mark.gpu
def test_gltf_scene_renders_like_the_oracle(gsb, orc, tmp_path):
    """The glTF scene (diffuse floor, emissive triangle, rough-dielectric tetrahedron) through the CUDA path: exact hits
    on camera rays and an equal-seed image that agrees with the oracle per pixel."""
    s = gsb.Scene.load(_build(tmp_path, True))
    o = orc.OracleScene.from_host_scene(s)
    w, h, spp = 160, 90, 16
    with gsb.Context(0) as ctx:
        ctx.upload(s)
        ys, xs = np.mgrid[0:h, 0:w]
        rays = o.camera_rays(w, h, xs.ravel(), ys.ravel())
        got, ref = ctx.trace_batch(rays), o.trace_batch(rays, mode=1)
        for k in ("inst", "prim", "t", "u", "v"):
            assert np.array_equal(got[k], ref[k]), k
        assert (got["inst"] != gsb.NO_HIT).mean() > 0.2
        ctx.set_params(max_depth=6)
        o.set_params(max_depth=6)
        ctx.render(w, h, 0, 0); ctx.clear(); ctx.render(w, h, 0, spp)
        img = ctx.read_hdr()
        want, oc, _ = o.render(w, h, 0, spp)
        assert ctx.counters()["paths"] == w * h * spp == oc["paths"]
        diff = np.abs(img[..., :3] - want[..., :3]).max(axis=2)
        close = diff <= 2e-3 * np.maximum(want[..., :3].max(axis=2), 0.05)
        assert close.mean() > 0.98 and img[..., :3].mean() > 1e-3


def _one_triangle(tmp_path, emissive):
    """The smallest scene the loaders accept: one triangle facing the camera, lit by nothing or by itself."""
    p = np.float32([[-1, -1, 0], [1, -1, 0], [0, 1, 0]])
    mat = {"pbrMetallicRoughness": {"baseColorFactor": [0.5, 0.5, 0.5, 1], "metallicFactor": 0}}
    if emissive:
        mat["emissiveFactor"] = [1, 1, 1]
    doc = {
        "asset": {"version": "2.0"}, "scene": 0, "scenes": [{"nodes": [0, 1]}],
        "nodes": [{"mesh": 0}, {"camera": 0, "translation": [0, 0, 4]}],
        "cameras": [{"type": "perspective", "perspective": {"yfov": 0.8, "aspectRatio": 1.0, "znear": 0.1}}],
        "meshes": [{"primitives": [{"attributes": {"POSITION": 0}, "material": 0}]}],
        "materials": [mat],
        "buffers": [{"uri": "data:application/octet-stream;base64," + base64.b64encode(p.tobytes()).decode(), "byteLength": 36}],
        "bufferViews": [{"buffer": 0, "byteOffset": 0, "byteLength": 36}],
        "accessors": [{"bufferView": 0, "componentType": 5126, "count": 3, "type": "VEC3"}],
    }
    path = tmp_path / ("lit.gltf" if emissive else "dark.gltf")
    path.write_text(json.dumps(doc))
    return path


@pytest.mark.gpu
@pytest.mark.parametrize("emissive", [False, True])
def test_smallest_scenes_match_the_oracle(gsb, orc, tmp_path, emissive):
    """Edge cases of SURVEY 8(a) A9: a scene without any light (numLights == 0, a modulo by zero in the shader; both sides
    still make the three draws and add nothing) and a one-triangle scene whose only surface is its only light. One mesh,
    one instance, a node-less BVH: hits exact, ray counts exact, images equal."""
    s = gsb.Scene.load(_one_triangle(tmp_path, emissive))
    assert s.lights[1] == (1 if emissive else 0) and s.instances[1] == 1 and s.instanced_triangles == 1
    o = orc.OracleScene.from_host_scene(s)
    w = h = 96
    with gsb.Context(0) as ctx:
        ctx.upload(s)
        ys, xs = np.mgrid[0:h, 0:w]
        rays = o.camera_rays(w, h, xs.ravel(), ys.ravel())
        got, ref = ctx.trace_batch(rays), o.trace_batch(rays, mode=1)
        for k in ("inst", "prim", "t", "u", "v"):
            assert np.array_equal(got[k], ref[k]), k
        assert 0.05 < (got["inst"] != gsb.NO_HIT).mean() < 0.5
        ctx.set_params(max_depth=5)
        o.set_params(max_depth=5)
        ctx.render(w, h, 0, 0); ctx.clear(); ctx.render(w, h, 0, 8)
        img, gc = ctx.read_hdr(), ctx.counters()
        want, oc, _ = o.render(w, h, 0, 8)
        for k in ("paths", "extension_rays"):
            assert gc[k] == oc[k], k
        assert np.isfinite(img).all()
        assert np.isclose(img, want, rtol=1e-5, atol=1e-6).all(axis=2).mean() > 0.995
        hits = gc["extension_rays"] - gc["paths"]
        if emissive:
            # the light sample lies in the plane of the shaded point, so the gate `dot(N, L) > 0` (rayhit.rchit:736) is the
            # sign of a rounding error and the two sides cast different numbers of (fruitless) shadow rays
            assert 0 < gc["shadow_rays"] <= hits and 0 < oc["shadow_rays"] <= hits
            assert img[..., :3].max() == pytest.approx(1.0, rel=1e-5)      # the triangle shows its own radiance
        else:
            assert not img[..., :3].any() and gc["shadow_rays"] == 0 == oc["shadow_rays"]

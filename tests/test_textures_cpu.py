"""Reflectance textures (extension; dead code in the reference, engine/Loader.cpp:122-143): texture coordinates from OBJ files
and the built-in shapes, the image decoders (own PNG inflate, PPM, PFM), the Mitsuba <texture> elements, and the oracle's
texture lookup against a numpy restatement."""
import ctypes as C
import os
import struct
import zlib

import numpy as np
import pytest

_XML = """<scene version="0.6.0"><sensor type="perspective"><float name="fov" value="40"/>
<transform name="to_world"><lookat origin="0,0,4" target="0,0,0" up="0,1,0"/></transform></sensor>
%s
<shape type="rectangle"><emitter type="area"><rgb name="radiance" value="10,10,10"/></emitter>
<transform name="to_world"><scale value="0.5"/><rotate x="1" angle="90"/><translate y="1.9"/></transform>
<bsdf type="diffuse"><rgb name="reflectance" value="0,0,0"/></bsdf></shape></scene>"""


def _png(path, rgb8, filters):
    """An 8-bit RGB PNG with the given per-row filter types, deflated by zlib (dynamic Huffman blocks at level 9)."""
    h, w, _ = rgb8.shape
    raw = bytearray()
    prev = np.zeros((w, 3), np.int32)
    for y in range(h):
        cur = rgb8[y].astype(np.int32)
        f = filters[y % len(filters)]
        left = np.vstack([np.zeros((1, 3), np.int32), cur[:-1]])
        upleft = np.vstack([np.zeros((1, 3), np.int32), prev[:-1]])
        if f == 0: pred = np.zeros_like(cur)
        elif f == 1: pred = left
        elif f == 2: pred = prev
        elif f == 3: pred = (left + prev) // 2
        else:
            p = left + prev - upleft
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - upleft)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, upleft))
        raw.append(f)
        raw += ((cur - pred) % 256).astype(np.uint8).tobytes()
        prev = cur

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body))
    open(path, "wb").write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)) +
                           chunk(b"IDAT", zlib.compress(bytes(raw), 9)) + chunk(b"IEND", b""))


def _texels(tex):
    return np.ctypeslib.as_array(tex.texels, shape=(tex.height, tex.width, 3)).copy()


def _srgb_to_linear(v8):
    c = v8.astype(np.float64) / 255.0
    return np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)


def test_png_ppm_pfm_bitmaps_and_checkerboard_from_xml(gsb, tmp_path):
    rng = np.random.default_rng(4)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    img[:, :20] = (img[:, :20] // 64) * 64          # runs, so that the deflate stream has matches as well as literals
    _png(tmp_path / "a.png", img, [0, 1, 2, 3, 4])
    open(tmp_path / "b.ppm", "wb").write(b"P6\n# comment\n53 37\n255\n" + img.tobytes())
    hdr = rng.random((9, 7, 4)).astype(np.float32)
    gsb.write_pfm(tmp_path / "c.pfm", hdr)
    open(tmp_path / "d.jpg", "wb").write(b"\xff\xd8\xff not really a jpeg")
    shapes = ""
    for k, name in enumerate(["a.png", "b.ppm", "c.pfm", "d.jpg"]):
        shapes += f'<shape type="rectangle"><transform name="to_world"><translate x="{k}"/></transform><bsdf type="diffuse">' \
                  f'<texture name="reflectance" type="bitmap"><string name="filename" value="{name}"/><float name="uscale" value="2"/></texture></bsdf></shape>'
    shapes += '<shape type="cube"><bsdf type="twosided"><bsdf type="diffuse"><texture name="reflectance" type="checkerboard">' \
              '<rgb name="color0" value="0.9,0.1,0.1"/><rgb name="color1" value="0.1,0.1,0.9"/><float name="uscale" value="3"/>' \
              '<float name="voffset" value="0.25"/></texture></bsdf></bsdf></shape>'
    s = gsb.Scene.from_xml_string(_XML % shapes, str(tmp_path))
    tex, nt = s.textures
    assert nt == 4                                   # the jpeg is refused with a warning, its record keeps 0.5
    assert any("jpeg" in w or "unsupported" in w for w in s.warnings)
    want = _srgb_to_linear(img)
    assert np.allclose(_texels(tex[0]), want, atol=2e-6) and np.allclose(_texels(tex[1]), want, atol=2e-6)
    assert np.array_equal(_texels(tex[2]), hdr[..., :3])
    assert tex[0].type == 1 and tex[0].uscale == 2.0 and (tex[0].width, tex[0].height) == (53, 37)
    assert tex[3].type == 0 and list(tex[3].color0) == pytest.approx([0.9, 0.1, 0.1]) and tex[3].uscale == 3.0 and tex[3].voffset == 0.25
    recs = [struct.unpack("3fi", s.bsdfs(0)[0][16 * i:16 * i + 16]) for i in range(s.bsdfs(0)[1])]
    assert [r[3] for r in recs[:5]] == [1, 2, 3, 0, 4] and recs[3][:3] == (0.5, 0.5, 0.5)
    # texture coordinates of the built-in shapes: the vt lines of assets/rect.obj; the unit square on every cube face
    uv = s.mesh_uv(0)
    assert uv.shape == (6, 2) and sorted(map(tuple, np.unique(uv, axis=0))) == [(0, 0), (0, 1), (1, 0), (1, 1)]


def test_obj_texture_coordinates(gsb, tmp_path):
    open(tmp_path / "q.obj", "w").write("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvt 0 0\nvt 2 0\nvt 2 3\nvt 0 3\nvn 0 0 1\n"
                                        "f 1/1/1 2/2/1 3/3/1 4/4/1\nf 1//1 2//1 3//1\n")
    xml = _XML % '<shape type="obj"><string name="filename" value="q.obj"/><bsdf type="diffuse"><rgb name="reflectance" value="0.5,0.5,0.5"/></bsdf></shape>'
    s = gsb.Scene.from_xml_string(xml, str(tmp_path))
    uv = s.mesh_uv(0)
    assert uv.tolist() == [[0, 0], [2, 0], [2, 3], [0, 0], [2, 3], [0, 3], [0, 0], [0, 0], [0, 0]]


def _lookup(tex, texels, u, v):
    """numpy restatement of the lookup rule stated in include/gpuspectral_b200.h"""
    uu, vv = u * tex.uscale + tex.uoffset, v * tex.vscale + tex.voffset
    if tex.type == 0:
        even = (np.floor(2 * uu) + np.floor(2 * vv)) % 2 == 0
        return np.where(even[:, None], np.array(list(tex.color0))[None], np.array(list(tex.color1))[None])
    h, w, _ = texels.shape
    x, y = (uu - np.floor(uu)) * w - 0.5, (1 - (vv - np.floor(vv))) * h - 0.5
    x0, y0 = np.floor(x).astype(int), np.floor(y).astype(int)
    ax, ay = (x - x0)[:, None], (y - y0)[:, None]
    px = lambda xx, yy: texels[yy % h, xx % w]
    return (px(x0, y0) * (1 - ax) + px(x0 + 1, y0) * ax) * (1 - ay) + (px(x0, y0 + 1) * (1 - ax) + px(x0 + 1, y0 + 1) * ax) * ay


def test_oracle_texture_lookup_through_a_rendered_quad(gsb, orc):
    """The textured Cornell box: primary hits on the back wall (bitmap) and the floor (checkerboard) carry kD / pi from the
    texture at the hit's interpolated coordinate.  Checked through shallow same-seed renders: the ratio of the textured to
    the untextured frame at a pixel follows texture(uv) / reflectance (numpy restatement of the lookup rule)."""
    t, p = gsb.Scene.builtin("cornell_textured"), gsb.Scene.builtin("cornell")
    ot, op = orc.OracleScene.from_host_scene(t), orc.OracleScene.from_host_scene(p)
    for o in (ot, op):
        o.set_params(max_depth=0)          # direct light only: pixel = kD/pi x (the same light term on both sides)
    w = h = 96
    a, _, _ = ot.render(w, h, 0, 64)
    b, _, _ = op.render(w, h, 0, 64)
    tex, nt = t.textures
    assert nt == 3
    # back wall: instance 2, a rectangle; its pixels in the middle of the frame.  Same seeds -> same light samples.
    ys, xs = np.mgrid[30:60:3, 30:66:3]
    rays = op.camera_rays(w, h, xs.ravel(), ys.ravel())
    hits = op.trace_batch(rays, mode=1)
    sel = (hits["inst"] == 2) & (b[ys.ravel(), xs.ravel(), 0] > 1e-3)
    assert sel.sum() > 50
    uvm = t.mesh_uv(0).reshape(2, 3, 2)
    bary = np.stack([1 - hits["u"] - hits["v"], hits["u"], hits["v"]], 1)
    uv = np.einsum("nk,nkc->nc", bary, uvm[hits["prim"] % 2])
    kd = _lookup(tex[1], _texels(tex[1]), uv[:, 0], uv[:, 1])
    ratio = a[ys.ravel(), xs.ravel(), :3][sel] / b[ys.ravel(), xs.ravel(), :3][sel]
    want = kd[sel] / np.array([0.725, 0.71, 0.68])
    # (the path's second vertex may land on another textured surface, so the ratio is the texture value only up to the
    # indirect share of the pixel: a few per cent; a flipped or transposed lookup is off by integer factors)
    assert np.allclose(ratio, want, rtol=0.1, atol=0.03)
    assert np.corrcoef(ratio.ravel(), want.ravel())[0, 1] > 0.995


def test_baseline_jpeg_decoder_against_opencv(gsb, tmp_path):
    """Own baseline-JPEG decoder (the format of the reference's staircase2 textures) against OpenCV's libjpeg on files
    written here: 4:4:4 within a level of libjpeg's integer IDCT, 4:2:0 apart only by the chroma upsampling filter
    (nearest here, interpolating there); grey; a progressive file is refused with a warning."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(6)
    yy, xx = np.mgrid[0:75, 0:101]
    img = np.stack([127 + 120 * np.sin(xx / 9.0), 127 + 120 * np.cos(yy / 7.0), (xx * 2 + yy) % 256], 2)
    img = np.clip(img + rng.normal(0, 6, img.shape), 0, 255).astype(np.uint8)
    cases = {"a444.jpg": [cv2.IMWRITE_JPEG_QUALITY, 92, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_444],
             "a420.jpg": [cv2.IMWRITE_JPEG_QUALITY, 85, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_420, cv2.IMWRITE_JPEG_RST_INTERVAL, 3],
             "prog.jpg": [cv2.IMWRITE_JPEG_PROGRESSIVE, 1]}
    for name, flags in cases.items():
        assert cv2.imwrite(str(tmp_path / name), img[..., ::-1], flags)
    assert cv2.imwrite(str(tmp_path / "grey.jpg"), img[..., 0])
    shapes = "".join(f'<shape type="rectangle"><bsdf type="diffuse"><texture name="reflectance" type="bitmap">'
                     f'<string name="filename" value="{n}"/></texture></bsdf></shape>' for n in ("a444.jpg", "a420.jpg", "grey.jpg", "prog.jpg"))
    s = gsb.Scene.from_xml_string(_XML % shapes, str(tmp_path))
    tex, nt = s.textures
    assert nt == 3 and any("baseline" in w for w in s.warnings)

    def levels(t):      # back to 8-bit sRGB levels
        a = _texels(t).astype(np.float64)
        return np.where(a <= 0.0031308, a * 12.92, 1.055 * np.maximum(a, 1e-12) ** (1 / 2.4) - 0.055) * 255
    for t, name, mean_tol, p99_tol in ((tex[0], "a444.jpg", 0.3, 2.0), (tex[1], "a420.jpg", 5.0, 40.0)):
        ref = cv2.imread(str(tmp_path / name), cv2.IMREAD_COLOR)[..., ::-1].astype(np.float64)
        d = np.abs(levels(t) - ref)
        assert (t.width, t.height) == (101, 75)
        assert d.mean() < mean_tol and np.percentile(d, 99) <= p99_tol, (name, d.mean(), np.percentile(d, 99))
    g = cv2.imread(str(tmp_path / "grey.jpg"), cv2.IMREAD_GRAYSCALE).astype(np.float64)
    d = np.abs(levels(tex[2])[..., 1] - g)
    assert d.mean() < 0.3 and d.max() <= 2.0

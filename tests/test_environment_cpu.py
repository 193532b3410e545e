"""Environment emitter (extension; dead code in the reference, engine/Loader.cpp:338-346): the oracle's restatement pinned
by its own invariants — the sampled direction's radiance equals the lookup for that direction, the density integrates to
one and follows luminance x sin(theta), a uniform sky over a diffuse plane returns the plane's albedo (the furnace
test; exactly so without NEE, within the reference's MIS quirk with it), and the Mitsuba <emitter type="constant|envmap"> elements load."""
import numpy as np
import pytest

_XML = """<scene version="0.6.0"><sensor type="perspective"><float name="fov" value="50"/>
<transform name="to_world"><lookat origin="0,1,6" target="0,0.9,0" up="0,1,0"/></transform></sensor>
<shape type="rectangle"><transform name="to_world"><scale value="50"/><rotate x="1" angle="-90"/></transform>
<bsdf type="diffuse"><rgb name="reflectance" value="0.5,0.25,0.75"/></bsdf></shape>
%s</scene>"""


def test_uniform_sky_furnace(gsb, orc):
    s = gsb.Scene.from_xml_string(_XML % '<emitter type="constant"><rgb name="radiance" value="1,1,1"/></emitter>', ".")
    assert s.environment.type == 1 and list(s.environment.scale) == [1.0, 1.0, 1.0] and s.warnings == []
    o = orc.OracleScene.from_host_scene(s)
    for nee in (1, 0):
        o.set_params(max_depth=4, nee=nee)
        img, c, _ = o.render(24, 24, 0, 256)
        mean = img[18:24, :, :3].mean(axis=(0, 1))                               # the rows that see the plane
        # albedo x sky radiance, nothing else.  Exact without NEE; with NEE the reference's MIS weights (rayhit.rchit:751,786
        # pair the light sample's density with the density of the BSDF's OWN sampled direction, a quirk the port keeps,
        # SURVEY 8a) do not add up to one and the estimate sits a few per cent high — for triangle lights just the same
        assert np.allclose(mean, [0.5, 0.25, 0.75], rtol=0.03 if nee == 0 else 0.15), (nee, mean)
        assert np.allclose(img[:6, :, :3], 1.0)                                  # the rows that see the sky itself
        assert c["shadow_rays"] > 0 if nee else c["shadow_rays"] == 0


def test_sky_sampling_density_and_lookup_agree(gsb, orc):
    s = gsb.Scene.builtin("field_env")
    env = s.environment
    assert env.type == 2 and (env.width, env.height) == (128, 64)
    o = orc.OracleScene.from_host_scene(s)
    rng = np.random.default_rng(8)
    n = 400000
    d, rad, pdf = o.env_sample(rng.random(n), rng.random(n))
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-5) and (pdf > 0).all()
    # what is sampled is what a ray in that direction sees (same texel, same rotation convention)
    same = np.isclose(o.env_radiance(d), rad, rtol=1e-5).all(axis=1)
    assert same.mean() > 0.999                       # the rest sits exactly on a texel edge
    # the density integrates to one: E[1 / pdf] over its own samples is the measure of its support, the whole sphere
    assert np.mean(1.0 / pdf.astype(np.float64)) == pytest.approx(4 * np.pi, rel=0.02)
    # importance: the sun (a few texels out of 8192) draws a share of the samples equal to its share of luminance x sin(theta)
    tex = np.ctypeslib.as_array(env.texels, shape=(64, 128, 3))
    lum = tex @ np.array([0.2126, 0.7152, 0.0722]) * np.sin(np.pi * (np.arange(64) + 0.5) / 64)[:, None]
    sun = tex[..., 0] > 100
    assert 0 < sun.sum() < 40
    assert (rad[:, 0] > 100).mean() == pytest.approx(lum[sun].sum() / lum.sum(), rel=0.03)
    # constant sky: uniform sphere, density 1 / 4 pi
    o2 = orc.OracleScene()
    e = type(env)(1, 0, 0, None, (C3 := (2.0, 3.0, 4.0)), (1, 0, 0, 0, 1, 0, 0, 0, 1))
    o2.set_environment(e)
    d2, rad2, pdf2 = o2.env_sample(rng.random(1000), rng.random(1000))
    assert np.allclose(pdf2, 1 / (4 * np.pi)) and np.allclose(rad2, C3) and abs(d2.mean(axis=0)).max() < 0.1


def test_envmap_emitter_from_xml(gsb, tmp_path):
    sky = np.ones((8, 16, 4), np.float32)
    sky[:4, :, :3] = [0.2, 0.4, 1.0]
    gsb.write_pfm(tmp_path / "sky.pfm", sky)
    xml = _XML % ('<emitter type="envmap"><string name="filename" value="sky.pfm"/><float name="scale" value="2.5"/>'
                  '<transform name="to_world"><rotate y="1" angle="90"/></transform></emitter>')
    s = gsb.Scene.from_xml_string(xml, str(tmp_path))
    e = s.environment
    assert e.type == 2 and (e.width, e.height) == (16, 8) and list(e.scale) == [2.5, 2.5, 2.5]
    R = np.array(list(e.to_world)).reshape(3, 3)
    assert np.allclose(R @ R.T, np.eye(3), atol=1e-6) and np.allclose(R @ [0, 0, -1], [-1, 0, 0], atol=1e-6)
    assert np.allclose(np.ctypeslib.as_array(e.texels, shape=(8, 16, 3))[0, 0], [0.2, 0.4, 1.0])
    s2 = gsb.Scene.from_xml_string(_XML % '<emitter type="envmap"><string name="filename" value="missing.pfm"/></emitter>', str(tmp_path))
    assert s2.environment.type == 0 and any("envmap" in w for w in s2.warnings)

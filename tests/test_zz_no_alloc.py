"""gsb_trace_batch_static, the ray-batch entry that allocates nothing (scratch in a static device array of the library):
through the C-ABI, against gsb_trace_batch on the same rays, bit for bit (the tolerance of the intersection tests: 0),
on calls in a row that share the array.  Named test_zz_* so that pytest collects it after every other file."""
import numpy as np
import pytest

from conftest import host_scene, random_rays

pytestmark = pytest.mark.gpu
FIELDS = ("inst", "prim", "t", "u", "v")


def _same(a, b):
    for k in FIELDS:
        assert np.array_equal(a[k], b[k]), k


@pytest.mark.parametrize("name", ["cornell", "field_small"])
def test_static_batch_equals_trace_batch_on_calls_in_a_row(gsb, name):
    s = host_scene(name)
    limit = gsb.Context.STATIC_BATCH_MAX_RAYS
    with gsb.Context(0) as c:
        c.upload(s)
        first = random_rays(s, limit, seed=21)[:limit]
        second = random_rays(s, 9000, seed=22)[:7777]          # a shorter, different batch: finds the first one's rays and hits in the array
        assert len(first) == limit
        want_first, want_second = c.trace_batch(first), c.trace_batch(second)
        got_first = c.trace_batch_static(first)
        got_second = c.trace_batch_static(second)
        _same(got_first, want_first)
        _same(got_second, want_second)
        assert (got_first["inst"] != gsb.NO_HIT).mean() > 0.3
        # occlusion rays through the same array, then the first batch again
        occl = second.copy()
        occl["tmin"] = 0.01
        occl["tmax"] = np.random.default_rng(5).random(len(occl)).astype(np.float32) * 4.0
        assert np.array_equal(c.trace_batch_static(occl, any_hit=True)["inst"], c.trace_batch(occl, any_hit=True)["inst"])
        _same(c.trace_batch_static(first), want_first)
        assert len(c.trace_batch_static(np.zeros(0, gsb.RAY_DTYPE))) == 0


def test_static_batch_between_renders_and_across_contexts(gsb):
    """The array belongs to the device, not to a context: a second context's call, and a render in between, leave the results equal."""
    s = host_scene("cornell")
    rays = random_rays(s, 30000, seed=23)
    with gsb.Context(0) as a, gsb.Context(0) as b:
        a.upload(s)
        b.upload(host_scene("gallery"))
        want = a.trace_batch(rays)
        _same(a.trace_batch_static(rays), want)
        other = b.trace_batch_static(rays[:5000])
        _same(other, b.trace_batch(rays[:5000]))
        a.set_params(max_depth=4)
        a.render(64, 48, 0, 2)
        try:                                                     # an asynchronous render still in flight is refused, not waited for
            _same(a.trace_batch_static(rays), want)
        except gsb.GsbError as e:
            assert "pending" in str(e)
        a.synchronize()
        _same(a.trace_batch_static(rays), want)


def test_static_batch_rejects_more_rays_than_the_array_holds(gsb):
    s = host_scene("cornell")
    with gsb.Context(0) as c:
        c.upload(s)
        rays = np.zeros(gsb.Context.STATIC_BATCH_MAX_RAYS + 1, gsb.RAY_DTYPE)
        with pytest.raises(Exception, match="GSB_STATIC_BATCH_MAX_RAYS"):
            c.trace_batch_static(rays)
        ok = random_rays(s, 4000, seed=24)                       # the refusal left the entry usable
        _same(c.trace_batch_static(ok), c.trace_batch(ok))

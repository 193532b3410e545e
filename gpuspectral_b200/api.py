"""Python view of the host layer and the CUDA core (thin: every method is one native call).

Scene       the host scene model (renderer/Scene.h of the reference) loaded from Mitsuba XML
            or generated; exposes the flat arrays the C ABI consumes
Context     one gsb_context = one GPU: the calls of include/gpuspectral_b200.h
PathTracer  the render pass object shaped like the reference's PathTracer plugin
            (renderer/PathTracer.h:48-68): setup(), create_render_pass(scene), timestamp
"""
import ctypes as C

import numpy as np

from . import _native as N


class GsbError(RuntimeError):
    pass


def _check(status, what):
    if status != 0:
        raise GsbError(f"{what}: {N.core.gsb_last_error().decode(errors='replace')} (status {status})")


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


RAY_DTYPE = np.dtype([("origin", np.float32, 3), ("tmin", np.float32), ("dir", np.float32, 3), ("tmax", np.float32)])
HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("inst", np.uint32), ("prim", np.uint32)])


def make_rays(origins, dirs, tmin=0.0, tmax=1e10):
    origins = np.asarray(origins, np.float32)
    dirs = np.asarray(dirs, np.float32)
    rays = np.zeros(len(origins), RAY_DTYPE)
    rays["origin"] = origins
    rays["dir"] = dirs
    rays["tmin"] = tmin
    rays["tmax"] = tmax
    return rays


class Scene:
    """Host scene. Owns a gsbh_scene."""

    def __init__(self, handle):
        if not handle:
            raise GsbError("scene: " + N.host.gsbh_last_error().decode(errors="replace"))
        self._h = C.c_void_p(handle)

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h and N is not None and getattr(N, "host", None) is not None:
            N.host.gsbh_scene_free(h)

    @staticmethod
    def load_xml(path):
        return Scene(N.host.gsbh_scene_load_xml(str(path).encode()))

    @staticmethod
    def load_gltf(path):
        return Scene(N.host.gsbh_scene_load_gltf(str(path).encode()))

    @staticmethod
    def load(path):
        """Mitsuba XML or glTF (.gltf / .glb), by extension."""
        return Scene.load_gltf(path) if str(path).lower().endswith((".gltf", ".glb")) else Scene.load_xml(path)

    @staticmethod
    def from_xml_string(xml, base_dir="."):
        return Scene(N.host.gsbh_scene_load_xml_string(xml.encode(), str(base_dir).encode()))

    @staticmethod
    def builtin(name, *params):
        arr = (C.c_int * max(1, len(params)))(*params)
        return Scene(N.host.gsbh_scene_builtin(name.encode(), arr, len(params)))

    # ---- flat arrays ------------------------------------------------------------------
    @property
    def mesh_count(self):
        return N.host.gsbh_scene_mesh_count(self._h)

    def mesh(self, i):
        """(positions[n,3], normals[n,3]) float32 copies of mesh i (non-indexed, 3 vertices per triangle)."""
        n = N.host.gsbh_scene_mesh_vertex_count(self._h, i)
        if n == 0:
            return np.zeros((0, 3), np.float32), np.zeros((0, 3), np.float32)
        p = np.ctypeslib.as_array(N.host.gsbh_scene_mesh_positions(self._h, i), shape=(n, 3)).copy()
        q = np.ctypeslib.as_array(N.host.gsbh_scene_mesh_normals(self._h, i), shape=(n, 3)).copy()
        return p, q

    def mesh_uv(self, i):
        """(n, 2) float32 texture coordinates of mesh i, or None when the mesh has none."""
        n = N.host.gsbh_scene_mesh_vertex_count(self._h, i)
        p = N.host.gsbh_scene_mesh_uvs(self._h, i)
        if n == 0 or not p:
            return None
        return np.ctypeslib.as_array(p, shape=(n, 2)).copy()

    @property
    def textures(self):
        """(gsb_texture array, count); bitmap texel pointers belong to this Scene."""
        n = N.host.gsbh_scene_texture_count(self._h)
        arr = (N.Texture * max(1, n))()
        if n:
            N.host.gsbh_scene_copy_textures(self._h, arr)
        return arr, n

    @property
    def environment(self):
        """gsb_environment record (type 0 = none); the texel pointer belongs to this Scene."""
        e = N.Environment()
        N.host.gsbh_scene_environment(self._h, C.byref(e))
        return e

    def mesh_name(self, i):
        return N.host.gsbh_scene_mesh_name(self._h, i).decode()

    @property
    def instances(self):
        n = N.host.gsbh_scene_instance_count(self._h)
        arr = (N.Instance * max(1, n))()
        if n:
            N.host.gsbh_scene_copy_instances(self._h, arr)
        return arr, n

    def bsdfs(self, btype):
        """(raw bytes of the parameter records, count) for one BSDF type."""
        n = N.host.gsbh_scene_bsdf_count(self._h, btype)
        if n == 0:
            return b"", 0
        ptr = N.host.gsbh_scene_bsdf_data(self._h, btype)
        return C.string_at(ptr, n * N.BSDF_STRIDES[btype]), n

    @property
    def lights(self):
        n = N.host.gsbh_scene_light_count(self._h)
        arr = (N.TriangleLight * max(1, n))()
        if n:
            C.memmove(arr, N.host.gsbh_scene_lights(self._h), n * C.sizeof(N.TriangleLight))
        return arr, n

    @property
    def camera(self):
        eye = np.zeros(3, np.float32)
        tw = np.zeros(16, np.float32)
        fov = C.c_float()
        N.host.gsbh_scene_camera(self._h, _fp(eye), _fp(tw), C.byref(fov))
        return eye, tw, fov.value

    @property
    def film(self):
        w, h = C.c_uint32(), C.c_uint32()
        N.host.gsbh_scene_film(self._h, C.byref(w), C.byref(h))
        return w.value, h.value

    @property
    def instanced_triangles(self):
        return N.host.gsbh_scene_instanced_triangles(self._h)

    @property
    def warnings(self):
        return [N.host.gsbh_scene_warning(self._h, i).decode() for i in range(N.host.gsbh_scene_warning_count(self._h))]


class Context:
    """One GPU context: the C ABI, call for call."""

    def __init__(self, device=0, _borrowed=None):
        self._own = _borrowed is None
        if _borrowed is not None:
            self._c = C.c_void_p(_borrowed)
        else:
            h = C.c_void_p()
            _check(N.core.gsb_create(device, C.byref(h)), "gsb_create")
            self._c = h
        self._mesh_map = []
        self._size = None

    def close(self):
        c, self._c = getattr(self, "_c", None), None
        if c and self._own and N is not None and getattr(N, "core", None) is not None:
            N.core.gsb_destroy(c)

    def __del__(self):
        self.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- scene ----------------------------------------------------------------------------
    def mesh_create(self, positions, normals, uv=None):
        p = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        q = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        out = C.c_uint32()
        if uv is None:
            _check(N.core.gsb_mesh_create(self._c, _fp(p), _fp(q), len(p), C.byref(out)), "gsb_mesh_create")
        else:
            t = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
            _check(N.core.gsb_mesh_create_uv(self._c, _fp(p), _fp(q), _fp(t), len(p), C.byref(out)), "gsb_mesh_create_uv")
        return out.value

    def set_environment(self, env=None):
        """env: _native.Environment record or None (no environment emitter)."""
        _check(N.core.gsb_scene_set_environment(self._c, C.byref(env) if env is not None else None), "gsb_scene_set_environment")

    def set_textures(self, textures, count):
        _check(N.core.gsb_scene_set_textures(self._c, textures, count), "gsb_scene_set_textures")

    def mesh_destroy(self, mesh):
        _check(N.core.gsb_mesh_destroy(self._c, mesh), "gsb_mesh_destroy")

    def set_instances(self, instances, count):
        _check(N.core.gsb_scene_set_instances(self._c, instances, count), "gsb_scene_set_instances")

    def update_transforms(self, first, transforms):
        """transforms: (n, 16) column-major 4x4 matrices for instances first .. first+n-1 (moved instances: refit, no re-upload)."""
        t = np.ascontiguousarray(transforms, np.float32).reshape(-1, 16)
        _check(N.core.gsb_scene_update_transforms(self._c, first, len(t), _fp(t)), "gsb_scene_update_transforms")

    def set_bsdfs(self, btype, raw, count):
        buf = C.create_string_buffer(raw, max(1, len(raw)))
        _check(N.core.gsb_scene_set_bsdfs(self._c, btype, C.cast(buf, C.c_void_p), count), "gsb_scene_set_bsdfs")

    def set_lights(self, lights, count):
        _check(N.core.gsb_scene_set_lights(self._c, lights, count), "gsb_scene_set_lights")

    def set_camera(self, eye, to_world, fov):
        eye = np.ascontiguousarray(eye, np.float32)
        tw = np.ascontiguousarray(to_world, np.float32).reshape(16)
        _check(N.core.gsb_set_camera(self._c, _fp(eye), _fp(tw), fov), "gsb_set_camera")

    @staticmethod
    def default_params():
        p = N.Params()
        N.core.gsb_default_params(C.byref(p))
        return p

    def set_params(self, max_depth=50, rr_start=10, cutoff=20.0, nee=1, pixel_filter=0):
        """pixel_filter: 0 none (the reference), 1 box jitter, 2 tent of radius 1 pixel."""
        p = N.Params(max_depth, rr_start, cutoff, nee, pixel_filter)
        _check(N.core.gsb_set_params(self._c, C.byref(p)), "gsb_set_params")

    def upload(self, scene):
        """Uploads a whole host Scene through the scene calls of the C ABI (meshes, BSDFs, lights, instances, camera)."""
        for m in self._mesh_map:
            N.core.gsb_mesh_destroy(self._c, m)
        has_uv = hasattr(scene, "mesh_uv")
        self._mesh_map = [self.mesh_create(*scene.mesh(i), uv=scene.mesh_uv(i) if has_uv else None) for i in range(scene.mesh_count)]
        if hasattr(scene, "textures"):
            tex, nt = scene.textures
            self.set_textures(tex, nt)
        self.set_environment(scene.environment if hasattr(scene, "environment") else None)
        for t in range(N.BSDF_TYPE_COUNT):
            raw, n = scene.bsdfs(t)
            self.set_bsdfs(t, raw, n)
        lights, nl = scene.lights
        self.set_lights(lights, nl)
        inst, ni = scene.instances
        for i in range(ni):
            inst[i].mesh = self._mesh_map[inst[i].mesh]
        self.set_instances(inst, ni)
        self.set_camera(*scene.camera)

    def bvh_stats(self):
        s = N.BvhStats()
        _check(N.core.gsb_get_bvh_stats(self._c, C.byref(s)), "gsb_get_bvh_stats")
        return {n: getattr(s, n) for n, _ in N.BvhStats._fields_}

    # ---- render ---------------------------------------------------------------------------
    def render(self, width, height, first_sample, num_samples, sample_stride=1):
        _check(N.core.gsb_render(self._c, width, height, first_sample, num_samples, sample_stride), "gsb_render")
        self._size = (width, height)

    def clear(self):
        _check(N.core.gsb_clear(self._c), "gsb_clear")

    def _film(self, what):
        if self._size is None:
            raise GsbError(f"{what}: nothing rendered yet")
        return self._size

    def synchronize(self):
        _check(N.core.gsb_synchronize(self._c), "gsb_synchronize")

    def read_hdr(self, out=None):
        w, h = self._film("gsb_read_hdr")
        img = out if out is not None else np.empty((h, w, 4), np.float32)
        _check(N.core.gsb_read_hdr(self._c, _fp(img)), "gsb_read_hdr")
        return img

    def read_ldr(self):
        w, h = self._film("gsb_read_ldr")
        img = np.empty((h, w, 4), np.uint8)
        _check(N.core.gsb_read_ldr(self._c, img.ctypes.data_as(C.POINTER(C.c_uint8))), "gsb_read_ldr")
        return img

    def read_sum(self):
        """Checkpoint: the raw accumulation buffer (H, W, 4) = rgb sums + sample counts."""
        w, h = self._film("gsb_read_sum")
        out = np.empty((h, w, 4), np.float32)
        _check(N.core.gsb_read_sum(self._c, _fp(out)), "gsb_read_sum")
        return out

    def write_sum(self, sums):
        """Resume: replaces the accumulation buffer by a checkpoint taken with read_sum (any context, any GPU)."""
        sums = np.ascontiguousarray(sums, np.float32)
        h, w = sums.shape[:2]
        _check(N.core.gsb_write_sum(self._c, w, h, _fp(sums)), "gsb_write_sum")
        self._size = (w, h)

    def device_sum_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        _check(N.core.gsb_device_sum_ptr(self._c, C.byref(p), C.byref(n)), "gsb_device_sum_ptr")
        return p.value, n.value

    def counters(self):
        c = N.Counters()
        _check(N.core.gsb_get_counters(self._c, C.byref(c)), "gsb_get_counters")
        return {n: getattr(c, n) for n, _ in N.Counters._fields_}

    def reset_counters(self):
        _check(N.core.gsb_reset_counters(self._c), "gsb_reset_counters")

    def set_profiling(self, enabled):
        _check(N.core.gsb_set_profiling(self._c, int(enabled)), "gsb_set_profiling")

    def set_instrumented(self, mode):
        """0 off, 1 extension rays, 2 shadow rays, 3 both."""
        _check(N.core.gsb_set_instrumented(self._c, int(mode)), "gsb_set_instrumented")

    def set_stream(self, cuda_stream):
        """cuda_stream: integer value of a cudaStream_t (e.g. torch.cuda.current_stream().cuda_stream); None = private stream."""
        _check(N.core.gsb_set_stream(self._c, C.c_void_p(cuda_stream or 0)), "gsb_set_stream")

    def launch_info(self):
        a, b = C.c_uint64(), C.c_uint64()
        _check(N.core.gsb_get_launch_info(self._c, C.byref(a), C.byref(b)), "gsb_get_launch_info")
        return a.value, b.value

    def last_render_ms(self):
        ms = (C.c_double * 5)()
        _check(N.core.gsb_last_render_ms(self._c, ms), "gsb_last_render_ms")
        return dict(zip(("total", "extend", "shade", "shadow", "finish"), ms))

    # ---- ray batches ----------------------------------------------------------------------
    def trace_batch(self, rays, any_hit=False, stats=False):
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.zeros(len(rays), HIT_DTYPE)
        fn = N.core.gsb_trace_batch_stats if stats else N.core.gsb_trace_batch
        _check(fn(self._c, rays.ctypes.data_as(C.POINTER(N.Ray)), len(rays), hits.ctypes.data_as(C.POINTER(N.Hit)), int(any_hit)),
               "gsb_trace_batch")
        return hits

    STATIC_BATCH_MAX_RAYS = 65536

    def trace_batch_static(self, rays, any_hit=False):
        """gsb_trace_batch_static: the same hits as trace_batch with no device allocation (scratch in a static
        device array of the library, at most STATIC_BATCH_MAX_RAYS rays per call; calls on one device must not overlap)."""
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.zeros(len(rays), HIT_DTYPE)
        _check(N.core.gsb_trace_batch_static(self._c, rays.ctypes.data_as(C.POINTER(N.Ray)), len(rays),
                                             hits.ctypes.data_as(C.POINTER(N.Hit)), int(any_hit)), "gsb_trace_batch_static")
        return hits

    def trace_batch_timed(self, rays, any_hit=False, repeats=10):
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        ms = C.c_double()
        _check(N.core.gsb_trace_batch_timed(self._c, rays.ctypes.data_as(C.POINTER(N.Ray)), len(rays), int(any_hit), repeats, C.byref(ms)),
               "gsb_trace_batch_timed")
        return ms.value

    # ---- diagnostics ----------------------------------------------------------------------
    def test_rng(self, idx, timestamp, n_draws):
        idx = np.ascontiguousarray(idx, np.uint32)
        ts = np.ascontiguousarray(timestamp, np.uint32)
        out = np.zeros((len(idx), 1 + n_draws), np.uint32)
        _check(N.core.gsb_test_rng(self._c, _up(idx), _up(ts), len(idx), n_draws, _up(out)), "gsb_test_rng")
        return out

    def test_camera_rays(self, width, height, px, py):
        px = np.ascontiguousarray(px, np.uint32)
        py = np.ascontiguousarray(py, np.uint32)
        out = np.zeros(len(px), RAY_DTYPE)
        _check(N.core.gsb_test_camera_rays(self._c, width, height, _up(px), _up(py), len(px), out.ctypes.data_as(C.POINTER(N.Ray))),
               "gsb_test_camera_rays")
        return out

    def test_bsdf_sample(self, handle, wo, seed):
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        seed = np.ascontiguousarray(seed, np.uint32)
        n = len(wo)
        wi, f = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        pdf, delta, seed_out = np.zeros(n, np.float32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        _check(N.core.gsb_test_bsdf_sample(self._c, handle, _fp(wo), _up(seed), n, _fp(wi), _fp(f), _fp(pdf), _up(delta), _up(seed_out)),
               "gsb_test_bsdf_sample")
        return wi, f, pdf, delta, seed_out

    def test_bsdf_eval(self, handle, wo, wi):
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, np.float32).reshape(-1, 3)
        n = len(wo)
        f, pdf, delta = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint32)
        _check(N.core.gsb_test_bsdf_eval(self._c, handle, _fp(wo), _fp(wi), n, _fp(f), _fp(pdf), _up(delta)), "gsb_test_bsdf_eval")
        return f, pdf, delta

    def test_sample_light(self, pos, seed):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        seed = np.ascontiguousarray(seed, np.uint32)
        n = len(pos)
        lpos, em = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        pdf, seed_out = np.zeros(n, np.float32), np.zeros(n, np.uint32)
        _check(N.core.gsb_test_sample_light(self._c, _fp(pos), _up(seed), n, _fp(lpos), _fp(em), _fp(pdf), _up(seed_out)),
               "gsb_test_sample_light")
        return lpos, em, pdf, seed_out


class PathTracer:
    """The reference's render-pass plugin shape over the CUDA core (host/path_tracer.h)."""

    def __init__(self, device=0):
        h = N.host.gsbh_tracer_create(device)
        if not h:
            raise GsbError("PathTracer: " + N.host.gsbh_last_error().decode(errors="replace"))
        self._t = C.c_void_p(h)
        self.context = Context(_borrowed=N.host.gsbh_tracer_context(self._t))
        self.width = self.height = 0

    def close(self):
        t, self._t = getattr(self, "_t", None), None
        if t and N is not None and getattr(N, "host", None) is not None:
            N.host.gsbh_tracer_destroy(t)

    def __del__(self):
        self.close()

    def _hcheck(self, rc, what):
        if rc != 0:
            raise GsbError(f"{what}: {N.host.gsbh_last_error().decode(errors='replace')}")

    def setup(self, width, height):
        self._hcheck(N.host.gsbh_tracer_setup(self._t, width, height), "setup")
        self.width, self.height = width, height
        self.context._size = (width, height)

    def set_params(self, max_depth=50, rr_start=10, cutoff=20.0, nee=1, pixel_filter=0):
        """pixel_filter: 0 none (the reference), 1 box jitter, 2 tent of radius 1 pixel."""
        p = N.Params(max_depth, rr_start, cutoff, nee, pixel_filter)
        self._hcheck(N.host.gsbh_tracer_set_params(self._t, C.byref(p)), "set_params")

    def set_timestamp(self, ts):
        N.host.gsbh_tracer_set_timestamp(self._t, ts)

    def upload(self, scene):
        self._hcheck(N.host.gsbh_tracer_upload(self._t, scene._h), "upload")

    def create_render_pass(self, scene):
        """One frame of the reference: 1 sample per pixel, timestamp++ (PathTracer.cpp:9-93)."""
        self._hcheck(N.host.gsbh_tracer_create_render_pass(self._t, scene._h), "create_render_pass")

    def render(self, scene, spp, stride=1):
        self._hcheck(N.host.gsbh_tracer_render(self._t, scene._h, spp, stride), "render")

    def read_hdr(self, out=None):
        img = out if out is not None else np.empty((self.height, self.width, 4), np.float32)
        self._hcheck(N.host.gsbh_tracer_read_hdr(self._t, _fp(img)), "read_hdr")
        return img

    def read_ldr(self):
        img = np.empty((self.height, self.width, 4), np.uint8)
        self._hcheck(N.host.gsbh_tracer_read_ldr(self._t, img.ctypes.data_as(C.POINTER(C.c_uint8))), "read_ldr")
        return img

    def synchronize(self):
        self._hcheck(N.host.gsbh_tracer_synchronize(self._t), "synchronize")


def read_pfm(path):
    """PFM file -> (H, W, 4) float32, rows top-to-bottom (the host layer's reader; what `gsb_render --diff` uses)."""
    w, h = C.c_uint32(), C.c_uint32()
    if N.host.gsbh_read_pfm(str(path).encode(), None, 0, C.byref(w), C.byref(h)) != 0:
        raise GsbError(N.host.gsbh_last_error().decode())
    img = np.empty((h.value, w.value, 4), np.float32)
    if N.host.gsbh_read_pfm(str(path).encode(), _fp(img), w.value * h.value, C.byref(w), C.byref(h)) != 0:
        raise GsbError(N.host.gsbh_last_error().decode())
    return img


def write_pfm(path, rgba):
    rgba = np.ascontiguousarray(rgba, np.float32)
    if N.host.gsbh_write_pfm(str(path).encode(), _fp(rgba), rgba.shape[1], rgba.shape[0]) != 0:
        raise GsbError(N.host.gsbh_last_error().decode())


def image_diff(a, b):
    """relMSE / RMSE / means / max |a - b| of two (H, W, 4) images over rgb; b is the reference (north_star check 3)."""
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    if a.shape != b.shape:
        raise ValueError("image sizes differ")
    out = (C.c_double * 5)()
    if N.host.gsbh_image_diff(_fp(a), _fp(b), a.shape[1], a.shape[0], out) != 0:
        raise GsbError(N.host.gsbh_last_error().decode())
    return dict(zip(("rel_mse", "rmse", "mean_a", "mean_b", "max_abs"), out))

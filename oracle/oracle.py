"""ctypes binding of the CPU oracle (oracle/liboracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package (gpuspectral_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")

RAY_DTYPE = np.dtype([("origin", np.float32, 3), ("tmin", np.float32), ("dir", np.float32, 3), ("tmax", np.float32)])
HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("inst", np.uint32), ("prim", np.uint32)])
BOUNCE_DTYPE = np.dtype([("inst", np.uint32), ("prim", np.uint32), ("t", np.float32), ("emitted", np.float32, 3),
                         ("weight", np.float32, 3), ("seed", np.uint32), ("shadow_cast", np.uint32), ("shadow_visible", np.uint32)])
NO_HIT = 0xFFFFFFFF


class Params(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("rr_start", C.c_int32), ("cutoff", C.c_float), ("nee", C.c_int32), ("pixel_filter", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("paths", "extension_rays", "shadow_rays", "rng_draws")]


def build(force=False):
    """Compiles liboracle.so with oracle/Makefile when it is missing or older than its sources."""
    srcs = [os.path.join(_HERE, f) for f in ("oracle.cpp", "oracle.h", "Makefile")]
    srcs.append(os.path.join(_HERE, "..", "include", "gpuspectral_b200.h"))
    stale = force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if stale:
        r = subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], capture_output=True, text=True)
        if r.returncode != 0:
            if force or not os.path.exists(_SO):
                raise RuntimeError("oracle: make failed\n" + r.stdout[-2000:] + r.stderr[-2000:])
            # a prebuilt library travels with the repository snapshot; file times on the receiving box say nothing about
            # staleness, and a box without a host compiler must still be able to run the checker
            import warnings
            warnings.warn("oracle: rebuild failed, using the existing liboracle.so")
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        P, u32p, f32p = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_float)
        sig = {
            "or_scene_create": (P, []),
            "or_scene_destroy": (None, [P]),
            "or_scene_add_mesh": (C.c_uint32, [P, f32p, f32p, C.c_uint32]),
            "or_scene_add_mesh_uv": (C.c_uint32, [P, f32p, f32p, f32p, C.c_uint32]),
            "or_scene_set_textures": (C.c_int, [P, P, C.c_uint32]),
            "or_scene_set_environment": (C.c_int, [P, P]),
            "or_env_sample": (None, [P, f32p, f32p, C.c_uint32, f32p, f32p, f32p]),
            "or_env_radiance": (None, [P, f32p, C.c_uint32, f32p]),
            "or_scene_set_instances": (C.c_int, [P, P, C.c_uint32]),
            "or_scene_set_bsdfs": (C.c_int, [P, C.c_uint32, P, C.c_uint32]),
            "or_scene_set_lights": (C.c_int, [P, P, C.c_uint32]),
            "or_scene_set_camera": (None, [P, f32p, f32p, C.c_float]),
            "or_scene_set_params": (None, [P, C.POINTER(Params)]),
            "or_render": (C.c_int, [P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, f32p, C.c_int, C.c_int, C.POINTER(Counters)]),
            "or_path_probe": (C.c_int, [P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, f32p, P, C.c_uint32]),
            "or_trace_batch": (C.c_int, [P, P, C.c_uint32, P, C.c_int, C.c_int, C.c_int]),
            "or_intersect": (C.c_int, [P, f32p, f32p, C.c_float, C.c_float, C.c_int, P]),
            "or_tea": (C.c_uint32, [C.c_uint32, C.c_uint32]),
            "or_pcg_hash": (C.c_uint32, [C.c_uint32]),
            "or_rng": (None, [u32p, u32p, C.c_uint32, C.c_uint32, u32p]),
            "or_rand_to_float": (C.c_float, [C.c_uint32]),
            "or_camera_rays": (None, [P, C.c_uint32, C.c_uint32, u32p, u32p, C.c_uint32, P]),
            "or_bsdf_sample": (C.c_int, [P, C.c_uint32, f32p, u32p, C.c_uint32, f32p, f32p, f32p, u32p, u32p]),
            "or_bsdf_eval": (C.c_int, [P, C.c_uint32, f32p, f32p, C.c_uint32, f32p, f32p, u32p]),
            "or_sample_light": (C.c_int, [P, f32p, u32p, C.c_uint32, f32p, f32p, f32p, u32p]),
            "or_encode_srgb8": (None, [f32p, C.c_uint32, C.POINTER(C.c_uint8)]),
        }
        for name, (res, args) in sig.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


def _up(a):
    return a.ctypes.data_as(C.POINTER(C.c_uint32))


class OracleScene:
    """The oracle's scene, fed with the same flat records as the product's C ABI."""

    def __init__(self):
        self._s = C.c_void_p(lib().or_scene_create())
        self.set_params()

    def __del__(self):
        s, self._s = getattr(self, "_s", None), None
        if s and _lib is not None:
            _lib.or_scene_destroy(s)

    @staticmethod
    def from_host_scene(scene):
        """scene: gpuspectral_b200.Scene (the host layer's flat arrays; no product compute involved)."""
        o = OracleScene()
        has_uv = hasattr(scene, "mesh_uv")
        for i in range(scene.mesh_count):
            o.add_mesh(*scene.mesh(i), uv=scene.mesh_uv(i) if has_uv else None)
        if hasattr(scene, "textures"):
            tex, nt = scene.textures
            o.set_textures(tex, nt)
        if hasattr(scene, "environment"):
            o.set_environment(scene.environment)
        for t in range(9):   # 8 reference types + the rough dielectric extension
            raw, n = scene.bsdfs(t)
            o.set_bsdfs(t, raw, n)
        lights, nl = scene.lights
        o.set_lights(lights, nl)
        inst, ni = scene.instances
        o.set_instances(inst, ni)
        o.set_camera(*scene.camera)
        return o

    # ---- scene archives: the flat records of a host scene as one .npz, so that the CPU arm of bench.py can run without the
    # product package (tests/golden/make_scene_npz.py writes them; tests/test_oracle_cpu.py checks archive == host scene)
    @staticmethod
    def dump_host_scene(scene, path, max_texture=0):
        """max_texture > 0: bitmap textures are box-filtered down until neither side exceeds it (keeps archives small)."""
        import ctypes
        arrays = {"mesh_count": np.array([scene.mesh_count], np.uint32)}
        tex, nt = scene.textures if hasattr(scene, "textures") else (None, 0)
        arrays["texture_count"] = np.array([nt], np.uint32)
        for t in range(nt):
            d = tex[t]
            arrays[f"tex{t}_desc"] = np.array([d.type, d.uscale, d.vscale, d.uoffset, d.voffset, *d.color0, *d.color1], np.float64)
            if d.type == 1:
                px = np.ctypeslib.as_array(d.texels, shape=(d.height, d.width, 3)).astype(np.float32)
                f = 1
                while max_texture and max(d.width, d.height) // f > max_texture:
                    f *= 2
                if f > 1:
                    h, w = d.height // f * f, d.width // f * f
                    px = px[:h, :w].reshape(h // f, f, w // f, f, 3).mean(axis=(1, 3)).astype(np.float32)
                arrays[f"tex{t}_px"] = px.astype(np.float16)
        for i in range(scene.mesh_count):
            p, q = scene.mesh(i)
            arrays[f"mesh{i}_p"], arrays[f"mesh{i}_n"] = p, q
            uv = scene.mesh_uv(i) if (nt and hasattr(scene, "mesh_uv")) else None
            if uv is not None:
                arrays[f"mesh{i}_uv"] = uv
        for t in range(9):
            raw, n = scene.bsdfs(t)
            arrays[f"bsdf{t}"] = np.frombuffer(raw, np.uint8)
            arrays[f"bsdf{t}_count"] = np.array([n], np.uint32)
        lights, nl = scene.lights
        arrays["lights"] = np.frombuffer(ctypes.string_at(lights, ctypes.sizeof(lights)), np.uint8)
        arrays["lights_count"] = np.array([nl], np.uint32)
        inst, ni = scene.instances
        arrays["instances"] = np.frombuffer(ctypes.string_at(inst, ctypes.sizeof(inst)), np.uint8)
        arrays["instances_count"] = np.array([ni], np.uint32)
        eye, tw, fov = scene.camera
        arrays["eye"], arrays["to_world"], arrays["fov"] = eye, tw, np.array([fov], np.float32)
        np.savez_compressed(path, **arrays)

    @staticmethod
    def from_npz(path):
        z = np.load(path)
        o = OracleScene()
        for i in range(int(z["mesh_count"][0])):
            o.add_mesh(z[f"mesh{i}_p"], z[f"mesh{i}_n"], uv=z[f"mesh{i}_uv"] if f"mesh{i}_uv" in z else None)
        tex, nt, _keep = archive_textures(z)
        if nt:
            o.set_textures(tex, nt)
        for t in range(9):
            o.set_bsdfs(t, z[f"bsdf{t}"].tobytes(), int(z[f"bsdf{t}_count"][0]))
        lights = C.create_string_buffer(z["lights"].tobytes(), max(1, z["lights"].size))
        o.set_lights(lights, int(z["lights_count"][0]))
        inst = C.create_string_buffer(z["instances"].tobytes(), max(1, z["instances"].size))
        o.set_instances(inst, int(z["instances_count"][0]))
        o.set_camera(z["eye"], z["to_world"], float(z["fov"][0]))
        return o

    def add_mesh(self, positions, normals, uv=None):
        p = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        q = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        if uv is None:
            return lib().or_scene_add_mesh(self._s, _fp(p), _fp(q), len(p))
        t = np.ascontiguousarray(uv, np.float32).reshape(-1, 2)
        return lib().or_scene_add_mesh_uv(self._s, _fp(p), _fp(q), _fp(t), len(p))

    def set_environment(self, env=None):
        """env: ctypes gsb_environment record (gpuspectral_b200._native.Environment layout) or None."""
        if lib().or_scene_set_environment(self._s, C.cast(C.pointer(env), C.c_void_p) if env is not None else None) != 0:
            raise ValueError("oracle: the environment image has no texels")

    def env_sample(self, e1, e2):
        e1, e2 = np.ascontiguousarray(e1, np.float32), np.ascontiguousarray(e2, np.float32)
        n = len(e1)
        d, r, p = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.float32)
        lib().or_env_sample(self._s, _fp(e1), _fp(e2), n, _fp(d), _fp(r), _fp(p))
        return d, r, p

    def env_radiance(self, dirs):
        dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        out = np.zeros((len(dirs), 3), np.float32)
        lib().or_env_radiance(self._s, _fp(dirs), len(dirs), _fp(out))
        return out

    def set_textures(self, textures, count):
        """textures: ctypes array of gsb_texture records (gpuspectral_b200._native.Texture layout)."""
        if lib().or_scene_set_textures(self._s, C.cast(textures, C.c_void_p) if textures is not None else None, count) != 0:
            raise ValueError("oracle: a bitmap texture has no texels")

    def set_instances(self, instances, count):
        if lib().or_scene_set_instances(self._s, C.cast(instances, C.c_void_p), count) != 0:
            raise ValueError("oracle: instance references an unknown mesh")

    def set_bsdfs(self, btype, raw, count):
        buf = C.create_string_buffer(raw, max(1, len(raw)))
        lib().or_scene_set_bsdfs(self._s, btype, C.cast(buf, C.c_void_p), count)

    def set_lights(self, lights, count):
        lib().or_scene_set_lights(self._s, C.cast(lights, C.c_void_p), count)

    def set_camera(self, eye, to_world, fov):
        eye = np.ascontiguousarray(eye, np.float32)
        tw = np.ascontiguousarray(to_world, np.float32).reshape(16)
        lib().or_scene_set_camera(self._s, _fp(eye), _fp(tw), fov)

    def set_params(self, max_depth=50, rr_start=10, cutoff=20.0, nee=1, pixel_filter=0):
        p = Params(max_depth, rr_start, cutoff, nee, pixel_filter)
        lib().or_scene_set_params(self._s, C.byref(p))

    def render(self, width, height, first=0, n=1, stride=1, image=None, sum_mode=False, threads=0):
        """Returns (image[h,w,4] float32, counters dict, threads used). Running mean unless sum_mode."""
        img = image if image is not None else np.zeros((height, width, 4), np.float32)
        c = Counters()
        used = lib().or_render(self._s, width, height, first, n, stride, _fp(img), int(sum_mode), threads, C.byref(c))
        if used < 0:
            raise ValueError("oracle: an instance references a BSDF record that was not set")
        return img, {k: getattr(c, k) for k, _ in Counters._fields_}, used

    def path_probe(self, width, height, x, y, timestamp, max_bounces=64):
        rgb = np.zeros(3, np.float32)
        b = np.zeros(max_bounces, BOUNCE_DTYPE)
        n = lib().or_path_probe(self._s, width, height, x, y, timestamp, _fp(rgb), b.ctypes.data_as(C.c_void_p), max_bounces)
        return rgb, b[:min(n, max_bounces)]

    def trace_batch(self, rays, any_hit=False, mode=0, threads=0):
        rays = np.ascontiguousarray(rays, RAY_DTYPE)
        hits = np.zeros(len(rays), HIT_DTYPE)
        lib().or_trace_batch(self._s, rays.ctypes.data_as(C.c_void_p), len(rays), hits.ctypes.data_as(C.c_void_p), int(any_hit), mode, threads)
        return hits

    def camera_rays(self, width, height, px, py):
        px = np.ascontiguousarray(px, np.uint32)
        py = np.ascontiguousarray(py, np.uint32)
        out = np.zeros(len(px), RAY_DTYPE)
        lib().or_camera_rays(self._s, width, height, _up(px), _up(py), len(px), out.ctypes.data_as(C.c_void_p))
        return out

    def bsdf_sample(self, handle, wo, seed):
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        seed = np.ascontiguousarray(seed, np.uint32)
        n = len(wo)
        wi, f = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        pdf, delta, seed_out = np.zeros(n, np.float32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        if lib().or_bsdf_sample(self._s, handle, _fp(wo), _up(seed), n, _fp(wi), _fp(f), _fp(pdf), _up(delta), _up(seed_out)) != 0:
            raise ValueError("oracle: BSDF handle out of range")
        return wi, f, pdf, delta, seed_out

    def bsdf_eval(self, handle, wo, wi):
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, np.float32).reshape(-1, 3)
        n = len(wo)
        f, pdf, delta = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint32)
        if lib().or_bsdf_eval(self._s, handle, _fp(wo), _fp(wi), n, _fp(f), _fp(pdf), _up(delta)) != 0:
            raise ValueError("oracle: BSDF handle out of range")
        return f, pdf, delta

    def sample_light(self, pos, seed):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        seed = np.ascontiguousarray(seed, np.uint32)
        n = len(pos)
        lpos, em = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        pdf, seed_out = np.zeros(n, np.float32), np.zeros(n, np.uint32)
        lib().or_sample_light(self._s, _fp(pos), _up(seed), n, _fp(lpos), _fp(em), _fp(pdf), _up(seed_out))
        return lpos, em, pdf, seed_out


class _Texture(C.Structure):     # gsb_texture (include/gpuspectral_b200.h), declared here so that the archive reader needs no product import
    _fields_ = [("type", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("texels", C.POINTER(C.c_float)),
                ("color0", C.c_float * 3), ("color1", C.c_float * 3),
                ("uscale", C.c_float), ("vscale", C.c_float), ("uoffset", C.c_float), ("voffset", C.c_float)]


def archive_textures(z):
    """(gsb_texture array, count, texel arrays to keep alive) of a scene archive."""
    nt = int(z["texture_count"][0]) if "texture_count" in z else 0
    arr = (_Texture * max(1, nt))()
    keep = []
    for t in range(nt):
        d = z[f"tex{t}_desc"]
        arr[t].type = int(d[0])
        arr[t].uscale, arr[t].vscale, arr[t].uoffset, arr[t].voffset = (float(x) for x in d[1:5])
        arr[t].color0[:] = [float(x) for x in d[5:8]]
        arr[t].color1[:] = [float(x) for x in d[8:11]]
        if arr[t].type == 1:
            px = np.ascontiguousarray(z[f"tex{t}_px"], np.float32)
            keep.append(px)
            arr[t].height, arr[t].width = px.shape[:2]
            arr[t].texels = px.ctypes.data_as(C.POINTER(C.c_float))
    return arr, nt, keep


def rng(idx, timestamp, n_draws):
    idx = np.ascontiguousarray(idx, np.uint32)
    ts = np.ascontiguousarray(timestamp, np.uint32)
    out = np.zeros((len(idx), 1 + n_draws), np.uint32)
    lib().or_rng(_up(idx), _up(ts), len(idx), n_draws, _up(out))
    return out


def tea(v0, v1):
    return lib().or_tea(v0, v1)


def pcg_hash(v):
    return lib().or_pcg_hash(v)


def rand_to_float(word):
    return lib().or_rand_to_float(word)


def encode_srgb8(rgba):
    rgba = np.ascontiguousarray(rgba, np.float32)
    out = np.zeros(rgba.shape, np.uint8)
    lib().or_encode_srgb8(_fp(rgba), rgba.size // 4, out.ctypes.data_as(C.POINTER(C.c_uint8)))
    return out

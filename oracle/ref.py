"""ctypes binding of oracle/_ref/libgsref.so: THE REFERENCE'S OWN SHADERS compiled as C++.  TEST INFRASTRUCTURE ONLY.

libgsref.so is built from /root/reference's shader sources where they lie (oracle/Makefile, oracle/ref/glsl2cpp.py,
oracle/ref/ref_glue.cpp) and exists to pin oracle/oracle.cpp: tests compare the restatement with the shader text
itself.  The ray/triangle search has no source in the reference (it is the Vulkan driver's), so whole-shader runs take
the oracle's intersector as the traceRayEXT callback.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import oracle as orc

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libgsref.so")
REFERENCE_SHADERS = "/root/reference/src/GPUSpectral/assets/shaders"

BOUNCE_DTYPE = orc.BOUNCE_DTYPE
PAYLOAD_WORDS = 17   # HitPayload, pt_common.glsl:4-14


def available():
    """True when the library exists (prebuilt) or can be built here (the reference checkout is present)."""
    return os.path.exists(_SO) or os.path.exists(os.path.join(REFERENCE_SHADERS, "rayhit.rchit"))


def build(force=False):
    if os.path.exists(os.path.join(REFERENCE_SHADERS, "rayhit.rchit")):
        cmd = ["make", "-C", _HERE, "_ref/libgsref.so"] + (["-B"] if force else [])
        subprocess.run(cmd, check=True, capture_output=True)
    return _SO


_lib = None
INTERSECT_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_float, C.c_float, C.c_int, C.c_void_p)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        P, u32p, f32p = C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_float)
        u64p = C.POINTER(C.c_uint64)
        sig = {
            "ref_scene_create": (P, []),
            "ref_scene_destroy": (None, [P]),
            "ref_scene_add_mesh": (C.c_uint32, [P, f32p, f32p, C.c_uint32]),
            "ref_scene_set_instances": (C.c_int, [P, P, C.c_uint32]),
            "ref_scene_set_bsdfs": (C.c_int, [P, C.c_uint32, P, C.c_uint32]),
            "ref_scene_set_lights": (C.c_int, [P, P, C.c_uint32]),
            "ref_scene_set_camera": (None, [P, f32p, f32p, C.c_float]),
            "ref_tea": (C.c_uint32, [C.c_uint32, C.c_uint32]),
            "ref_pcg_hash": (C.c_uint32, [C.c_uint32]),
            "ref_rng": (None, [u32p, u32p, C.c_uint32, C.c_uint32, u32p]),
            "ref_rand_uniform": (C.c_float, [C.c_uint32, u32p]),
            "ref_onb": (None, [f32p, f32p, C.c_uint32, f32p, f32p, f32p]),
            "ref_bsdf_sample": (C.c_int, [P, C.c_uint32, f32p, u32p, C.c_uint32, f32p, f32p, f32p, u32p, u32p]),
            "ref_bsdf_eval": (C.c_int, [P, C.c_uint32, f32p, f32p, C.c_uint32, f32p, f32p, u32p]),
            "ref_sample_light": (C.c_int, [P, f32p, u32p, C.c_uint32, f32p, f32p, f32p, u32p]),
            "ref_camera_rays": (None, [P, C.c_uint32, C.c_uint32, u32p, u32p, C.c_uint32, P]),
            "ref_render": (C.c_int, [P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, f32p, P, P, C.c_int, u64p, u64p]),
            "ref_path_probe": (C.c_int, [P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, f32p, P, C.c_uint32, P, P]),
            "ref_closest_hit": (C.c_int, [P, f32p, P, f32p, P, P]),
            "ref_set_max_depth": (C.c_int, [C.c_int]),
        }
        for name, (res, args) in sig.items():
            f = getattr(L, name)
            f.restype, f.argtypes = res, args
        _lib = L
    return _lib


_fp, _up = orc._fp, orc._up


def set_max_depth(depth=0):
    """raygen.rgen's MAX_DEPTH for the following renders (process-wide); 0 restores the shader's literal (50)."""
    return lib().ref_set_max_depth(int(depth))


def _oracle_intersector():
    """Address of the oracle's or_intersect (the traceRayEXT callback)."""
    return C.cast(orc.lib().or_intersect, C.c_void_p)


class RefScene:
    """The scene as PathTracer::prepareScene hands it to the shaders, fed with the same flat records as the oracle."""

    def __init__(self):
        self._s = C.c_void_p(lib().ref_scene_create())

    def __del__(self):
        s, self._s = getattr(self, "_s", None), None
        if s and _lib is not None:
            _lib.ref_scene_destroy(s)

    @staticmethod
    def from_npz(path):
        """The flat records of a scene archive (oracle.OracleScene.dump_host_scene)."""
        z = np.load(path)
        r = RefScene()
        for i in range(int(z["mesh_count"][0])):
            r.add_mesh(z[f"mesh{i}_p"], z[f"mesh{i}_n"])
        for t in range(8):
            r.set_bsdfs(t, z[f"bsdf{t}"].tobytes(), int(z[f"bsdf{t}_count"][0]))
        r.set_lights(C.create_string_buffer(z["lights"].tobytes(), max(1, z["lights"].size)), int(z["lights_count"][0]))
        r.set_instances(C.create_string_buffer(z["instances"].tobytes(), max(1, z["instances"].size)), int(z["instances_count"][0]))
        r.set_camera(z["eye"], z["to_world"], float(z["fov"][0]))
        return r

    @staticmethod
    def from_host_scene(scene):
        r = RefScene()
        for i in range(scene.mesh_count):
            r.add_mesh(*scene.mesh(i))
        for t in range(8):   # the reference's eight models; the rough-dielectric extension has no reference source
            raw, n = scene.bsdfs(t)
            r.set_bsdfs(t, raw, n)
        lights, nl = scene.lights
        r.set_lights(lights, nl)
        inst, ni = scene.instances
        r.set_instances(inst, ni)
        r.set_camera(*scene.camera)
        return r

    def add_mesh(self, positions, normals):
        p = np.ascontiguousarray(positions, np.float32).reshape(-1, 3)
        q = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
        return lib().ref_scene_add_mesh(self._s, _fp(p), _fp(q), len(p))

    def set_instances(self, instances, count):
        if lib().ref_scene_set_instances(self._s, C.cast(instances, C.c_void_p), count) != 0:
            raise ValueError("ref: instance references an unknown mesh")

    def set_bsdfs(self, btype, raw, count):
        buf = C.create_string_buffer(raw, max(1, len(raw)))
        if lib().ref_scene_set_bsdfs(self._s, btype, C.cast(buf, C.c_void_p), count) != 0:
            raise ValueError("ref: the reference has BSDF types 0..7 only")

    def set_lights(self, lights, count):
        lib().ref_scene_set_lights(self._s, C.cast(lights, C.c_void_p), count)

    def set_camera(self, eye, to_world, fov):
        eye = np.ascontiguousarray(eye, np.float32)
        tw = np.ascontiguousarray(to_world, np.float32).reshape(16)
        lib().ref_scene_set_camera(self._s, _fp(eye), _fp(tw), fov)

    # ---- unit level -------------------------------------------------------------------------------------------
    def bsdf_sample(self, handle, wo, seed):
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        seed = np.ascontiguousarray(seed, np.uint32)
        n = len(wo)
        wi, f = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        pdf, delta, seed_out = np.zeros(n, np.float32), np.zeros(n, np.uint32), np.zeros(n, np.uint32)
        lib().ref_bsdf_sample(self._s, handle, _fp(wo), _up(seed), n, _fp(wi), _fp(f), _fp(pdf), _up(delta), _up(seed_out))
        return wi, f, pdf, delta, seed_out

    def bsdf_eval(self, handle, wo, wi):
        wo = np.ascontiguousarray(wo, np.float32).reshape(-1, 3)
        wi = np.ascontiguousarray(wi, np.float32).reshape(-1, 3)
        n = len(wo)
        f, pdf, delta = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.uint32)
        lib().ref_bsdf_eval(self._s, handle, _fp(wo), _fp(wi), n, _fp(f), _fp(pdf), _up(delta))
        return f, pdf, delta

    def sample_light(self, pos, seed):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        seed = np.ascontiguousarray(seed, np.uint32)
        n = len(pos)
        lpos, em = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        pdf, seed_out = np.zeros(n, np.float32), np.zeros(n, np.uint32)
        if lib().ref_sample_light(self._s, _fp(pos), _up(seed), n, _fp(lpos), _fp(em), _fp(pdf), _up(seed_out)) != 0:
            raise ValueError("ref: sampleLight with numLights == 0 divides by zero in the shader")
        return lpos, em, pdf, seed_out

    def camera_rays(self, width, height, px, py):
        px = np.ascontiguousarray(px, np.uint32)
        py = np.ascontiguousarray(py, np.uint32)
        out = np.zeros(len(px), orc.RAY_DTYPE)
        lib().ref_camera_rays(self._s, width, height, _up(px), _up(py), len(px), out.ctypes.data_as(C.c_void_p))
        return out

    # ---- whole shaders (the oracle scene supplies traceRayEXT's ray/triangle search) ------------------------------
    def render(self, oracle_scene, width, height, first=0, n=1, image=None, threads=0):
        """raygen.rgen main over the frame for timestamps first..first+n-1; returns (image, ext rays, shadow rays)."""
        img = image if image is not None else np.zeros((height, width, 4), np.float32)
        ext, sh = C.c_uint64(0), C.c_uint64(0)
        threads = threads or (os.cpu_count() or 1)
        lib().ref_render(self._s, width, height, first, n, _fp(img), _oracle_intersector(), oracle_scene._s, threads,
                         C.byref(ext), C.byref(sh))
        return img, ext.value, sh.value

    def path_probe(self, oracle_scene, width, height, x, y, timestamp, max_bounces=64):
        rgb = np.zeros(3, np.float32)
        b = np.zeros(max_bounces, BOUNCE_DTYPE)
        n = lib().ref_path_probe(self._s, width, height, x, y, timestamp, _fp(rgb), b.ctypes.data_as(C.c_void_p), max_bounces,
                                 _oracle_intersector(), oracle_scene._s)
        return rgb, b[:min(n, max_bounces)]

    def closest_hit(self, oracle_scene, payload17, hit):
        """rayhit.rchit main once.  payload17: float32[17] viewed words of HitPayload; hit: one HIT_DTYPE record."""
        pin = np.ascontiguousarray(payload17, np.float32).reshape(PAYLOAD_WORDS)
        pout = np.zeros(PAYLOAD_WORDS, np.float32)
        h = np.ascontiguousarray(hit, orc.HIT_DTYPE).reshape(1)
        lib().ref_closest_hit(self._s, _fp(pin), h.ctypes.data_as(C.c_void_p), _fp(pout), _oracle_intersector(), oracle_scene._s)
        return pout


def rng(idx, timestamp, n_draws):
    idx = np.ascontiguousarray(idx, np.uint32)
    ts = np.ascontiguousarray(timestamp, np.uint32)
    out = np.zeros((len(idx), 1 + n_draws), np.uint32)
    lib().ref_rng(_up(idx), _up(ts), len(idx), n_draws, _up(out))
    return out


def tea(v0, v1):
    return lib().ref_tea(v0, v1)


def pcg_hash(v):
    return lib().ref_pcg_hash(v)


def rand_uniform(state):
    out = C.c_uint32(0)
    u = lib().ref_rand_uniform(state, C.byref(out))
    return u, out.value


def onb(normals, vectors):
    n = np.ascontiguousarray(normals, np.float32).reshape(-1, 3)
    v = np.ascontiguousarray(vectors, np.float32).reshape(-1, 3)
    frame, a, b = np.zeros((len(n), 9), np.float32), np.zeros((len(n), 3), np.float32), np.zeros((len(n), 3), np.float32)
    lib().ref_onb(_fp(n), _fp(v), len(n), _fp(frame), _fp(a), _fp(b))
    return frame, a, b

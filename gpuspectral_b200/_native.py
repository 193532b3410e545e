"""ctypes bindings of the two native libraries.

libgpuspectral_b200.so  the CUDA core behind include/gpuspectral_b200.h
libgsb_host.so          the C++ host layer behind include/gpuspectral_b200_host.h

There is no Python or CPU fallback: if the libraries are missing the import fails, and every
call that needs a GPU fails with the library's own error message when none is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBDIR = os.path.join(_HERE, "lib")

BSDF_TYPE_COUNT = 9
NO_HIT = 0xFFFFFFFF
(BSDF_DIFFUSE, BSDF_SMOOTH_DIELECTRIC, BSDF_SMOOTH_CONDUCTOR, BSDF_SMOOTH_PLASTIC,
 BSDF_ROUGH_CONDUCTOR, BSDF_SMOOTH_FLOOR, BSDF_ROUGH_FLOOR, BSDF_ROUGH_PLASTIC, BSDF_ROUGH_DIELECTRIC) = range(9)
BSDF_NAMES = ("diffuse", "smooth_dielectric", "smooth_conductor", "smooth_plastic",
              "rough_conductor", "smooth_floor", "rough_floor", "rough_plastic", "rough_dielectric")
# sizeof of the parameter records, renderer/Scene.h:29-82 of the reference
BSDF_STRIDES = (16, 8, 8, 24, 44, 16, 20, 32, 16)   # last: the rough dielectric extension


class Instance(C.Structure):
    _fields_ = [("transform", C.c_float * 16), ("mesh", C.c_uint32), ("emission", C.c_float * 3),
                ("bsdf", C.c_uint32), ("twofaced", C.c_uint32)]


class TriangleLight(C.Structure):
    _fields_ = [("positions", (C.c_float * 4) * 3), ("radiance", C.c_float * 4)]


class Texture(C.Structure):
    _fields_ = [("type", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("texels", C.POINTER(C.c_float)),
                ("color0", C.c_float * 3), ("color1", C.c_float * 3),
                ("uscale", C.c_float), ("vscale", C.c_float), ("uoffset", C.c_float), ("voffset", C.c_float)]


class Environment(C.Structure):
    _fields_ = [("type", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("texels", C.POINTER(C.c_float)),
                ("scale", C.c_float * 3), ("to_world", C.c_float * 9)]


class Ray(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("tmin", C.c_float), ("dir", C.c_float * 3), ("tmax", C.c_float)]


class Hit(C.Structure):
    _fields_ = [("t", C.c_float), ("u", C.c_float), ("v", C.c_float), ("inst", C.c_uint32), ("prim", C.c_uint32)]


class Params(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("rr_start", C.c_int32), ("cutoff", C.c_float), ("nee", C.c_int32), ("pixel_filter", C.c_int32)]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("paths", "extension_rays", "shadow_rays", "iterations",
                                          "nodes_visited", "tris_tested", "inst_entered", "rays_counted", "shaded_vertices")]


class BvhStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("num_meshes", "num_instances", "unique_triangles", "instanced_triangles",
                                          "blas_nodes", "tlas_nodes", "bvh_bytes")] + \
               [("build_ms", C.c_double), ("sah_cost_before", C.c_double), ("sah_cost_after", C.c_double)] + \
               [(n, C.c_uint64) for n in ("baked_instances", "baked_triangles", "world_nodes", "refits")] + [("refit_ms", C.c_double)]


assert C.sizeof(Instance) == 88 and C.sizeof(TriangleLight) == 64 and C.sizeof(Ray) == 32 and C.sizeof(Hit) == 20

# name -> (restype, argtypes); the complete export list of include/gpuspectral_b200.h
_P = C.c_void_p
_u32p = C.POINTER(C.c_uint32)
_f32p = C.POINTER(C.c_float)
CORE_API = {
    "gsb_create": (C.c_int, [C.c_int, C.POINTER(_P)]),
    "gsb_destroy": (C.c_int, [_P]),
    "gsb_last_error": (C.c_char_p, []),
    "gsb_version": (C.c_char_p, []),
    "gsb_mesh_create": (C.c_int, [_P, _f32p, _f32p, C.c_uint32, _u32p]),
    "gsb_mesh_destroy": (C.c_int, [_P, C.c_uint32]),
    "gsb_scene_set_instances": (C.c_int, [_P, C.POINTER(Instance), C.c_uint32]),
    "gsb_mesh_create_uv": (C.c_int, [_P, _f32p, _f32p, _f32p, C.c_uint32, C.POINTER(C.c_uint32)]),
    "gsb_scene_set_textures": (C.c_int, [_P, C.POINTER(Texture), C.c_uint32]),
    "gsb_scene_set_environment": (C.c_int, [_P, C.POINTER(Environment)]),
    "gsb_scene_update_transforms": (C.c_int, [_P, C.c_uint32, C.c_uint32, C.POINTER(C.c_float)]),
    "gsb_scene_set_bsdfs": (C.c_int, [_P, C.c_uint32, _P, C.c_uint32]),
    "gsb_scene_set_lights": (C.c_int, [_P, C.POINTER(TriangleLight), C.c_uint32]),
    "gsb_set_camera": (C.c_int, [_P, _f32p, _f32p, C.c_float]),
    "gsb_default_params": (C.c_int, [C.POINTER(Params)]),
    "gsb_set_params": (C.c_int, [_P, C.POINTER(Params)]),
    "gsb_get_bvh_stats": (C.c_int, [_P, C.POINTER(BvhStats)]),
    "gsb_render": (C.c_int, [_P, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]),
    "gsb_clear": (C.c_int, [_P]),
    "gsb_synchronize": (C.c_int, [_P]),
    "gsb_read_hdr": (C.c_int, [_P, _f32p]),
    "gsb_read_ldr": (C.c_int, [_P, C.POINTER(C.c_uint8)]),
    "gsb_device_sum_ptr": (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "gsb_read_sum": (C.c_int, [_P, C.POINTER(C.c_float)]),
    "gsb_write_sum": (C.c_int, [_P, C.c_uint32, C.c_uint32, C.POINTER(C.c_float)]),
    "gsb_get_counters": (C.c_int, [_P, C.POINTER(Counters)]),
    "gsb_reset_counters": (C.c_int, [_P]),
    "gsb_set_profiling": (C.c_int, [_P, C.c_int]),
    "gsb_last_render_ms": (C.c_int, [_P, C.POINTER(C.c_double)]),
    "gsb_set_instrumented": (C.c_int, [_P, C.c_int]),
    "gsb_set_stream": (C.c_int, [_P, _P]),
    "gsb_get_launch_info": (C.c_int, [_P, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "gsb_trace_batch": (C.c_int, [_P, C.POINTER(Ray), C.c_uint32, C.POINTER(Hit), C.c_int]),
    "gsb_trace_batch_stats": (C.c_int, [_P, C.POINTER(Ray), C.c_uint32, C.POINTER(Hit), C.c_int]),
    "gsb_trace_batch_timed": (C.c_int, [_P, C.POINTER(Ray), C.c_uint32, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "gsb_trace_batch_static": (C.c_int, [_P, C.POINTER(Ray), C.c_uint32, C.POINTER(Hit), C.c_int]),
    "gsb_test_rng": (C.c_int, [_P, _u32p, _u32p, C.c_uint32, C.c_uint32, _u32p]),
    "gsb_test_camera_rays": (C.c_int, [_P, C.c_uint32, C.c_uint32, _u32p, _u32p, C.c_uint32, C.POINTER(Ray)]),
    "gsb_test_bsdf_sample": (C.c_int, [_P, C.c_uint32, _f32p, _u32p, C.c_uint32, _f32p, _f32p, _f32p, _u32p, _u32p]),
    "gsb_test_bsdf_eval": (C.c_int, [_P, C.c_uint32, _f32p, _f32p, C.c_uint32, _f32p, _f32p, _u32p]),
    "gsb_test_sample_light": (C.c_int, [_P, _f32p, _u32p, C.c_uint32, _f32p, _f32p, _f32p, _u32p]),
}

HOST_API = {
    "gsbh_last_error": (C.c_char_p, []),
    "gsbh_scene_load_xml": (_P, [C.c_char_p]),
    "gsbh_scene_load_xml_string": (_P, [C.c_char_p, C.c_char_p]),
    "gsbh_scene_load_gltf": (_P, [C.c_char_p]),
    "gsbh_scene_builtin": (_P, [C.c_char_p, C.POINTER(C.c_int), C.c_int]),
    "gsbh_scene_free": (None, [_P]),
    "gsbh_scene_mesh_count": (C.c_uint32, [_P]),
    "gsbh_scene_mesh_vertex_count": (C.c_uint32, [_P, C.c_uint32]),
    "gsbh_scene_mesh_positions": (_f32p, [_P, C.c_uint32]),
    "gsbh_scene_mesh_normals": (_f32p, [_P, C.c_uint32]),
    "gsbh_scene_mesh_uvs": (_f32p, [_P, C.c_uint32]),
    "gsbh_scene_texture_count": (C.c_uint32, [_P]),
    "gsbh_scene_environment": (None, [_P, C.POINTER(Environment)]),
    "gsbh_scene_copy_textures": (None, [_P, C.POINTER(Texture)]),
    "gsbh_scene_mesh_name": (C.c_char_p, [_P, C.c_uint32]),
    "gsbh_scene_instance_count": (C.c_uint32, [_P]),
    "gsbh_scene_copy_instances": (None, [_P, C.POINTER(Instance)]),
    "gsbh_scene_bsdf_count": (C.c_uint32, [_P, C.c_uint32]),
    "gsbh_scene_bsdf_data": (_P, [_P, C.c_uint32]),
    "gsbh_scene_bsdf_stride": (C.c_uint32, [C.c_uint32]),
    "gsbh_scene_light_count": (C.c_uint32, [_P]),
    "gsbh_scene_lights": (C.POINTER(TriangleLight), [_P]),
    "gsbh_scene_camera": (None, [_P, _f32p, _f32p, _f32p]),
    "gsbh_scene_film": (None, [_P, _u32p, _u32p]),
    "gsbh_scene_instanced_triangles": (C.c_uint64, [_P]),
    "gsbh_scene_warning_count": (C.c_uint32, [_P]),
    "gsbh_scene_warning": (C.c_char_p, [_P, C.c_uint32]),
    "gsbh_tracer_create": (_P, [C.c_int]),
    "gsbh_tracer_destroy": (None, [_P]),
    "gsbh_tracer_context": (_P, [_P]),
    "gsbh_tracer_setup": (C.c_int, [_P, C.c_uint32, C.c_uint32]),
    "gsbh_tracer_set_params": (C.c_int, [_P, C.POINTER(Params)]),
    "gsbh_tracer_set_timestamp": (C.c_int, [_P, C.c_int]),
    "gsbh_tracer_upload": (C.c_int, [_P, _P]),
    "gsbh_tracer_create_render_pass": (C.c_int, [_P, _P]),
    "gsbh_tracer_render": (C.c_int, [_P, _P, C.c_uint32, C.c_uint32]),
    "gsbh_tracer_read_hdr": (C.c_int, [_P, _f32p]),
    "gsbh_tracer_read_ldr": (C.c_int, [_P, C.POINTER(C.c_uint8)]),
    "gsbh_tracer_synchronize": (C.c_int, [_P]),
    "gsbh_write_pfm": (C.c_int, [C.c_char_p, _f32p, C.c_uint32, C.c_uint32]),
    "gsbh_read_pfm": (C.c_int, [C.c_char_p, _f32p, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "gsbh_image_diff": (C.c_int, [_f32p, _f32p, C.c_uint32, C.c_uint32, C.POINTER(C.c_double)]),
    "gsbh_write_png": (C.c_int, [C.c_char_p, C.POINTER(C.c_uint8), C.c_uint32, C.c_uint32]),
}


def _load(name, api):
    path = os.path.join(_LIBDIR, name)
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build the native libraries first (python -c 'import __graft_entry__ as g; g.build()' "
            f"or make -C {_HERE}). gpuspectral_b200 has no Python/CPU fallback.")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    for fn, (res, args) in api.items():
        f = getattr(lib, fn)  # AttributeError here = the library does not export what the header declares
        f.restype = res
        f.argtypes = args
    return lib


core = _load("libgpuspectral_b200.so", CORE_API)
host = _load("libgsb_host.so", HOST_API)
CORE_PATH = os.path.join(_LIBDIR, "libgpuspectral_b200.so")
HOST_PATH = os.path.join(_LIBDIR, "libgsb_host.so")

/*
 * gpuspectral_b200_host.h — flat C view of the C++ host layer (libgsb_host.so).
 *
 * The host layer is the caller side of the render-path boundary: the scene model of
 * renderer/Scene.h, the Mitsuba-XML/OBJ loader of engine/Loader.cpp:253-349 (headless) and
 * the PathTracer render pass of renderer/PathTracer.cpp.  These entry points let a harness
 * (tests, bench, the Python package) load or generate a scene once and hand the very same
 * arrays to the CUDA core and to the checker.
 */
#ifndef GPUSPECTRAL_B200_HOST_H
#define GPUSPECTRAL_B200_HOST_H

#include "gpuspectral_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gsbh_scene gsbh_scene;
typedef struct gsbh_tracer gsbh_tracer;

GSB_API const char* gsbh_last_error(void);

/* Mitsuba XML (0.5 / 2 dialects) with obj / rectangle / cube shapes. NULL on failure. */
GSB_API gsbh_scene* gsbh_scene_load_xml(const char* path);
GSB_API gsbh_scene* gsbh_scene_load_xml_string(const char* xml, const char* base_dir);
/* glTF 2.0 (.gltf with external or data-URI buffers, .glb): triangle meshes, node transforms, the first perspective
 * camera, pbrMetallicRoughness factors mapped onto the shader's BSDFs (host/gltf.cpp). NULL on failure. */
GSB_API gsbh_scene* gsbh_scene_load_gltf(const char* path);
/* Built-in scenes: "cornell" | "gallery" | "field" (params: seed, unique meshes, subdiv, grid n;
 * missing trailing params take defaults 1234, 16, 7, 13). NULL on failure. */
GSB_API gsbh_scene* gsbh_scene_builtin(const char* name, const int* params, int nparams);
GSB_API void gsbh_scene_free(gsbh_scene* s);

GSB_API uint32_t gsbh_scene_mesh_count(const gsbh_scene* s);
GSB_API uint32_t gsbh_scene_mesh_vertex_count(const gsbh_scene* s, uint32_t mesh);
GSB_API const float* gsbh_scene_mesh_positions(const gsbh_scene* s, uint32_t mesh);
GSB_API const float* gsbh_scene_mesh_normals(const gsbh_scene* s, uint32_t mesh);
GSB_API const float* gsbh_scene_mesh_uvs(const gsbh_scene* s, uint32_t mesh);   /* 2 floats per vertex, or NULL */
GSB_API void gsbh_scene_environment(const gsbh_scene* s, gsb_environment* out);  /* type 0 = none; texel pointer belongs to the scene */
GSB_API uint32_t gsbh_scene_texture_count(const gsbh_scene* s);
GSB_API void gsbh_scene_copy_textures(const gsbh_scene* s, gsb_texture* out);   /* texel pointers belong to the scene */
GSB_API const char* gsbh_scene_mesh_name(const gsbh_scene* s, uint32_t mesh);
GSB_API uint32_t gsbh_scene_instance_count(const gsbh_scene* s);
GSB_API void gsbh_scene_copy_instances(const gsbh_scene* s, gsb_instance* out); /* mesh = scene mesh index */
GSB_API uint32_t gsbh_scene_bsdf_count(const gsbh_scene* s, uint32_t type);
GSB_API const void* gsbh_scene_bsdf_data(const gsbh_scene* s, uint32_t type);
GSB_API uint32_t gsbh_scene_bsdf_stride(uint32_t type);
GSB_API uint32_t gsbh_scene_light_count(const gsbh_scene* s);
GSB_API const gsb_triangle_light* gsbh_scene_lights(const gsbh_scene* s);
GSB_API void gsbh_scene_camera(const gsbh_scene* s, float eye[3], float to_world[16], float* fov_radians);
GSB_API void gsbh_scene_film(const gsbh_scene* s, uint32_t* width, uint32_t* height);
GSB_API uint64_t gsbh_scene_instanced_triangles(const gsbh_scene* s);
GSB_API uint32_t gsbh_scene_warning_count(const gsbh_scene* s);
GSB_API const char* gsbh_scene_warning(const gsbh_scene* s, uint32_t i);

/* PathTracer render pass (renderer/PathTracer.h:48-68) over the CUDA core. */
GSB_API gsbh_tracer* gsbh_tracer_create(int device_ordinal);
GSB_API void gsbh_tracer_destroy(gsbh_tracer* t);
GSB_API gsb_context* gsbh_tracer_context(gsbh_tracer* t);
GSB_API int gsbh_tracer_setup(gsbh_tracer* t, uint32_t width, uint32_t height);
GSB_API int gsbh_tracer_set_params(gsbh_tracer* t, const gsb_params* p);
GSB_API int gsbh_tracer_set_timestamp(gsbh_tracer* t, int timestamp);
GSB_API int gsbh_tracer_upload(gsbh_tracer* t, const gsbh_scene* s);
GSB_API int gsbh_tracer_create_render_pass(gsbh_tracer* t, const gsbh_scene* s);            /* one frame, 1 spp */
GSB_API int gsbh_tracer_render(gsbh_tracer* t, const gsbh_scene* s, uint32_t spp, uint32_t stride);
GSB_API int gsbh_tracer_read_hdr(gsbh_tracer* t, float* out_rgba);
GSB_API int gsbh_tracer_read_ldr(gsbh_tracer* t, uint8_t* out_rgba8);
GSB_API int gsbh_tracer_synchronize(gsbh_tracer* t);

GSB_API int gsbh_write_pfm(const char* path, const float* rgba, uint32_t width, uint32_t height);
/* Reads a PFM into RGBA32F rows top-to-bottom.  rgba may be NULL to query the size; capacity_pixels = pixels rgba can hold. */
GSB_API int gsbh_read_pfm(const char* path, float* rgba, uint32_t capacity_pixels, uint32_t* out_width, uint32_t* out_height);
/* out5 = relMSE (mean of (a-b)^2 / (b^2 + 1e-2)), RMSE, mean of a, mean of b, max |a-b| over rgb; b is the reference image. */
GSB_API int gsbh_image_diff(const float* rgba_a, const float* rgba_b, uint32_t width, uint32_t height, double out5[5]);
GSB_API int gsbh_write_png(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height);

#ifdef __cplusplus
}
#endif
#endif

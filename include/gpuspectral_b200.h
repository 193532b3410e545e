/*
 * gpuspectral_b200.h — C ABI of the B200-native path-tracing core.
 *
 * This library stands where the reference's ray-tracing driver operators and
 * its PathTracer plugin glue stand (paths relative to the reference checkout):
 *
 *   HwDriver::createBLAS / createTLAS / destroyBLAS / destroyTLAS / traceRays
 *       src/GPUSpectral/backend/Command.inc:83-87
 *   RTInstance / RTSceneDescriptor / Ray / RayHit
 *       src/GPUSpectral/backend/DriverTypes.h:253-276
 *   PathTracer::setup / createRenderPass / prepareScene, RenderState
 *       src/GPUSpectral/renderer/PathTracer.cpp:5-93, PathTracer.h:9-46
 *   raygen.rgen / rayhit.rchit / miss.rmiss / shadowmiss.rmiss (the device code)
 *       src/GPUSpectral/assets/shaders/
 *
 * Conventions
 *   - plain C, no exceptions: every call returns a gsb_status; the message of
 *     the last failure on the calling thread is gsb_last_error().
 *   - the library copies every input during the call; the caller may free its
 *     arrays as soon as the call returns.  Device memory belongs to the context.
 *   - one context per GPU, externally synchronised (single-threaded use),
 *     mirroring the reference's single-threaded frame loop.
 *   - matrices are column-major float[16] (glm::mat4 memory order).
 */
#ifndef GPUSPECTRAL_B200_H
#define GPUSPECTRAL_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define GSB_API __attribute__((visibility("default")))
#else
#define GSB_API
#endif

typedef struct gsb_context gsb_context;

typedef enum gsb_status {
    GSB_OK = 0,
    GSB_ERR_INVALID = 1, /* bad argument / handle / range */
    GSB_ERR_CUDA = 2,    /* a CUDA runtime call or kernel failed */
    GSB_ERR_STATE = 3,   /* call order violated (e.g. render before scene) */
    GSB_ERR_OOM = 4
} gsb_status;

/* BSDF type ids = enum order of assets/shaders/BSDF.inc (rayhit.rchit:332-339). */
enum {
    GSB_BSDF_DIFFUSE = 0,
    GSB_BSDF_SMOOTH_DIELECTRIC = 1,
    GSB_BSDF_SMOOTH_CONDUCTOR = 2,
    GSB_BSDF_SMOOTH_PLASTIC = 3,
    GSB_BSDF_ROUGH_CONDUCTOR = 4,
    GSB_BSDF_SMOOTH_FLOOR = 5,
    GSB_BSDF_ROUGH_FLOOR = 6,
    GSB_BSDF_ROUGH_PLASTIC = 7,
    /* Extension: not in BSDF.inc.  The reference only has the dead include of a rough dielectric (engine/Loader.cpp:88-143
     * never instantiates one); north_star lists it, so it is built against Mitsuba's `roughdielectric` semantics
     * (GGX microfacet reflection + refraction, Walter et al. 2007) with the checker's own restatement as the oracle. */
    GSB_BSDF_ROUGH_DIELECTRIC = 8,
    GSB_BSDF_TYPE_COUNT = 9
};

/* BSDF handle = type<<16 | index (renderer/Scene.h:84-97, pt_common.glsl:32-42). */
#define GSB_BSDF_HANDLE(type, index) ((((uint32_t)(type)) << 16) | ((uint32_t)(index) & 0xffffu))

/* Parameter records: byte-for-byte the host structs of renderer/Scene.h:29-82
 * ("scalar" block layout on the shader side, rayhit.rchit:22-70). */
typedef struct gsb_diffuse_bsdf { float reflectance[3]; int32_t hasTexture; } gsb_diffuse_bsdf;                 /* 16 B */
typedef struct gsb_smooth_dielectric_bsdf { float iorIn; float iorOut; } gsb_smooth_dielectric_bsdf;           /*  8 B */
typedef struct gsb_smooth_conductor_bsdf { float iorIn; float iorOut; } gsb_smooth_conductor_bsdf;             /*  8 B */
typedef struct gsb_smooth_floor_bsdf { float diffuse[3]; float R0; } gsb_smooth_floor_bsdf;                    /* 16 B */
typedef struct gsb_smooth_plastic_bsdf { float diffuse[3]; float iorIn; float iorOut; float R0; } gsb_smooth_plastic_bsdf; /* 24 B */
typedef struct gsb_rough_conductor_bsdf {
    float eta[3]; float k[3]; float reflectance[3]; float alpha; int32_t hasTexture;
} gsb_rough_conductor_bsdf;                                                                                     /* 44 B */
typedef struct gsb_rough_plastic_bsdf {
    float diffuse[3]; float iorIn; float iorOut; float R0; float alpha; int32_t hasTexture;
} gsb_rough_plastic_bsdf;                                                                                       /* 32 B */
typedef struct gsb_rough_floor_bsdf { float diffuse[3]; float R0; float alpha; } gsb_rough_floor_bsdf;         /* 20 B */
/* extension (no reference struct): GGX roughness alpha as in Mitsuba, no sqrt(2) rescale */
typedef struct gsb_rough_dielectric_bsdf { float iorIn; float iorOut; float alpha; int32_t pad; } gsb_rough_dielectric_bsdf; /* 16 B */

/* renderer/Scene.h:106-109: world-space triangle light, 64 B. radiance[3] unused. */
typedef struct gsb_triangle_light { float positions[3][4]; float radiance[4]; } gsb_triangle_light;

/* One render object: RTInstance{blas, transform} (DriverTypes.h:253-256) joined with
 * RenderState::Instance (PathTracer.h:27-34).  `mesh` replaces both the BLAS handle and
 * the position/normal device addresses; the inverse-transpose normal matrix
 * (PathTracer.cpp:62) is derived inside the library.  Instance index = array order
 * = gl_InstanceID. */
typedef struct gsb_instance {
    float transform[16]; /* object -> world, column-major */
    uint32_t mesh;       /* handle from gsb_mesh_create */
    float emission[3];
    uint32_t bsdf;       /* GSB_BSDF_HANDLE */
    uint32_t twofaced;
} gsb_instance;          /* 88 B */

/* DriverTypes.h:263-270 (32 B). */
typedef struct gsb_ray { float origin[3]; float tmin; float dir[3]; float tmax; } gsb_ray;

/* Result of one traced ray.  Closest hit: t, barycentrics (u = weight of vertex 1,
 * v = weight of vertex 2: the hitAttributeEXT convention of rayhit.rchit:690),
 * inst = gl_InstanceID, prim = gl_PrimitiveID.  Miss: inst = prim = GSB_NO_HIT, t = tmax.
 * Any-hit mode: inst = 0 when occluded, GSB_NO_HIT when the segment is free. */
typedef struct gsb_hit { float t; float u; float v; uint32_t inst; uint32_t prim; } gsb_hit;
#define GSB_NO_HIT 0xffffffffu

/* Integrator constants that are shader literals in the reference, with the
 * reference's values as defaults (gsb_default_params):
 *   max_depth 50 (raygen.rgen:27,73), rr_start 10 (:66), cutoff 20 (:60), nee 1 (rayhit.rchit:656). */
/* pixel_filter (extension; raygen.rgen:38 holds the sub-pixel jitter only as a comment): 0 = the reference's behaviour, one
 * ray through every pixel's integer coordinate; 1 = box filter, the primary ray goes through a uniformly drawn point of the
 * pixel's unit square; 2 = tent filter of radius 1 pixel (importance-sampled, so samples keep unit weight).  The two
 * uniform draws are the first two of the pixel-sample's RNG stream, right after seeding (raygen.rgen:37). */
typedef struct gsb_params { int32_t max_depth; int32_t rr_start; float cutoff; int32_t nee; int32_t pixel_filter; } gsb_params;

/* Device-side statistics since the last gsb_clear / gsb_reset_counters. */
typedef struct gsb_counters {
    uint64_t paths;           /* pixel-samples completed */
    uint64_t extension_rays;  /* closest-hit rays traced */
    uint64_t shadow_rays;     /* any-hit rays traced */
    uint64_t iterations;      /* wavefront iterations launched */
    /* filled only by instrumented traversal (gsb_trace_batch_stats) */
    uint64_t nodes_visited;   /* 8-wide nodes fetched (80 B each) */
    uint64_t tris_tested;     /* triangles fetched (48 B each) */
    uint64_t inst_entered;    /* instance records fetched (64 B each) */
    uint64_t rays_counted;
    uint64_t shaded_vertices; /* path vertices shaded by k_shade (extension rays that hit a surface) */
} gsb_counters;

/* Acceleration-structure statistics of the current scene. */
typedef struct gsb_bvh_stats {
    uint64_t num_meshes, num_instances;
    uint64_t unique_triangles, instanced_triangles;
    uint64_t blas_nodes, tlas_nodes;   /* 8-wide nodes */
    uint64_t bvh_bytes;                /* nodes + packed triangles + instance records */
    double   build_ms;                 /* device time of the last build (all BLAS + TLAS) */
    double   sah_cost_before, sah_cost_after; /* binary-tree SAH before/after treelet refinement, summed over meshes */
    /* world level: instances of small meshes are stored as world-space triangles under one BVH of their own (no ray
     * transform, no instance record on the way to them); the TLAS covers the remaining instances */
    uint64_t baked_instances, baked_triangles, world_nodes;
    /* animated instances: refits since the last (re)build and the device time of the last one (gsb_scene_update_transforms) */
    uint64_t refits;
    double   refit_ms;
} gsb_bvh_stats;

/* Reflectance textures (extension: the reference's loader keeps `bitmap` and `checkerboard` only as commented-out code,
 * engine/Loader.cpp:122-143, and its shaders pass a `uv` to every BSDF that none of them uses, rayhit.rchit:341-654;
 * Mitsuba semantics).  A diffuse BSDF record refers to texture i by hasTexture = i + 1 (0 = the constant reflectance). */
#define GSB_TEXTURE_CHECKERBOARD 0u  /* color0 where floor(2 u') + floor(2 v') is even, color1 elsewhere; (u', v') = (u * uscale + uoffset, v * vscale + voffset) */
#define GSB_TEXTURE_BITMAP 1u        /* width x height linear-RGB texels, row 0 = top of the image (v = 1), bilinear, repeat */
typedef struct gsb_texture {
    uint32_t type;
    uint32_t width, height;      /* bitmap only */
    const float* texels;         /* bitmap only: width * height * 3 floats, copied by gsb_scene_set_textures */
    float color0[3], color1[3];  /* checkerboard only */
    float uscale, vscale, uoffset, voffset;
} gsb_texture;

/* Environment emitter (extension: engine/Loader.cpp:338-346 holds `envmap` only as commented-out code; Mitsuba's
 * `constant` and `envmap` emitters).  Rays that leave the scene gather its radiance — weighted by the path's directWeight
 * exactly like a surface emitter reached by a BSDF sample (rayhit.rchit:760-768) — and light sampling treats it as one more
 * light next to the triangle lights: chosen with the same raw draw, its direction drawn with the same two uniform draws.
 * Bitmap: latitude-longitude image, row 0 = +Y pole, u = atan2(x, -z) / 2 pi + 1/2; piecewise-constant radiance per texel,
 * importance-sampled proportionally to texel luminance x sin(theta).  to_world: row-major 3x3 rotation, local -> world. */
#define GSB_ENVIRONMENT_NONE 0u
#define GSB_ENVIRONMENT_CONSTANT 1u
#define GSB_ENVIRONMENT_LATLONG 2u
typedef struct gsb_environment {
    uint32_t type;
    uint32_t width, height;   /* latlong only */
    const float* texels;      /* latlong only: width * height * 3 floats of linear RGB, copied */
    float scale[3];           /* constant: the radiance; latlong: factor per channel */
    float to_world[9];
} gsb_environment;

/* ---- context ------------------------------------------------------------------- */
GSB_API gsb_status gsb_create(int device_ordinal, gsb_context** out_ctx);
GSB_API gsb_status gsb_destroy(gsb_context* ctx);
GSB_API const char* gsb_last_error(void);
GSB_API const char* gsb_version(void);

/* ---- scene (replaces Mesh uploads + createBLAS/createTLAS + prepareScene) -------- */
/* Non-indexed triangle soup, 3 consecutive vertices per triangle, tightly packed float3
 * positions and normals (renderer/Mesh.cpp:7-22; VulkanRays.cpp:17-35).  nverts % 3 == 0.
 * Uploads the mesh; its bottom-level BVH is built by the first gsb_scene_set_instances that
 * needs it, together with those of the scene's other meshes (one batched build, like a
 * vkCmdBuildAccelerationStructuresKHR with several infos; a mesh that is only baked into the
 * world level gets none). */
GSB_API gsb_status gsb_mesh_create(gsb_context* ctx, const float* positions3, const float* normals3,
                                   uint32_t nverts, uint32_t* out_mesh);
GSB_API gsb_status gsb_mesh_destroy(gsb_context* ctx, uint32_t mesh);
/* Replaces the instance list and builds the top-level BVH once (the reference rebuilds
 * its TLAS every frame, PathTracer.cpp:10-19). */
/* gsb_mesh_create with texture coordinates (2 floats per vertex; NULL = all zero).  Replaces: Mesh::Mesh's `texcoord`
 * attribute (renderer/Mesh.cpp:7-22,37,50), which the reference uploads and never reads. */
GSB_API gsb_status gsb_mesh_create_uv(gsb_context* ctx, const float* positions3, const float* normals3, const float* uv2,
                                      uint32_t num_vertices, uint32_t* out_mesh);
/* Sets (or, with NULL / type NONE, removes) the environment emitter. */
GSB_API gsb_status gsb_scene_set_environment(gsb_context* ctx, const gsb_environment* env);
/* Replaces the scene's texture table (may be called before or after gsb_scene_set_instances). */
GSB_API gsb_status gsb_scene_set_textures(gsb_context* ctx, const gsb_texture* textures, uint32_t count);
GSB_API gsb_status gsb_scene_set_instances(gsb_context* ctx, const gsb_instance* instances, uint32_t count);
/* Replaces the transforms (16 floats each, column-major, as gsb_instance::transform) of instances [first, first + count)
 * and refits the acceleration structures over their existing topology; falls back to a rebuild when the refit tree has
 * degraded (summed node area > 2x its value at build time).  Replaces: the per-frame TLAS rebuild from {blas, mat4} of
 * renderer/PathTracer.cpp:10-19 / backend/vulkan/VulkanRays.cpp:92-118 for scenes whose instances move. */
GSB_API gsb_status gsb_scene_update_transforms(gsb_context* ctx, uint32_t first, uint32_t count, const float* transforms);
/* type in [0, GSB_BSDF_TYPE_COUNT); params = array of the matching gsb_*_bsdf record. */
GSB_API gsb_status gsb_scene_set_bsdfs(gsb_context* ctx, uint32_t type, const void* params, uint32_t count);
GSB_API gsb_status gsb_scene_set_lights(gsb_context* ctx, const gsb_triangle_light* lights, uint32_t count);
/* eye = toWorld[3].xyz in the reference (PathTracer.cpp:88-90); passed separately as there. */
GSB_API gsb_status gsb_set_camera(gsb_context* ctx, const float eye[3], const float to_world[16], float fov_radians);
GSB_API gsb_status gsb_default_params(gsb_params* out);
GSB_API gsb_status gsb_set_params(gsb_context* ctx, const gsb_params* params);
GSB_API gsb_status gsb_get_bvh_stats(gsb_context* ctx, gsb_bvh_stats* out);

/* ---- render (replaces traceRays(pipeline, w, h), one launch = one sample/pixel) --- */
/* Renders samples with timestamp = first_sample + k*sample_stride, k in [0, num_samples),
 * for every pixel of a width x height frame and adds them to the context's accumulation
 * buffer (allocated/cleared when the size changes).  Asynchronous on the context's stream: the whole call is one
 * device-side loop (a conditional CUDA graph); gsb_synchronize, the read-back and counter calls wait for it, and
 * an error raised inside the loop (traversal stack overflow) is reported by the first of those calls. */
GSB_API gsb_status gsb_render(gsb_context* ctx, uint32_t width, uint32_t height,
                              uint32_t first_sample, uint32_t num_samples, uint32_t sample_stride);
GSB_API gsb_status gsb_clear(gsb_context* ctx);       /* zero accumulation + counters */
GSB_API gsb_status gsb_synchronize(gsb_context* ctx);
/* Mean radiance, RGBA32F, row 0 = top (raygen.rgen:84-108); alpha = 1.  Synchronises. */
GSB_API gsb_status gsb_read_hdr(gsb_context* ctx, float* out_rgba);
/* 8-bit sRGB encode of the clamped mean, no tone map: what presenting the RGBA32F image
 * to the B8G8R8A8_SRGB swapchain does (DrawTexture.frag:9-13).  RGBA byte order. */
GSB_API gsb_status gsb_read_ldr(gsb_context* ctx, uint8_t* out_rgba8);
/* Device pointer to the float4 accumulation buffer (rgb = sum of samples, a = number of
 * samples added) for an external reduction across GPUs (NCCL), and its size in bytes. */
GSB_API gsb_status gsb_device_sum_ptr(gsb_context* ctx, void** out_ptr, size_t* out_bytes);
/* Checkpoint / resume of a progressive render: the raw accumulation buffer (width x height float4: rgb = sum of samples,
 * a = samples added) out to and in from host memory.  gsb_write_sum (re)sizes the frame like gsb_render does and replaces
 * its contents; continuing with gsb_render(first_sample = samples already in the checkpoint) gives the image an
 * uninterrupted render would have given.  The reference keeps its running mean in the RGBA32F image of
 * PathTracer.cpp:5-7 and has no way to save it. */
GSB_API gsb_status gsb_read_sum(gsb_context* ctx, float* out_rgba_sum);
GSB_API gsb_status gsb_write_sum(gsb_context* ctx, uint32_t width, uint32_t height, const float* rgba_sum);
GSB_API gsb_status gsb_get_counters(gsb_context* ctx, gsb_counters* out);
GSB_API gsb_status gsb_reset_counters(gsb_context* ctx);
/* Device time (ms, CUDA events on the context's stream) of the last gsb_render, split by
 * stage: [0] total, [1] extend traversal, [2] shade, [3] shadow traversal, [4] finish/regen.
 * Stage times are only collected when profiling is enabled. */
GSB_API gsb_status gsb_set_profiling(gsb_context* ctx, int enabled);
GSB_API gsb_status gsb_last_render_ms(gsb_context* ctx, double out_ms[5]);
/* Instrumented rendering: gsb_render runs the counting variants of the two traversal kernels and
 * adds nodes / triangles / instance records fetched (and rays counted) to the counters, so the
 * algorithmic bytes per ray are those of the rays the render really traced.
 * mode: 0 off (default), 1 extension rays only, 2 shadow rays only, 3 both. */
GSB_API gsb_status gsb_set_instrumented(gsb_context* ctx, int mode);
/* Runs all of the context's work on a caller-owned CUDA stream (a cudaStream_t passed as void*;
 * NULL = back to a private stream), so that a host framework's events and collectives (NCCL on
 * the same stream) order with the render without extra synchronisation. */
GSB_API gsb_status gsb_set_stream(gsb_context* ctx, void* cuda_stream);
/* Kernels launched by gsb_render since gsb_reset_counters / gsb_clear, and the number of wavefront
 * iterations of the last gsb_render (= launches of each stage kernel in it). */
GSB_API gsb_status gsb_get_launch_info(gsb_context* ctx, uint64_t* out_launches, uint64_t* out_last_render_iterations);

/* ---- ray batches (the intersection-parity entry; Ray/RayHit of DriverTypes.h) ----- */
/* Host arrays in, host arrays out.  any_hit = 0: closest hit; 1: occlusion only. */
GSB_API gsb_status gsb_trace_batch(gsb_context* ctx, const gsb_ray* rays, uint32_t count,
                                   gsb_hit* out_hits, int any_hit);
/* Same rays through the instrumented traversal kernel: also accumulates nodes/tris/inst
 * fetch counts into the context counters (algorithmic bytes per ray). */
GSB_API gsb_status gsb_trace_batch_stats(gsb_context* ctx, const gsb_ray* rays, uint32_t count,
                                         gsb_hit* out_hits, int any_hit);
/* Device-resident variant for timing traversal alone: uploads once, then launches the
 * traversal kernel `repeats` times and reports the mean kernel time in ms. */
GSB_API gsb_status gsb_trace_batch_timed(gsb_context* ctx, const gsb_ray* rays, uint32_t count,
                                         int any_hit, int repeats, double* out_ms_per_launch);
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same rays in, same hits out as gsb_trace_batch, but the device copies of the rays and hits
 * and the kernel's work counter live in a file-scope `static __device__` array of the library
 * (256 B + GSB_STATIC_BATCH_MAX_RAYS * (32 + 20) B = 3.25 MiB, held by every GPU that loads the
 * library until the process exits): the call makes no cudaMalloc / cudaFree / cudaMallocAsync and
 * no other allocating CUDA or library call.  count > GSB_STATIC_BATCH_MAX_RAYS is rejected with
 * GSB_ERR_INVALID (split the batch, or use gsb_trace_batch); a pending asynchronous gsb_render is
 * rejected with GSB_ERR_STATE (call gsb_synchronize first).  Every call on a device shares the
 * array and finds the previous call's contents in it (the control words are cleared per call, the
 * used ray/hit entries are overwritten): calls that overlap on the GPU - two contexts or two host
 * threads on the same device - race and must be serialised by the caller. */
#define GSB_STATIC_BATCH_MAX_RAYS 65536u
GSB_API gsb_status gsb_trace_batch_static(gsb_context* ctx, const gsb_ray* rays, uint32_t count,
                                          gsb_hit* out_hits, int any_hit);

/* ---- diagnostic entry points (unit parity against the oracle) ---------------------- */
/* Seeds exactly as raygen.rgen:37 for pixel-linear index idx and frame `timestamp`, then
 * draws n_draws values of randPcg(); out has n*(1+n_draws) words: seed, draws... */
GSB_API gsb_status gsb_test_rng(gsb_context* ctx, const uint32_t* idx, const uint32_t* timestamp, uint32_t n,
                                uint32_t n_draws, uint32_t* out);
/* Primary rays of a width x height launch (raygen.rgen:20-35) for pixels px[i], py[i]. */
GSB_API gsb_status gsb_test_camera_rays(gsb_context* ctx, uint32_t width, uint32_t height,
                                        const uint32_t* px, const uint32_t* py, uint32_t n, gsb_ray* out);
/* sampleBSDF / evalBSDF (rayhit.rchit:630-654) on the scene's BSDF tables.
 * sample: in  wo[3n], seed[n];  out wi[3n], f[3n], pdf[n], is_delta[n], seed_out[n].
 * eval:   in  wo[3n], wi[3n];   out f[3n], pdf[n], is_delta[n]. */
GSB_API gsb_status gsb_test_bsdf_sample(gsb_context* ctx, uint32_t bsdf_handle, const float* wo, const uint32_t* seed,
                                        uint32_t n, float* wi, float* f, float* pdf, uint32_t* is_delta, uint32_t* seed_out);
GSB_API gsb_status gsb_test_bsdf_eval(gsb_context* ctx, uint32_t bsdf_handle, const float* wo, const float* wi,
                                      uint32_t n, float* f, float* pdf, uint32_t* is_delta);
/* sampleLight (rayhit.rchit:123-153): in pos[3n], seed[n]; out lpos[3n], emission[3n], pdf[n], seed_out[n]. */
GSB_API gsb_status gsb_test_sample_light(gsb_context* ctx, const float* pos, const uint32_t* seed, uint32_t n,
                                         float* lpos, float* emission, float* pdf, uint32_t* seed_out);

#ifdef __cplusplus
}
#endif
#endif /* GPUSPECTRAL_B200_H */

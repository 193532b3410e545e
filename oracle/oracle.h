/*
 * oracle.h — C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  The oracle is a CPU restatement of the reference's shader
 * math (pt_common.glsl, raygen.rgen, rayhit.rchit, miss.rmiss, shadowmiss.rmiss) used as
 * the checker for the CUDA path.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it; the product never links or calls it.
 *
 * It consumes the same flat scene records as the product's C ABI (the struct layouts of
 * include/gpuspectral_b200.h, which are themselves the reference's Scene.h layouts).
 */
#ifndef GSB_ORACLE_H
#define GSB_ORACLE_H

#include "../include/gpuspectral_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct or_scene or_scene;

typedef struct or_counters {
    uint64_t paths;
    uint64_t extension_rays;
    uint64_t shadow_rays;
    uint64_t rng_draws;     /* randPcg() calls made by closest-hit shading */
} or_counters;

/* Per-bounce record of one path (debug / parity probes). */
typedef struct or_bounce {
    uint32_t inst, prim;      /* GSB_NO_HIT on miss */
    float t;
    float emitted[3];         /* prd.emitted after this bounce */
    float weight[3];          /* prd.weight after this bounce */
    uint32_t seed;            /* prd.seed after this bounce */
    uint32_t shadow_cast;     /* 1 if a shadow ray was traced */
    uint32_t shadow_visible;  /* 1 if it reached the light */
} or_bounce;

or_scene* or_scene_create(void);
void or_scene_destroy(or_scene* s);
uint32_t or_scene_add_mesh(or_scene* s, const float* positions3, const float* normals3, uint32_t nverts);
/* texture extension: texture coordinates per vertex (NULL = zero) and the texture table, as gsb_mesh_create_uv / gsb_scene_set_textures */
uint32_t or_scene_add_mesh_uv(or_scene* s, const float* positions3, const float* normals3, const float* uv2, uint32_t nverts);
int or_scene_set_textures(or_scene* s, const gsb_texture* textures, uint32_t count);
int or_scene_set_environment(or_scene* s, const gsb_environment* env);   /* NULL / type 0 removes it */
void or_env_sample(const or_scene* s, const float* e1, const float* e2, uint32_t n, float* dirs3, float* radiance3, float* pdf);
void or_env_radiance(const or_scene* s, const float* dirs3, uint32_t n, float* radiance3);
int or_scene_set_instances(or_scene* s, const gsb_instance* instances, uint32_t count);
/* number of instances whose triangles are stored in world space (the bake rule of oracle.cpp: kBakeMaxTris / kBakeBudget) */
uint32_t or_scene_baked_instances(const or_scene* s);
int or_scene_set_bsdfs(or_scene* s, uint32_t type, const void* params, uint32_t count);
int or_scene_set_lights(or_scene* s, const gsb_triangle_light* lights, uint32_t count);
void or_scene_set_camera(or_scene* s, const float eye[3], const float to_world[16], float fov_radians);
void or_scene_set_params(or_scene* s, const gsb_params* p);

/* Renders timestamps first + k*stride, k in [0,n), into `image` (RGBA32F, row 0 = top) with
 * the reference's running mean (raygen.rgen:84-108): frame index used for the blend weight
 * is the timestamp itself, exactly as in the shader, so start at first = 0, stride = 1 for
 * the reference's behaviour.  `image` must hold the previous frames' mean when first > 0.
 * If sum_mode != 0 the image instead receives rgb = plain sum of the samples and
 * a = number of samples added (the layout of the product's accumulation buffer).
 * threads <= 0: all OpenMP threads. Returns the number of threads used. */
int or_render(or_scene* s, uint32_t width, uint32_t height, uint32_t first, uint32_t n, uint32_t stride,
              float* image_rgba, int sum_mode, int threads, or_counters* counters);

/* One pixel-sample; returns radiance in out_rgb and up to max_bounces records. */
int or_path_probe(or_scene* s, uint32_t width, uint32_t height, uint32_t x, uint32_t y, uint32_t timestamp,
                  float out_rgb[3], or_bounce* bounces, uint32_t max_bounces);

/* mode 0: float32 brute force over every triangle of every instance (the parity definition)
 * mode 1: float32 through the oracle's CPU BVH (must equal mode 0)
 * mode 2: float64 Moeller-Trumbore brute force in world space (independent cross-check) */
int or_trace_batch(or_scene* s, const gsb_ray* rays, uint32_t count, gsb_hit* hits, int any_hit, int mode, int threads);

/* One ray query, callable as the `ref_intersect_fn` of oracle/ref/ref_glue.cpp (scene = or_scene*).  Returns 1 on a hit. */
int or_intersect(void* scene, const float* origin, const float* dir, float tmin, float tmax, int any_hit, gsb_hit* out);

/* unit-level functions */
uint32_t or_tea(uint32_t v0, uint32_t v1);
uint32_t or_pcg_hash(uint32_t v);
void or_rng(const uint32_t* idx, const uint32_t* timestamp, uint32_t n, uint32_t n_draws, uint32_t* out);
float or_rand_to_float(uint32_t word);
void or_camera_rays(or_scene* s, uint32_t width, uint32_t height, const uint32_t* px, const uint32_t* py,
                    uint32_t n, gsb_ray* out);
int or_bsdf_sample(or_scene* s, uint32_t handle, const float* wo, const uint32_t* seed, uint32_t n,
                   float* wi, float* f, float* pdf, uint32_t* is_delta, uint32_t* seed_out);
int or_bsdf_eval(or_scene* s, uint32_t handle, const float* wo, const float* wi, uint32_t n,
                 float* f, float* pdf, uint32_t* is_delta);
int or_sample_light(or_scene* s, const float* pos, const uint32_t* seed, uint32_t n,
                    float* lpos, float* emission, float* pdf, uint32_t* seed_out);
/* sRGB 8-bit encode used for the LDR framebuffer (linear -> sRGB OETF, clamp, round). */
void or_encode_srgb8(const float* rgba, uint32_t npixels, uint8_t* out_rgba8);

#ifdef __cplusplus
}
#endif
#endif

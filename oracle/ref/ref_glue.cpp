// ref_glue.cpp — hosts the REFERENCE'S OWN shaders, compiled as C++, behind a small C interface.
//
// TEST INFRASTRUCTURE ONLY (oracle/_ref/libgsref.so).  It pins oracle/oracle.cpp: the restatement is compared with
// this library, which executes the reference's shader text itself — pt_common.glsl, raygen.rgen, rayhit.rchit,
// miss.rmiss, shadowmiss.rmiss, read from /root/reference at build time, rewritten syntactically by
// oracle/ref/glsl2cpp.py (rules listed there) and compiled against the reference's vendored external/glm with
// -ffp-contract=off.  Nothing under gpuspectral_b200/ links, loads or calls it.
//
// What this file supplies is exactly what the Vulkan ray-tracing pipeline supplies to those shaders and nothing more:
//   - the interface variables: payload 0 (`prd`), payload 2 (`shadowed`), `attribs`, the gl_* built-ins, `renderState`,
//     `image` (imageLoad / imageStore on an RGBA32F host array);
//   - buffer references (a uvec2 device address -> a host pointer);
//   - traceRayEXT: the ray/triangle search itself lives inside the Vulkan driver / RT cores and has no source in the
//     reference, so it is a CALLBACK the test passes in (the oracle's intersector).  The dispatch around it is
//     the pipeline's: closest hit -> rayhit.rchit main, miss index 0 -> miss.rmiss main, miss index 1 -> shadowmiss.rmiss
//     main, TerminateOnFirstHit | SkipClosestHitShader -> no shader on a hit (VulkanPipelineCache.cpp:379-452 builds the
//     groups in that order; raygen.rgen:54-58, rayhit.rchit:737-748 are the two call sites);
//   - the host-side fill of RenderState (PathTracer.cpp:58-93): transformInvT = glm::inverse(glm::transpose(M)).
#define GLM_FORCE_SWIZZLE
#include <glm/glm.hpp>

#include <cmath>
#include <cstdint>
#include <cstring>
#include <vector>

#include "../../include/gpuspectral_b200.h"

#ifdef M_PI
#undef M_PI
#endif

// A ray query: returns 1 on a hit and fills `hit` (closest hit: t, u, v, inst, prim; any hit: anything).
typedef int (*ref_intersect_fn)(void* user, const float* origin, const float* dir, float tmin, float tmax, int any_hit, gsb_hit* hit);

// per-bounce record, same layout as oracle.h's or_bounce
struct ref_bounce {
    uint32_t inst, prim;
    float t;
    float emitted[3];
    float weight[3];
    uint32_t seed;
    uint32_t shadow_cast;
    uint32_t shadow_visible;
};

namespace gsref {
using namespace glm;

// GLSL converts an int operand to float implicitly; glm's operators are templates on one scalar type
inline vec3 operator*(int s, const vec3& v) { return float(s) * v; }
inline vec3 operator*(const vec3& v, int s) { return v * float(s); }

typedef int accelerationStructureEXT;
struct image2D {
    float* rgba = nullptr;   // RGBA32F, row 0 = top
    uint32_t width = 0;
};
inline vec4 imageLoad(const image2D& img, ivec2 p) {
    const float* q = img.rgba + 4 * ((size_t)p.y * img.width + (size_t)p.x);
    return vec4(q[0], q[1], q[2], q[3]);
}
inline void imageStore(const image2D& img, ivec2 p, vec4 v) {
    float* q = img.rgba + 4 * ((size_t)p.y * img.width + (size_t)p.x);
    q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = v.w;
}

inline const void* refPtr(uvec2 p) { return reinterpret_cast<const void*>(((uint64_t)p.y << 32) | (uint64_t)p.x); }
inline uvec2 toRef(const void* p) {
    const uint64_t a = reinterpret_cast<uint64_t>(p);
    return uvec2((uint32_t)(a & 0xffffffffull), (uint32_t)(a >> 32));
}
#define REF_BUFFER(NAME, T, FIELD)                                        \
    struct NAME {                                                         \
        const T* FIELD;                                                   \
        NAME() : FIELD(nullptr) {}                                        \
        explicit NAME(uvec2 p) : FIELD(static_cast<const T*>(refPtr(p))) {} \
    };

constexpr uint gl_RayFlagsOpaqueEXT = 1u;
constexpr uint gl_RayFlagsTerminateOnFirstHitEXT = 4u;
constexpr uint gl_RayFlagsSkipClosestHitShaderEXT = 8u;

// Pipeline state of one shader invocation chain (one per host thread).
struct Pipeline {
    alignas(16) unsigned char payload0[128];   // location 0: HitPayload
    bool payload2 = false;                     // location 2: shadowed
    vec3 attribs = vec3(0.0f);
    int primitiveID = 0, instanceID = 0;
    mat4x3 objectToWorld = mat4x3(1.0f);
    vec3 worldRayOrigin = vec3(0.0f), worldRayDirection = vec3(0.0f);
    float hitT = 0.0f;
    uvec3 launchID = uvec3(0u), launchSize = uvec3(0u);
    image2D image;
    const void* renderState = nullptr;
    const struct Scene* scene = nullptr;
    ref_intersect_fn intersect = nullptr;
    void* user = nullptr;
    // probes
    ref_bounce* bounces = nullptr;
    uint32_t maxBounces = 0, numBounces = 0;
    uint64_t extRays = 0, shadowRays = 0;
};
thread_local Pipeline P;
extern int g_maxDepth;   // raygen.rgen's MAX_DEPTH (glsl2cpp.py rule 11): the shader's literal unless ref_set_max_depth changed it

void traceRayEXT(accelerationStructureEXT tlas, uint rayFlags, uint cullMask, int sbtOffset, int sbtStride, int missIndex,
                 vec3 origin, float tmin, vec3 direction, float tmax, int payload);

#define prd (*reinterpret_cast<HitPayload*>(gsref::P.payload0))
#define shadowed (gsref::P.payload2)
#define attribs (gsref::P.attribs)
#define renderState (*static_cast<const RenderState*>(gsref::P.renderState))
#define image (gsref::P.image)
#define gl_PrimitiveID (gsref::P.primitiveID)
#define gl_InstanceID (gsref::P.instanceID)
#define gl_ObjectToWorldEXT (gsref::P.objectToWorld)
#define gl_WorldRayOriginEXT (gsref::P.worldRayOrigin)
#define gl_WorldRayDirectionEXT (gsref::P.worldRayDirection)
#define gl_HitTEXT (gsref::P.hitT)
#define gl_LaunchIDEXT (gsref::P.launchID)
#define gl_LaunchSizeEXT (gsref::P.launchSize)
static const accelerationStructureEXT topLevelAS = 0;

// Each stage is its own compilation of its source (and of pt_common.glsl), as in the SPIR-V modules: stage-private globals
// (rngState) are not shared between stages, only the payloads are.
namespace rgen {
#include "raygen.rgen.cpp"
}
namespace rchit {
#include "rayhit.rchit.cpp"
}
namespace rmiss {
#include "miss.rmiss.cpp"
}
namespace shadowmiss {
#include "shadowmiss.rmiss.cpp"
}

int g_maxDepth = REF_MAX_DEPTH_LITERAL;

#undef prd
#undef shadowed
#undef attribs
#undef renderState
#undef image
#undef gl_PrimitiveID
#undef gl_InstanceID
#undef gl_ObjectToWorldEXT
#undef gl_WorldRayOriginEXT
#undef gl_WorldRayDirectionEXT
#undef gl_HitTEXT
#undef gl_LaunchIDEXT
#undef gl_LaunchSizeEXT

static_assert(sizeof(rchit::HitPayload) <= sizeof(Pipeline::payload0), "payload storage too small");
static_assert(sizeof(rchit::HitPayload) == sizeof(rgen::HitPayload), "stages disagree on the payload");
static_assert(sizeof(rchit::RenderState) == sizeof(rgen::RenderState), "stages disagree on RenderState");
static_assert(sizeof(rchit::DiffuseBSDF) == sizeof(gsb_diffuse_bsdf), "DiffuseBSDF layout");
static_assert(sizeof(rchit::SmoothDielectricBSDF) == sizeof(gsb_smooth_dielectric_bsdf), "SmoothDielectricBSDF layout");
static_assert(sizeof(rchit::SmoothConductorBSDF) == sizeof(gsb_smooth_conductor_bsdf), "SmoothConductorBSDF layout");
static_assert(sizeof(rchit::SmoothFloorBSDF) == sizeof(gsb_smooth_floor_bsdf), "SmoothFloorBSDF layout");
static_assert(sizeof(rchit::SmoothPlasticBSDF) == sizeof(gsb_smooth_plastic_bsdf), "SmoothPlasticBSDF layout");
static_assert(sizeof(rchit::RoughConductorBSDF) == sizeof(gsb_rough_conductor_bsdf), "RoughConductorBSDF layout");
static_assert(sizeof(rchit::RoughPlasticBSDF) == sizeof(gsb_rough_plastic_bsdf), "RoughPlasticBSDF layout");
static_assert(sizeof(rchit::RoughFloorBSDF) == sizeof(gsb_rough_floor_bsdf), "RoughFloorBSDF layout");

// The scene as PathTracer::prepareScene (PathTracer.cpp:58-93) hands it to the shaders.
struct Scene {
    std::vector<std::vector<vec3>> positions, normals;   // per mesh, tightly packed float3 (Mesh.cpp:7-22)
    std::vector<gsb_instance> src;
    std::vector<rchit::Instance> instances;
    std::vector<mat4> transforms;
    std::vector<unsigned char> bsdf[8];
    std::vector<rchit::TriangleLight> lights;
    rchit::RenderState rs;

    void refresh() {
        rs.scene.instances = toRef(instances.data());
        rs.scene.triangleLights = toRef(lights.data());
        rs.scene.diffuseBSDFs = toRef(bsdf[GSB_BSDF_DIFFUSE].data());
        rs.scene.smoothDielectricBSDFs = toRef(bsdf[GSB_BSDF_SMOOTH_DIELECTRIC].data());
        rs.scene.smoothConductorBSDFs = toRef(bsdf[GSB_BSDF_SMOOTH_CONDUCTOR].data());
        rs.scene.smoothPlasticBSDFs = toRef(bsdf[GSB_BSDF_SMOOTH_PLASTIC].data());
        rs.scene.roughConductorBSDFs = toRef(bsdf[GSB_BSDF_ROUGH_CONDUCTOR].data());
        rs.scene.smoothFloorBSDFs = toRef(bsdf[GSB_BSDF_SMOOTH_FLOOR].data());
        rs.scene.roughFloorBSDFs = toRef(bsdf[GSB_BSDF_ROUGH_FLOOR].data());
        rs.scene.roughPlasticBSDFs = toRef(bsdf[GSB_BSDF_ROUGH_PLASTIC].data());
        rs.scene.numLights = (int)lights.size();
    }
};

void bind(const Scene* s, ref_intersect_fn fn, void* user) {
    P.scene = s;
    P.renderState = &s->rs;
    P.intersect = fn;
    P.user = user;
}

// vkCmdTraceRaysKHR's per-ray behaviour for the two call sites of the reference.
void traceRayEXT(accelerationStructureEXT, uint rayFlags, uint, int, int, int missIndex, vec3 origin, float tmin,
                 vec3 direction, float tmax, int payload) {
    const float o[3] = {origin.x, origin.y, origin.z}, d[3] = {direction.x, direction.y, direction.z};
    gsb_hit hit;
    hit.t = tmax; hit.u = 0.0f; hit.v = 0.0f; hit.inst = GSB_NO_HIT; hit.prim = GSB_NO_HIT;
    const bool anyHit = (rayFlags & gl_RayFlagsTerminateOnFirstHitEXT) != 0u;
    const bool found = P.intersect && P.intersect(P.user, o, d, tmin, tmax, anyHit ? 1 : 0, &hit) != 0;
    if (payload == 2) {
        // shadow query (rayhit.rchit:737-748): no closest-hit shader on a hit, miss index 1 otherwise
        P.shadowRays++;
        if (!found) {
            if (missIndex == 1) shadowmiss::shader_main();
            else rmiss::shader_main();
        }
        if (P.bounces && P.numBounces < P.maxBounces) {
            P.bounces[P.numBounces].shadow_cast = 1;
            P.bounces[P.numBounces].shadow_visible = found ? 0u : 1u;
        }
        return;
    }
    P.extRays++;
    ref_bounce* rec = (P.bounces && P.numBounces < P.maxBounces) ? &P.bounces[P.numBounces] : nullptr;
    if (rec) { std::memset(rec, 0, sizeof(*rec)); rec->inst = hit.inst; rec->prim = hit.prim; rec->t = hit.t; }
    if (found && !(rayFlags & gl_RayFlagsSkipClosestHitShaderEXT)) {
        P.instanceID = (int)hit.inst;
        P.primitiveID = (int)hit.prim;
        P.hitT = hit.t;
        P.attribs = vec3(hit.u, hit.v, 0.0f);
        P.worldRayOrigin = origin;
        P.worldRayDirection = direction;
        P.objectToWorld = mat4x3(P.scene->transforms[hit.inst]);   // gl_ObjectToWorldEXT: the instance's 4x3 transform
        rchit::shader_main();
    } else if (!found) {
        if (missIndex == 0) rmiss::shader_main();
        else shadowmiss::shader_main();
    }
    if (rec) {
        const rchit::HitPayload& p = *reinterpret_cast<const rchit::HitPayload*>(P.payload0);
        rec->emitted[0] = p.emitted.x; rec->emitted[1] = p.emitted.y; rec->emitted[2] = p.emitted.z;
        rec->weight[0] = p.weight.x; rec->weight[1] = p.weight.y; rec->weight[2] = p.weight.z;
        rec->seed = p.seed;
    }
    P.numBounces++;
}

}  // namespace gsref

using gsref::Scene;
typedef Scene ref_scene;

#define REF_API extern "C" __attribute__((visibility("default")))

REF_API ref_scene* ref_scene_create(void) {
    Scene* s = new Scene();
    std::memset(&s->rs, 0, sizeof(s->rs));
    s->refresh();
    return s;
}
REF_API void ref_scene_destroy(ref_scene* s) { delete s; }

REF_API uint32_t ref_scene_add_mesh(ref_scene* s, const float* pos3, const float* nrm3, uint32_t nverts) {
    std::vector<glm::vec3> p(nverts), n(nverts);
    for (uint32_t i = 0; i < nverts; i++) {
        p[i] = glm::vec3(pos3[3 * i], pos3[3 * i + 1], pos3[3 * i + 2]);
        n[i] = glm::vec3(nrm3[3 * i], nrm3[3 * i + 1], nrm3[3 * i + 2]);
    }
    s->positions.push_back(std::move(p));
    s->normals.push_back(std::move(n));
    return (uint32_t)s->positions.size() - 1u;
}

// PathTracer.cpp:59-71
REF_API int ref_scene_set_instances(ref_scene* s, const gsb_instance* inst, uint32_t count) {
    s->src.assign(inst, inst + count);
    s->instances.resize(count);
    s->transforms.resize(count);
    for (uint32_t i = 0; i < count; i++) {
        if (inst[i].mesh >= s->positions.size()) return 1;
        glm::mat4 M;
        std::memcpy(&M[0][0], inst[i].transform, sizeof(float) * 16);   // column-major, glm's memory order
        s->transforms[i] = M;
        gsref::rchit::Instance& ri = s->instances[i];
        ri.transformInvT = glm::inverse(glm::transpose(M));
        ri.positionBuffer = gsref::toRef(s->positions[inst[i].mesh].data());
        ri.normalBuffer = gsref::toRef(s->normals[inst[i].mesh].data());
        ri.emission = glm::vec4(inst[i].emission[0], inst[i].emission[1], inst[i].emission[2], 0.0f);
        ri.bsdf = inst[i].bsdf;
        ri.twofaced = inst[i].twofaced;
    }
    s->refresh();
    return 0;
}

REF_API int ref_scene_set_bsdfs(ref_scene* s, uint32_t type, const void* params, uint32_t count) {
    static const size_t sizes[8] = {16, 8, 8, 24, 44, 16, 20, 32};
    if (type >= 8) return 1;   // the reference has eight models (BSDF.inc)
    const unsigned char* p = static_cast<const unsigned char*>(params);
    s->bsdf[type].assign(p, p + sizes[type] * count);
    s->refresh();
    return 0;
}

REF_API int ref_scene_set_lights(ref_scene* s, const gsb_triangle_light* lights, uint32_t count) {
    s->lights.resize(count);
    for (uint32_t i = 0; i < count; i++) {
        for (int k = 0; k < 3; k++)
            s->lights[i].positions[k] = glm::vec4(lights[i].positions[k][0], lights[i].positions[k][1], lights[i].positions[k][2], lights[i].positions[k][3]);
        s->lights[i].emission = glm::vec3(lights[i].radiance[0], lights[i].radiance[1], lights[i].radiance[2]);
    }
    s->refresh();
    return 0;
}

// PathTracer.cpp:88-90: eye = camera position, view = toWorld, fov in radians
REF_API void ref_scene_set_camera(ref_scene* s, const float eye[3], const float to_world[16], float fov) {
    s->rs.camera.eye = glm::vec4(eye[0], eye[1], eye[2], 1.0f);
    std::memcpy(&s->rs.camera.view[0][0], to_world, sizeof(float) * 16);
    s->rs.camera.fov = fov;
}

// ---- unit-level entry points ------------------------------------------------------------------------------------
REF_API uint32_t ref_tea(uint32_t a, uint32_t b) { return gsref::rgen::tea(a, b); }
REF_API uint32_t ref_pcg_hash(uint32_t v) { return gsref::rgen::pcgHash(v); }

// seed as raygen.rgen:37 builds it, then n_draws x randPcg()
REF_API void ref_rng(const uint32_t* idx, const uint32_t* timestamp, uint32_t n, uint32_t n_draws, uint32_t* out) {
    for (uint32_t i = 0; i < n; i++) {
        gsref::rgen::rngState = gsref::rgen::pcgHash(gsref::rgen::tea(idx[i], timestamp[i]));
        out[(size_t)i * (n_draws + 1)] = gsref::rgen::rngState;
        for (uint32_t k = 0; k < n_draws; k++) out[(size_t)i * (n_draws + 1) + 1 + k] = gsref::rgen::randPcg();
    }
}

// randUniform() for a given state word; returns the float and the advanced state
REF_API float ref_rand_uniform(uint32_t state, uint32_t* state_out) {
    gsref::rchit::rngState = state;
    const float u = gsref::rchit::randUniform();
    if (state_out) *state_out = gsref::rchit::rngState;
    return u;
}

REF_API void ref_onb(const float* n3, const float* v3, uint32_t n, float* frame9, float* to_local3, float* to_world3) {
    for (uint32_t i = 0; i < n; i++) {
        const glm::vec3 nn(n3[3 * i], n3[3 * i + 1], n3[3 * i + 2]), v(v3[3 * i], v3[3 * i + 1], v3[3 * i + 2]);
        const gsref::rchit::Onb onb = gsref::rchit::onbCreate(nn);
        const glm::vec3 a = gsref::rchit::onbTransform(onb, v), b = gsref::rchit::onbUntransform(onb, v);
        const float f[9] = {onb.tangent.x, onb.tangent.y, onb.tangent.z, onb.binormal.x, onb.binormal.y, onb.binormal.z,
                            onb.normal.x, onb.normal.y, onb.normal.z};
        std::memcpy(frame9 + 9 * i, f, sizeof(f));
        to_local3[3 * i] = a.x; to_local3[3 * i + 1] = a.y; to_local3[3 * i + 2] = a.z;
        to_world3[3 * i] = b.x; to_world3[3 * i + 1] = b.y; to_world3[3 * i + 2] = b.z;
    }
}

// sampleBSDF(vec2(0), handle, wo, wi, res) through the reference's own switch (rayhit.rchit:630-641)
REF_API int ref_bsdf_sample(ref_scene* s, uint32_t handle, const float* wo, const uint32_t* seed, uint32_t n, float* wi,
                            float* f, float* pdf, uint32_t* is_delta, uint32_t* seed_out) {
    gsref::bind(s, nullptr, nullptr);
    for (uint32_t i = 0; i < n; i++) {
        gsref::rchit::rngState = seed[i];
        gsref::rchit::BSDFOutput res;
        res.bsdf = glm::vec3(0.0f); res.pdf = 0.0f; res.isDelta = false;
        glm::vec3 w(0.0f);
        gsref::rchit::sampleBSDF(glm::vec2(0.0f), handle, glm::vec3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), w, res);
        wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z;
        f[3 * i] = res.bsdf.x; f[3 * i + 1] = res.bsdf.y; f[3 * i + 2] = res.bsdf.z;
        pdf[i] = res.pdf;
        is_delta[i] = res.isDelta ? 1u : 0u;
        seed_out[i] = gsref::rchit::rngState;
    }
    return 0;
}

// evalBSDF(handle, vec2(0), wo, wi, res) (rayhit.rchit:643-654)
REF_API int ref_bsdf_eval(ref_scene* s, uint32_t handle, const float* wo, const float* wi, uint32_t n, float* f, float* pdf,
                          uint32_t* is_delta) {
    gsref::bind(s, nullptr, nullptr);
    for (uint32_t i = 0; i < n; i++) {
        gsref::rchit::BSDFOutput res;
        res.bsdf = glm::vec3(0.0f); res.pdf = 0.0f; res.isDelta = false;
        gsref::rchit::evalBSDF(handle, glm::vec2(0.0f), glm::vec3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]),
                               glm::vec3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), res);
        f[3 * i] = res.bsdf.x; f[3 * i + 1] = res.bsdf.y; f[3 * i + 2] = res.bsdf.z;
        pdf[i] = res.pdf;
        is_delta[i] = res.isDelta ? 1u : 0u;
    }
    return 0;
}

// sampleLight(pos) (rayhit.rchit:147-153)
REF_API int ref_sample_light(ref_scene* s, const float* pos, const uint32_t* seed, uint32_t n, float* lpos, float* emission,
                             float* pdf, uint32_t* seed_out) {
    if (s->lights.empty()) return 1;   // numLights == 0 is a division by zero in the shader
    gsref::bind(s, nullptr, nullptr);
    for (uint32_t i = 0; i < n; i++) {
        gsref::rchit::rngState = seed[i];
        const gsref::rchit::LightOutput r = gsref::rchit::sampleLight(glm::vec3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]));
        lpos[3 * i] = r.position.x; lpos[3 * i + 1] = r.position.y; lpos[3 * i + 2] = r.position.z;
        emission[3 * i] = r.emission.x; emission[3 * i + 1] = r.emission.y; emission[3 * i + 2] = r.emission.z;
        pdf[i] = r.pdf;
        seed_out[i] = gsref::rchit::rngState;
    }
    return 0;
}

// primary rays exactly as raygen.rgen:31-35 builds them
REF_API void ref_camera_rays(ref_scene* s, uint32_t width, uint32_t height, const uint32_t* px, const uint32_t* py, uint32_t n,
                             gsb_ray* out) {
    gsref::bind(s, nullptr, nullptr);
    using namespace gsref;
    for (uint32_t i = 0; i < n; i++) {
        const glm::uvec2 size(width, height), id(px[i], py[i]);
        glm::vec4 origin = s->rs.camera.eye;
        glm::vec4 direction = glm::vec4(rgen::rayDir(glm::vec2(size), glm::vec2(id), s->rs.camera.fov), 0.0f);
        direction = s->rs.camera.view * direction;
        direction.y *= -1;
        out[i].origin[0] = origin.x; out[i].origin[1] = origin.y; out[i].origin[2] = origin.z;
        out[i].dir[0] = direction.x; out[i].dir[1] = direction.y; out[i].dir[2] = direction.z;
        out[i].tmin = 0.0f; out[i].tmax = 1e10f;
    }
}

// ---- whole shader invocations -----------------------------------------------------------------------------------
// raygen.rgen main for timestamps first .. first+n-1 of every pixel, into `image` (RGBA32F, the reference's running
// mean, raygen.rgen:84-108).  The shader's literals apply: MAX_DEPTH 50, roulette after depth 10, cutoff 20, NEE on.
REF_API int ref_render(ref_scene* s, uint32_t width, uint32_t height, uint32_t first, uint32_t n, float* image_rgba,
                       ref_intersect_fn fn, void* user, int threads, uint64_t* ext_rays, uint64_t* shadow_rays) {
    uint64_t ext = 0, sh = 0;
    for (uint32_t k = 0; k < n; k++) {
        s->rs.params.timestamp = first + k;   // PathTracer.cpp:91-92
#pragma omp parallel for schedule(dynamic, 4) num_threads(threads > 0 ? threads : 1) reduction(+ : ext, sh)
        for (int y = 0; y < (int)height; y++) {
            gsref::bind(s, fn, user);
            gsref::P.image.rgba = image_rgba;
            gsref::P.image.width = width;
            gsref::P.bounces = nullptr;
            gsref::P.extRays = 0; gsref::P.shadowRays = 0;
            gsref::P.launchSize = glm::uvec3(width, height, 1u);
            for (uint32_t x = 0; x < width; x++) {
                gsref::P.launchID = glm::uvec3(x, (uint32_t)y, 0u);
                gsref::rgen::shader_main();
            }
            ext += gsref::P.extRays; sh += gsref::P.shadowRays;
        }
    }
    if (ext_rays) *ext_rays = ext;
    if (shadow_rays) *shadow_rays = sh;
    return 0;
}

// raygen.rgen's MAX_DEPTH for the following renders; <= 0 restores the shader's literal.  Returns the value in force.
REF_API int ref_set_max_depth(int depth) {
    gsref::g_maxDepth = depth > 0 ? depth : REF_MAX_DEPTH_LITERAL;
    return gsref::g_maxDepth;
}

// one pixel-sample with per-bounce records; the pixel's value goes to out_rgb (timestamp 0 semantics: the sample itself)
REF_API int ref_path_probe(ref_scene* s, uint32_t width, uint32_t height, uint32_t x, uint32_t y, uint32_t timestamp,
                           float out_rgb[3], ref_bounce* bounces, uint32_t max_bounces, ref_intersect_fn fn, void* user) {
    float px[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    gsref::bind(s, fn, user);
    // a 1-pixel image addressed at (x, y): hand the shader a base pointer such that its store lands in px.  With
    // timestamp > 0 the shader would blend with the previous mean; the probe wants the sample, so the blend input is the
    // sample's own storage zeroed and the result is rescaled below.
    std::vector<float> img((size_t)width * height * 4, 0.0f);
    gsref::P.image.rgba = img.data();
    gsref::P.image.width = width;
    gsref::P.bounces = bounces;
    gsref::P.maxBounces = max_bounces;
    gsref::P.numBounces = 0;
    gsref::P.launchSize = glm::uvec3(width, height, 1u);
    gsref::P.launchID = glm::uvec3(x, y, 0u);
    const uint32_t keep = s->rs.params.timestamp;
    // seed from the requested timestamp; accumulate as frame 0 so that the stored value is the sample itself
    // (raygen.rgen:37 reads params.timestamp for the seed, :86 for the blend) -> run with the real timestamp and undo the blend
    s->rs.params.timestamp = timestamp;
    gsref::rgen::shader_main();
    s->rs.params.timestamp = keep;
    const float* q = img.data() + 4 * ((size_t)y * width + x);
    const float a = timestamp > 0 ? 1.0f / float(timestamp + 1) : 1.0f;   // mix(0, sample, a) = sample * a
    px[0] = q[0] / a; px[1] = q[1] / a; px[2] = q[2] / a;
    out_rgb[0] = px[0]; out_rgb[1] = px[1]; out_rgb[2] = px[2];
    gsref::P.bounces = nullptr;
    return (int)gsref::P.numBounces;
}

// rayhit.rchit main once: payload in (17 words as pt_common.glsl:4-14 orders them), hit record, payload out
REF_API int ref_closest_hit(ref_scene* s, const float* payload_in17, const gsb_hit* hit, float* payload_out17,
                            ref_intersect_fn fn, void* user) {
    gsref::bind(s, fn, user);
    gsref::P.bounces = nullptr;
    gsref::rchit::HitPayload& p = *reinterpret_cast<gsref::rchit::HitPayload*>(gsref::P.payload0);
    static_assert(sizeof(gsref::rchit::HitPayload) == 17 * 4, "HitPayload is 17 words");
    std::memcpy(&p, payload_in17, sizeof(p));
    gsref::P.instanceID = (int)hit->inst;
    gsref::P.primitiveID = (int)hit->prim;
    gsref::P.hitT = hit->t;
    gsref::P.attribs = glm::vec3(hit->u, hit->v, 0.0f);
    gsref::P.worldRayOrigin = p.origin;
    gsref::P.worldRayDirection = p.direction;
    gsref::P.objectToWorld = glm::mat4x3(s->transforms[hit->inst]);
    gsref::rchit::shader_main();
    std::memcpy(payload_out17, &p, sizeof(p));
    return 0;
}

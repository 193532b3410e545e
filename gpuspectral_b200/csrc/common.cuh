// common.cuh — shared device/host definitions of the CUDA core: error handling, float3
// helpers with GLSL semantics, the reference's RNG, device-side scene records.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <string>

#include "../../include/gpuspectral_b200.h"

namespace gsb {

// ---- error plumbing ---------------------------------------------------------------------
void setLastError(const std::string& msg);
struct CudaFailure { cudaError_t err; const char* what; const char* file; int line; };

#define GSB_CUDA(call)                                                         \
    do {                                                                       \
        cudaError_t _e = (call);                                               \
        if (_e != cudaSuccess) throw gsb::CudaFailure{_e, #call, __FILE__, __LINE__}; \
    } while (0)

// ---- float3 helpers (GLSL semantics) -----------------------------------------------------
#define GSB_HD __host__ __device__ __forceinline__
#define GSB_D __device__ __forceinline__

GSB_HD float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
GSB_HD float3 f3(float s) { return make_float3(s, s, s); }
GSB_HD float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
GSB_HD float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
GSB_HD float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
GSB_HD float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
GSB_HD float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
GSB_HD float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
// one IEEE reciprocal + three products (<= 1 ulp from three divisions; GLSL does not promise correctly rounded division)
GSB_HD float3 operator/(float3 a, float s) { const float r = 1.0f / s; return f3(a.x * r, a.y * r, a.z * r); }
GSB_HD float3 operator/(float3 a, float3 b) { return f3(a.x / b.x, a.y / b.y, a.z / b.z); }
GSB_HD float3 operator+(float3 a, float s) { return f3(a.x + s, a.y + s, a.z + s); }
GSB_HD float3 operator-(float3 a, float s) { return f3(a.x - s, a.y - s, a.z - s); }
GSB_HD float3 operator-(float s, float3 a) { return f3(s - a.x, s - a.y, s - a.z); }
GSB_HD float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GSB_HD float3 cross(float3 a, float3 b) { return f3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
GSB_HD float length(float3 a) { return sqrtf(dot(a, a)); }
// GLSL leaves normalize() implementation-defined; one reciprocal and three products (the checker divides three times:
// results differ by at most 1 ulp per component, inside every stated tolerance) instead of three IEEE divisions
GSB_HD float3 normalize(float3 a) { return a * (1.0f / length(a)); }
GSB_HD float3 vsqrt(float3 a) { return f3(sqrtf(a.x), sqrtf(a.y), sqrtf(a.z)); }
// faceforward(N, I, Nref) = dot(Nref, I) < 0 ? N : -N
GSB_HD float3 faceforward(float3 N, float3 I, float3 Nref) { return dot(Nref, I) < 0.0f ? N : -N; }
GSB_HD float3 xyz(float4 v) { return f3(v.x, v.y, v.z); }

// "#define M_PI 3.14159265358979323846" is a 32-bit float literal inside the GLSL shaders.
#define GSB_PI 3.14159265358979323846f

// ---- RNG: pt_common.glsl:85-120 ------------------------------------------------------------
GSB_HD uint32_t pcgOutput(uint32_t state) {
    uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
    return (word >> 22u) ^ word;
}
// randPcg(): returns the hash of the pre-advance state and steps the LCG
GSB_HD uint32_t randPcg(uint32_t& rngState) {
    uint32_t state = rngState;
    rngState = rngState * 747796405u + 2891336453u;
    return pcgOutput(state);
}
GSB_HD uint32_t pcgHash(uint32_t v) { return pcgOutput(v * 747796405u + 2891336453u); }
// randUniform(): uint -> float (round to nearest even) times 2^-32, so u is in [0, 1] inclusive
GSB_HD float randUniform(uint32_t& rngState) { return (float)randPcg(rngState) * 2.3283064365386963e-10f; }
GSB_HD uint32_t tea(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
#pragma unroll
    for (uint32_t n = 0; n < 4; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

// ---- device-side scene records ---------------------------------------------------------------
// Per-instance shading record (128 B, eight 16-byte words), indexed by gl_InstanceID.
// The reference's Instance (pt_common.glsl:44-51, 112-B stride) with the object->world rows it
// gets from gl_ObjectToWorldEXT folded in and the two buffer addresses kept as 64-bit pointers.
struct ShadeInstance {
    float4 o2w[3];       // rows of the 3x4 object->world matrix
    float4 nrm[3];       // rows of inverse(transpose(M))'s 3x3; .w = emission r,g,b
    const float4* vpos;  // 3 float4 per triangle, object space, w unused
    const float4* vnrm;  // 3 float4 per triangle
    uint32_t bsdf;       // type<<16 | index
    uint32_t twofaced;
    uint32_t pad[2];
};
static_assert(sizeof(ShadeInstance) == 128, "ShadeInstance must stay 128 bytes");

// Traversal-side instance record (64 B), stored in TLAS leaf order.
struct TraceInstance {
    float4 w2o[3];          // rows of the 3x4 world->object matrix
    const uint4* blas;      // the mesh's bottom-level buffer: [nodes | triangles]
    uint32_t triOffset16;   // offset of the triangle array inside `blas`, in 16-byte units
    uint32_t instId;        // gl_InstanceID
};
static_assert(sizeof(TraceInstance) == 64, "TraceInstance must stay 64 bytes");

// The instance word a hit carries (WavefrontState::hitInst, world-level triangle records, TraceInstance::instId):
// bits 0..26 gl_InstanceID, bit 27 the instance is baked (its triangles sit at the world level), bits 28..31 the BSDF type
// of the instance — k_shade<TYPE> finds its work from this word alone.  GSB_NO_HIT (all ones) decodes to no valid type.
constexpr uint32_t kInstMask = 0x07ffffffu;
constexpr uint32_t kInstBaked = 1u << 27;
constexpr uint32_t kInstTypeShift = 28;
constexpr uint32_t kMaxInstances = kInstMask;   // gsb_scene_set_instances rejects more
constexpr uint32_t kNoLeaf = 0xffffffffu;

// Device-side texture record (gsb_texture with the texels on the device, one float4 per texel).
struct TextureDev {
    uint32_t type, width, height, pad;
    float uscale, vscale, uoffset, voffset;
    float4 color0, color1;
    const float4* texels;
    uint64_t pad2;
};
static_assert(sizeof(TextureDev) == 80, "TextureDev layout");

// Environment emitter on the device: texels (float4) + the sampling tables built by buildEnvironmentTables (api.cu):
// rowCdf[h + 1] over rows (weight = sum of the row's luminances x sin(theta) at the row centre), colCdf[h][w + 1] per row.
struct EnvDev {
    uint32_t type, width, height, pad;
    float scale[3];
    float distance;          // how far away a sampled point of the environment is placed (beyond the scene)
    float rot[9];            // local -> world
    float pad2[3];
    const float4* texels;
    const float* rowCdf;
    const float* colCdf;
    uint64_t pad3;
};

// Everything the kernels need to know about the scene; passed by value as a kernel parameter.
struct SceneDev {
    const uint4* worldNodes;         // world level: 8-wide nodes over the triangles of all baked instances (world space)
    const float4* worldTris;         //   their leaf records (3 rotated copies of 3 float4; v0.w = primitive, v1.w = instance word)
    const float4* worldNrm;          //   their vertex normals in world space (transformInvT * n), 3 float4 per triangle, leaf order
    const float4* instSmall;         // per instance: emission.xyz, bits(bsdf index | twofaced << 16) — all a baked hit needs of its instance
    const uint4* tlasNodes;          // 8-wide nodes of the top level over the instances that are not baked
    const TraceInstance* tlasInst;   // instance records in TLAS leaf order
    const ShadeInstance* shadeInst;  // instance records in gl_InstanceID order
    const uint32_t* instBsdf;        // bsdf handle per instance (compact copy for queue binning)
    const void* bsdf[GSB_BSDF_TYPE_COUNT];
    const gsb_triangle_light* lights;
    const TextureDev* textures;      // reflectance textures (diffuse records refer to them by hasTexture - 1)
    EnvDev env;                      // environment emitter (type 0 = none)
    uint32_t numLights;
    uint32_t numInstances;
    uint32_t numTextures;
    // camera: Camera of pt_common.glsl:16-21
    float eye[3];
    float view[16];  // column-major toWorld
    float fov;
    gsb_params params;
    // ray sorting: origin cell = (o - sortLo) * sortScale per axis, clamped to [0, 2^sortBits); sortBits = 0: no sorting
    float sortLo[3], sortScale[3];
    uint32_t sortBits;
    uint32_t sortMode;   // 0: cell, then octant; 1: octant, then cell; 2: cell, then 2 bits of direction per axis (A/B knob GSB_SORT_MODE)
};

}  // namespace gsb

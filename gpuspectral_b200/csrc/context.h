// context.h — the opaque gsb_context behind the C ABI (internal).
#pragma once

#include <string>
#include <vector>

#include "bvh.h"
#include "common.cuh"

namespace gsb {

struct MeshRec {
    bool alive = false;
    uint32_t ntris = 0;
    float4* vpos = nullptr;  // 3 float4 per triangle, original primitive order
    float4* vnrm = nullptr;
    DeviceBvh blas;
};

// Counter indices inside WavefrontState::qCount (two banks of Q_COUNT, selected by iteration parity):
// [0..TYPE_COUNT) unused (material sorting happens inside k_shade), extension rays traced, the shadow-ray queue.
enum { Q_MATERIAL0 = 0, Q_TRACED = GSB_BSDF_TYPE_COUNT, Q_SHADOW = GSB_BSDF_TYPE_COUNT + 1,
       Q_NEXT_EXTEND = GSB_BSDF_TYPE_COUNT + 2, Q_NEXT_SHADOW = GSB_BSDF_TYPE_COUNT + 3,  // work counters of the persistent traversal kernels
       Q_COUNT = GSB_BSDF_TYPE_COUNT + 4 };

// Per-slot path state, structure of arrays.  The 17-word HitPayload of pt_common.glsl:4-14
// (+ pixel sample index and the hit record) spread over 16-byte and 4-byte streams.
struct WavefrontState {
    uint32_t capacity = 0;      // slots allocated
    // the path's ray, 2 float4 per slot: origin.xyz + RNG state of the path (payload `seed`) in .w, direction.xyz + path flags
    // in .w (depth bits 0-15, wasDelta, countEmitted, dead).  One 32-byte record per slot (wavefront.cu: rayOrigin /
    // rayDirection), so that the scattered accesses — paths started by k_regen, rays gathered in sorted order — cost one
    // sector instead of two; GSB_RAY_AOS=0 builds the two-stream layout for comparison
    float4* ray = nullptr;
    float4* weightDW = nullptr; // weight.xyz, directWeight
    float4* result = nullptr;   // radiance of the current sample .xyz (only touched by vertices that add light); .w = bits of the pixel
                                // index the sample belongs to, 0xffffffff = slot retired (no work left)
    float4* hitTUVP = nullptr;    // t, u, v, primitive id bits
    uint32_t* hitInst = nullptr;
    float4* shadowO = nullptr;      // origin.xyz, tmax
    float4* shadowD = nullptr;      // direction.xyz, slot id bits
    float4* shadowBase = nullptr;   // emitted without NEE .xyz, directWeight if visible
    float4* shadowNee = nullptr;    // NEE contribution .xyz
    // ray sorting (optional): key per slot (written with the ray by k_regen / k_shade), the sorted keys, the slot order
    // k_extend follows, the identity permutation that is the sort's value input, cub's scratch space
    uint32_t* sortKey = nullptr;
    uint32_t* sortKeyOut = nullptr;
    uint32_t* order = nullptr;
    uint32_t* iota = nullptr;
    void* sortTemp = nullptr;
    size_t sortTempBytes = 0;
    uint32_t* qCount = nullptr;     // 2 x Q_COUNT queue counters
    // device statistics block: [0] live slots, [2..] totals
    unsigned long long* stats = nullptr;
};

enum { ST_LIVE = 0, ST_RETIRED = 1, ST_PATHS = 2, ST_EXT = 3, ST_SHADOW = 4, ST_ITERS = 5, ST_NODES = 6, ST_TRIS = 7, ST_INSTS = 8, ST_RAYS = 9, ST_OVERFLOW = 10, ST_NEXTWORK = 11, ST_BATCHNEXT = 12, ST_LOOPITERS = 13, ST_VERTICES = 14, ST_COUNT = 16 };

// A render call that was launched as a device-side loop and has not been waited for yet.
struct PendingRender {
    cudaGraphExec_t exec = nullptr;
    cudaGraph_t graph = nullptr;
    cudaEvent_t start = nullptr, stop = nullptr;
    uint32_t launchesPerIter = 0;
};

}  // namespace gsb

struct gsb_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    std::vector<gsb::MeshRec> meshes;
    std::vector<gsb_instance> instances;
    gsb::DeviceBvh world;              // world level: BVH over the triangles of all baked instances, in world space
    float4* dWorldVerts = nullptr;     //   their vertices (3 float4 per triangle; v0.w = primitive id, v1.w = instance id)
    float4* dWorldNrm = nullptr;       //   their vertex normals in world space, leaf order (shading reads them straight from here)
    float4* dInstSmall = nullptr;      // per instance: emission, bsdf index, twofaced (what a baked hit needs of its instance)
    uint32_t numBakedTriangles = 0;
    float4* dNrmSoup = nullptr;        //   the same normals in bake order (kept: a moved baked instance is re-baked in place)
    std::vector<uint64_t> bakeOffset;  // per instance: first world-level triangle (bake order), ~0 = not baked
    bool voteFromEnv = false, triMinFromEnv = false;   // GSB_VOTE / GSB_TRIMIN given: choosePolicy leaves those bits alone
    int worldSetting = -1;             // treelet setting that won the world level's build comparison (-1: not compared yet)
    uint64_t worldSettingTris = ~0ull; // ... for a world level of this many triangles
    std::vector<int32_t> traceIndex;   // per instance: its position in hostTrace, -1 = baked
    std::vector<gsb::TraceInstance> hostTrace;   // trace records and padded world bounds of the TLAS instances (refit input)
    std::vector<float> hostWlo, hostWhi;
    double lastRefitMs = 0.0;
    uint64_t refits = 0;
    gsb::DeviceBvh tlas;               // top level over the instances that are not baked
    uint32_t numTraceInstances = 0;
    gsb::ShadeInstance* dShade = nullptr;
    uint32_t* dInstBsdf = nullptr;
    void* dBsdf[GSB_BSDF_TYPE_COUNT] = {};
    uint32_t bsdfCount[GSB_BSDF_TYPE_COUNT] = {};
    bool typeUsed[GSB_BSDF_TYPE_COUNT] = {};
    gsb_triangle_light* dLights = nullptr;
    uint32_t numLights = 0;
    gsb::EnvDev env = {};                  // environment emitter; its device arrays:
    float4* dEnvTexels = nullptr;
    float* dEnvRowCdf = nullptr;
    float* dEnvColCdf = nullptr;
    gsb::TextureDev* dTextures = nullptr;
    std::vector<float4*> textureTexels;   // device texel arrays owned by the table
    uint32_t numTextures = 0;
    float eye[3] = {0, 0, 0};
    float view[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float fov = 1.0f;
    gsb_params params = {50, 10, 20.0f, 1, 0};
    bool haveInstances = false;
    double lastBuildMs = 0.0;

    uint32_t width = 0, height = 0;
    float4* dAccum = nullptr;  // rgb = sum of samples, a = samples added
    gsb::WavefrontState wf;
    bool profiling = false;
    uint32_t targetSlots = 1u << 24;  // paths in flight per wavefront (GSB_SLOTS overrides); 16 M x ~210 B = 3.5 GB of state.
                                      // Measured (r2f, four launches per iteration): 4 M -> 8 M +4 %, 8 M -> 16 M +1.5 % (Cornell) .. +3.5 % (gallery, field)
    // control word of the persistent traversal kernels (traverse.cuh: traceStream): idle lanes that trigger a refill
    // (bits 0..7, GSB_REFILL), lanes that make a quorum for a traversal phase (8..15, GSB_VOTE; closest-hit kernel only,
    // the any-hit kernel measured best without a quorum), lanes that keep the triangle loop spinning (16..23, GSB_TRIMIN)
    uint32_t refillMin = 24u | (4u << 8) | (4u << 16);
    uint32_t directRefill = 0u;    // idle lanes that trigger a refill straight from the source (GSB_REFILL); 0 = per-scene default (api.cu: choosePolicy)
    uint32_t shadowRefill = 0u;    // the same for the any-hit kernel (GSB_SHADOW_REFILL)
    uint32_t shadowControl = 16u | (4u << 16);  // the any-hit kernel's control word (api.cu: choosePolicy)
    uint32_t sortBits = 0;         // ray sorting: grid bits per axis of the origin cell in the sort key; 0 = rays are traced in slot order
    bool poolExtend = false, poolShadow = false;   // traverse.cuh: ray pool (true) or direct refill; chosen per scene in gsb_scene_set_instances, GSB_POOL overrides
    uint32_t shadeDirectMin = 32;  // k_shade: matching lanes from which a tile is shaded in place (GSB_SHADE_DIRECT; measured flat from 12 to 32)
    bool useGraphLoop = true;    // device-side render loop (conditional CUDA graph); GSB_GRAPH=0 or a failed build disables it
    int instrumented = 0;        // bit 0: count fetches of extension rays, bit 1: of shadow rays
    bool ownStream = true;       // false after gsb_set_stream
    uint64_t launches = 0;       // kernels launched by gsb_render since gsb_reset_counters
    uint64_t lastRenderIters = 0;
    double lastRenderMs[5] = {0, 0, 0, 0, 0};
    std::vector<gsb::PendingRender> pending;   // asynchronous render calls in flight on `stream`
    unsigned long long seenIters = 0;          // value of the device iteration counter already accounted for

    gsb::SceneDev sceneDev() const;
};

namespace gsb {
// wavefront.cu
void renderImpl(gsb_context* ctx, uint32_t width, uint32_t height, uint32_t first, uint32_t count, uint32_t stride);
void freeWavefront(gsb_context* ctx);
// Waits for the render calls in flight, releases their graphs and folds their iteration counts into the context.
void finishPending(gsb_context* ctx);
}  // namespace gsb

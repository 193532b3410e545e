// traverse.cuh — software ray traversal of the two-level, 8-wide compressed BVH.
//
// Replaces what the reference gets from the Vulkan driver / RT cores behind
//   traceRayEXT(topLevelAS, gl_RayFlagsOpaqueEXT, 0xff, ...)                 raygen.rgen:54-58
//   traceRayEXT(topLevelAS, TerminateOnFirstHit|Opaque|SkipClosestHit, ...)  rayhit.rchit:737-748
// over the TLAS/BLAS of backend/vulkan/VulkanRays.cpp:6-181 (opaque triangles, no culling).
//
// Node format (80 B = 5 x 16 B, fetched as five 128-bit loads), after Ylitie, Karras & Laine,
// "Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide BVHs" (HPG 2017):
//   w0: origin.x, origin.y, origin.z (f32), {ex, ey, ez, imask} (u8 x4)
//   w1: childBase (u32), leafBase (u32), meta[8] (u8 x8)
//   w2: qlo.x[8], qlo.y[8]      w3: qlo.z[8], qhi.x[8]      w4: qhi.y[8], qhi.z[8]
// child i box = origin + q * 2^(e - 127).  meta[i] = 0: empty; internal child: 0x20 | (24 + i);
// leaf child: (unary count 1/3/7) << 5 | offset of its first primitive inside the node's leaf range.
// Bottom level primitive = triangle (3 x float4: v0.xyz|primId, v1.xyz, v2.xyz), 48 B fetched per test; stored three
// times (144 B), once per dominant ray axis, pre-rotated so that the test needs no per-vertex selects.
// Top level primitive = TraceInstance (64 B).
//
// Intersection rule (shared with the checker; the reference has no source for it):
//   world ray -> object space with the instance's inverse matrix (direction not renormalised),
//   watertight test of Woop, Benthin & Wald (JCGT 2013) in float32 with individually rounded
//   operations, candidates need tmin < t < tmax, closest hit = lexicographic minimum of
//   (t, instance, primitive).  Box tests are conservative (padded boxes, t_near <= t_far).
#pragma once

#include "common.cuh"

namespace gsb {

constexpr int kStackSize = 48;      // total traversal stack entries per ray (world level, then TLAS + BLAS)
constexpr int kWorldStackSize = 32; // ... when the scene has the world level only (at most 2^22 baked triangles)
#ifndef GSB_SMEM_STACK
#define GSB_SMEM_STACK 8
#endif
constexpr int kSmemStack = GSB_SMEM_STACK;   // of which the first kSmemStack live in shared memory (the hot bottom of the stack)
constexpr int kTraceBlock = 128;    // every kernel that calls traceRay runs with this block size
constexpr uint32_t kDirectShift = 29;  // TraceInstance::triOffset16 bits 29..31: triangle count of a node-less mesh

struct TraceStats { uint32_t nodes, tris, insts; };

struct HitRecord {
    float t, u, v;
    uint32_t inst, prim;   // inst: the instance word (common.cuh: kInstMask / kInstBaked / type), GSB_NO_HIT on a miss
    uint32_t leaf;         // world-level hits: position of the triangle in the world level's leaf order; kNoLeaf otherwise
};

// object-space ray with the Woop shear constants
struct ObjRay {
    float ox, oy, oz;     // origin, rotated so that z is the dominant axis
    float Sx, Sy, Sz;
    uint32_t perm;        // kz
};

// Cyclic rotation that brings axis kz to z.  The checker additionally swaps x and y when d[kz] < 0 (to keep the
// winding); that swap negates U, V, W, det and T together (round-to-nearest is symmetric in sign), so the accept /
// reject decisions and the quotients t = T/det, u = V/det, v = W/det keep their bits without it — it is left out.
GSB_D void permute(uint32_t kz, float x, float y, float z, float& px, float& py, float& pz) {
    pz = kz == 0 ? x : (kz == 1 ? y : z);
    px = kz == 0 ? y : (kz == 1 ? z : x);
    py = kz == 0 ? z : (kz == 1 ? x : y);
}

GSB_D void makeObjRay(float ox, float oy, float oz, float dx, float dy, float dz, ObjRay& r) {
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    const uint32_t kz = (ax > ay) ? ((ax > az) ? 0u : 2u) : ((ay > az) ? 1u : 2u);
    r.perm = kz;
    float pdx, pdy, pdz;
    permute(r.perm, dx, dy, dz, pdx, pdy, pdz);
    r.Sx = __fdiv_rn(pdx, pdz);
    r.Sy = __fdiv_rn(pdy, pdz);
    r.Sz = __fdiv_rn(1.0f, pdz);
    permute(r.perm, ox, oy, oz, r.ox, r.oy, r.oz);
}

// Watertight ray/triangle test; every operation individually rounded (no FMA contraction).
// Returns false when the ray misses; otherwise the scaled distance T, the barycentric numerators V, W and det,
// so that callers pay for the divisions t = T/det, u = V/det, v = W/det only when they need them.
// v0, v1, v2: the triangle copy already rotated for the ray's dominant axis (bvh_build.cu: k_packTriangles).
GSB_D bool woopEdges(const ObjRay& r, float4 v0, float4 v1, float4 v2, float& T, float& V, float& W, float& det) {
    float Akx = v0.x, Aky = v0.y, Akz = v0.z, Bkx = v1.x, Bky = v1.y, Bkz = v1.z, Ckx = v2.x, Cky = v2.y, Ckz = v2.z;
    Akx = __fsub_rn(Akx, r.ox); Aky = __fsub_rn(Aky, r.oy); Akz = __fsub_rn(Akz, r.oz);
    Bkx = __fsub_rn(Bkx, r.ox); Bky = __fsub_rn(Bky, r.oy); Bkz = __fsub_rn(Bkz, r.oz);
    Ckx = __fsub_rn(Ckx, r.ox); Cky = __fsub_rn(Cky, r.oy); Ckz = __fsub_rn(Ckz, r.oz);
    const float Ax = __fsub_rn(Akx, __fmul_rn(r.Sx, Akz)), Ay = __fsub_rn(Aky, __fmul_rn(r.Sy, Akz));
    const float Bx = __fsub_rn(Bkx, __fmul_rn(r.Sx, Bkz)), By = __fsub_rn(Bky, __fmul_rn(r.Sy, Bkz));
    const float Cx = __fsub_rn(Ckx, __fmul_rn(r.Sx, Ckz)), Cy = __fsub_rn(Cky, __fmul_rn(r.Sy, Ckz));
    float U = __fsub_rn(__fmul_rn(Cx, By), __fmul_rn(Cy, Bx));
    V = __fsub_rn(__fmul_rn(Ax, Cy), __fmul_rn(Ay, Cx));
    W = __fsub_rn(__fmul_rn(Bx, Ay), __fmul_rn(By, Ax));
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        // products of two floats are exact in double; one rounding in the subtraction, one to float
        U = (float)__dsub_rn(__dmul_rn((double)Cx, (double)By), __dmul_rn((double)Cy, (double)Bx));
        V = (float)__dsub_rn(__dmul_rn((double)Ax, (double)Cy), __dmul_rn((double)Ay, (double)Cx));
        W = (float)__dsub_rn(__dmul_rn((double)Bx, (double)Ay), __dmul_rn((double)By, (double)Ax));
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    det = __fadd_rn(__fadd_rn(U, V), W);
    if (det == 0.0f) return false;
    const float Az = __fmul_rn(r.Sz, Akz), Bz = __fmul_rn(r.Sz, Bkz), Cz = __fmul_rn(r.Sz, Ckz);
    T = __fadd_rn(__fadd_rn(__fmul_rn(U, Az), __fmul_rn(V, Bz)), __fmul_rn(W, Cz));
    return true;
}

// The full test with all three quotients (diagnostics; the checker's definition of a hit).
GSB_D bool woopTest(const ObjRay& r, float4 v0, float4 v1, float4 v2, float& t, float& bu, float& bv) {
    float T, V, W, det;
    if (!woopEdges(r, v0, v1, v2, T, V, W, det)) return false;
    t = __fdiv_rn(T, det);
    bu = __fdiv_rn(V, det);
    bv = __fdiv_rn(W, det);
    return true;
}

// 1/d for the slab tests only (never for the triangle test): one MUFU.RCP.  |d| is clamped to 2^-80 as in the
// CWBVH paper so the reciprocal stays finite; its 1-ulp error is far inside the box padding (bvh.h: kBoxPadRel).
GSB_D float slabInv(float d) {
    const float eps = 8.271806125530277e-25f;
    const float c = fabsf(d) > eps ? d : copysignf(eps, d);
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(c));
    return r;
}

// plain approximate reciprocal (no clamping): inf / nan in, inf / nan out — callers treat those as "not clear"
GSB_D float slabInvRaw(float d) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(d));
    return r;
}

GSB_D uint32_t byteOf(uint32_t w, int j) { return (w >> (8 * j)) & 0xffu; }

// Intersects the 8 quantised child boxes of one node, branch-free.  Returns the node group
// (childBase, internal hit bits 24..31 | imask) and the leaf group (leafBase, hit bits 0..23).
// Hit-bit position of a child comes from its meta byte, four children at a time (CWBVH paper, listing 2):
// internal child in slot i has meta 0x20 | (24 + i) and lands on bit 24 + (i ^ oct7) so that a descending bit
// scan visits children front to back; a leaf child has meta (unary count << 5) | first-primitive offset.
GSB_D void intersectNode(const uint4* __restrict__ nodes, uint32_t nodeIndex, float ox, float oy, float oz, float idx,
                         float idy, float idz, float tmin, float tmax, uint32_t oct7, uint2& nodeGroup, uint2& leafGroup) {
    const uint4* n = nodes + (size_t)nodeIndex * 5;
    const uint4 w0 = __ldg(n + 0);
    const uint4 w1 = __ldg(n + 1);
    const uint4 w2 = __ldg(n + 2);
    const uint4 w3 = __ldg(n + 3);
    const uint4 w4 = __ldg(n + 4);
    const uint32_t e = w0.w;
    const uint32_t imask = e >> 24;
    // box = origin + q * scale  ->  t = q * (scale * idir) + (origin - o) * idir
    const float sx = __uint_as_float((e & 0xffu) << 23) * idx;
    const float sy = __uint_as_float(((e >> 8) & 0xffu) << 23) * idy;
    const float sz = __uint_as_float(((e >> 16) & 0xffu) << 23) * idz;
    const float bx = (__uint_as_float(w0.x) - ox) * idx;
    const float by = (__uint_as_float(w0.y) - oy) * idy;
    const float bz = (__uint_as_float(w0.z) - oz) * idz;
    const bool nx = idx < 0.0f, ny = idy < 0.0f, nz = idz < 0.0f;
    const uint32_t oct4 = oct7 * 0x01010101u;
    uint32_t hits = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = half ? w1.w : w1.z;
        const uint32_t isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;   // index field 24..31
        const uint32_t innerMask4 = (isInner4 >> 4) * 0xffu;
        const uint32_t bitIndex4 = (meta4 ^ (oct4 & innerMask4)) & 0x1f1f1f1fu;
        const uint32_t childBits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlox = half ? w2.y : w2.x, qloy = half ? w2.w : w2.z, qloz = half ? w3.y : w3.x;
        const uint32_t qhix = half ? w3.w : w3.z, qhiy = half ? w4.y : w4.x, qhiz = half ? w4.w : w4.z;
        const uint32_t nearx = nx ? qhix : qlox, farx = nx ? qlox : qhix;
        const uint32_t neary = ny ? qhiy : qloy, fary = ny ? qloy : qhiy;
        const uint32_t nearz = nz ? qhiz : qloz, farz = nz ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float t0x = fmaf((float)byteOf(nearx, j), sx, bx), t1x = fmaf((float)byteOf(farx, j), sx, bx);
            const float t0y = fmaf((float)byteOf(neary, j), sy, by), t1y = fmaf((float)byteOf(fary, j), sy, by);
            const float t0z = fmaf((float)byteOf(nearz, j), sz, bz), t1z = fmaf((float)byteOf(farz, j), sz, bz);
            const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            const uint32_t bits = byteOf(childBits4, j) << byteOf(bitIndex4, j);  // 0 for an empty slot
            hits |= (tn <= tf) ? bits : 0u;
        }
    }
    nodeGroup = make_uint2(w1.x, (hits & 0xff000000u) | imask);
    leafGroup = make_uint2(w1.y, hits & 0x00ffffffu);
}

// One ray as the traversal loop sees it.  `tag` is the caller's handle (slot / queue index / batch index).
struct RayJob {
    float ox, oy, oz, dx, dy, dz, tmin, tmax;
    uint32_t tag;
    uint32_t aux;   // a word of the source's own that travels with the ray and comes back in commit (extend: pixel, shadow: slot word)
};
constexpr uint32_t kNoTag = 0xffffffffu;

// Warp-private range of work indices [cur, end) carved out of a global counter in chunks, so that a refill costs
// no atomic most of the time and consecutive lanes get consecutive indices (coalesced state loads).  The chunk after
// the current one is reserved ahead of time (`next`), so the counter's round trip is off the critical path.
struct WorkRange {
    uint32_t cur = 0, end = 0;
    uint32_t next = 0;         // base of the chunk reserved ahead; valid when haveNext
    bool haveNext = false;
    bool exhausted = false;    // the counter has passed `total`: nothing beyond [cur, end) will come
#ifndef GSB_WORK_CHUNK
#define GSB_WORK_CHUNK 64
#endif
    static constexpr uint32_t kChunk = GSB_WORK_CHUNK;

    GSB_D void reserve(uint32_t* counter, uint32_t total) {
        if (haveNext || exhausted) return;
        uint32_t base = 0;
        if ((threadIdx.x & 31u) == 0) base = atomicAdd(counter, kChunk);
        next = __shfl_sync(0xffffffffu, base, 0);
        haveNext = true;
    }
    // Every lane of the warp calls this; lanes with `want` get distinct indices < total (or kNoTag when work ran out).
    GSB_D uint32_t take(uint32_t* counter, uint32_t total, bool want) {
        const uint32_t mask = __ballot_sync(0xffffffffu, want);
        const uint32_t n = (uint32_t)__popc(mask);
        const uint32_t lane = threadIdx.x & 31u;
        const uint32_t rank = (uint32_t)__popc(mask & ((1u << lane) - 1u));
        const uint32_t avail = end - cur;
        uint32_t idx;
        if (n <= avail) {
            idx = cur + rank;
            cur += n;
        } else {
            reserve(counter, total);
            uint32_t base = total;
            if (haveNext) { base = next; haveNext = false; }
            if (base >= total) exhausted = true;     // the counter only grows: nothing later can be valid
            idx = rank < avail ? cur + rank : base + (rank - avail);
            cur = base + (n - avail);
            end = base + kChunk;
            if (exhausted) { cur = 0; end = 0; }
        }
        if (cur >= total) { cur = 0; end = 0; exhausted = true; }
        reserve(counter, total);                     // issued now, consumed at a later refill
        return (want && idx < total) ? idx : kNoTag;
    }
    // Index the lane would get at the next refill if all lanes asked (for prefetching); kNoTag when unknown.
    GSB_D uint32_t peek(uint32_t total) const {
        const uint32_t i = cur + (threadIdx.x & 31u);
        return (i < end && i < total) ? i : kNoTag;
    }
};

GSB_D void prefetchL1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
GSB_D void prefetchL2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Traces a stream of rays with one persistent warp: the traversal loop never waits for the slowest ray of a
// batch — a lane whose ray is finished hands its result to `src.commit` and takes the next ray at the top of the next
// pass (dynamic ray fetch, Aila & Laine 2009), so all 32 lanes stay in flight until the stream runs dry.
//
// Source concept (all members are called by all 32 lanes together, warp-synchronously):
//   void commit(bool have, const HitRecord& hit, uint32_t tag, uint32_t aux)   result of a finished ray (have = lane has one)
//   bool fetch(bool want, RayJob& job)                           next ray for the lanes that want one; false = lane got none
//   bool drained()                                               warp-uniform: no more rays will ever come
//
// Ray pool.  Taking rays from the source is expensive (state loads from HBM, for k_extend also the end of the previous
// path and the start of the next one), so it is never done for a few idle lanes: the warp fetches 32 rays at a time with
// all lanes (coalesced loads, full lanes for the per-ray set-up: reciprocal direction, octant, the triangle test's shear
// constants) into a ring of kPoolSize prepared rays in shared memory, and a lane whose ray has ended picks its next one
// out of that ring at the top of the very next pass — four 128-bit shared-memory loads.  Round 1 refilled straight from
// the source once 24 lanes were idle, which left 20 of 32 lanes active on the Cornell box (r1k / r2b ncu tables).
//
// Levels.  Every ray first walks the WORLD level: one BVH over the triangles of all baked instances, stored in world
// space (api.cu: the bake rule; v0.w = gl_PrimitiveID, v1.w = gl_InstanceID per triangle), which the ray meets as it
// is — no transform, no instance record.  HASINST = false (every instance of the scene is baked: the Cornell box, the
// BSDF gallery): the ray is done when that stack drains, and none of the two-level state below exists in the kernel.
// HASINST = true: the ray then walks the TLAS over the remaining (large, instanced) meshes; on an instance leaf it is
// taken to object space (unnormalised direction, so t is shared between levels) and walks that mesh's BLAS; a stack mark
// returns it to the top level.
//
// Pass shape: the lanes of a persistent warp drift apart (different rays, different depths), so at any moment some
// want to (1) descend one node, some to (2a) enter an instance at the top level, some to (2b) test pending triangles,
// and the rest have nothing pending and (3) pop.  Running every phase in every pass under its
// own predicate would issue each phase for a handful of lanes.  Instead the warp votes: a phase runs in this pass only
// if at least `vote` lanes are waiting for it (if no phase has a quorum, the one with the most lanes runs, so every
// pass makes progress); a lane whose phase sits the pass out simply keeps its state.  Lanes that arrive at a phase
// while it runs (phase 1 hands triangles to 2b, 2a hands direct triangles to 2b) ride along.  Per-lane operation
// order is unchanged, so results and fetch counts do not depend on the vote threshold.
// `control` packs vote (bits 8..15) and triMin (16..23: lanes needed to keep the triangle loop of
// phase 2b spinning after its first round).  ANYHIT: a ray ends at the first triangle with tmin < t < tmax
// (hit.inst = 0).  The first kSmemStack stack entries live in shared memory ([entry][thread]: conflict-free 64-bit
// accesses); deeper entries spill to local memory.  Closest hit: u and v are kept as numerators (hit.u = V,
// hit.v = W, with `hitDet`) while the ray is in flight and divided once when it is handed in.
#ifndef GSB_POOL_SIZE
#define GSB_POOL_SIZE 48
#endif
#ifndef GSB_POOL_SIZE_INST
#define GSB_POOL_SIZE_INST 48   // two-level scenes carry one more row per entry: 48 entries keep 8 blocks per SM inside 227 KB
#endif

// POOL = false is round 1's scheme without the ring: rays go from the source straight into the lanes once `refillMin`
// (control bits 0..7) of them are idle.  It wins on scenes whose rays end after a handful of passes (the Cornell box:
// 1.6 nodes and 4.5 triangles per ray), where the pool's extra copy through shared memory costs more than the idle
// lanes; the pool wins as soon as rays live longer (gallery: -18 % k_extend time; numbers in profiles/README.md, r2f).
template <bool ANYHIT, bool STATS, bool HASINST, bool POOL, class Source>
GSB_D void traceStream(const SceneDev& sc, Source& src, uint32_t control, TraceStats* stats, uint32_t* overflowFlag) {
    const uint32_t refillMin = control & 0xffu;
    const int vote = (int)((control >> 8) & 0xffu);
    const int triMin = (int)((control >> 16) & 0xffu);
    constexpr int kLocalStack = HASINST ? kStackSize - kSmemStack : kWorldStackSize - kSmemStack;
    constexpr int kPoolWords = HASINST ? 5 : 4;
    constexpr uint32_t kPoolSize = HASINST ? GSB_POOL_SIZE_INST : GSB_POOL_SIZE;   // prepared rays per warp; 32 more are fetched when at most kPoolSize - 32 are left
    auto wrap = [](uint32_t e) { return e >= kPoolSize ? e - kPoolSize : e; };
    __shared__ uint2 s_stack[kSmemStack][kTraceBlock];
    // pool entry, one float4 per row: A = origin.xyz, tmax | B = 1/d .xyz (slab tests), octant + 8 * dominant axis |
    // C = shear constants Sx, Sy, Sz, tag | D = tmin, aux, -, - | E (two-level scenes) = direction.xyz, -
    __shared__ float4 s_pool[POOL ? kTraceBlock / 32 : 1][POOL ? kPoolWords : 1][POOL ? kPoolSize : 1];
    // Cold per-lane state lives in shared memory ([word][thread]: conflict-free), not in registers: it is written when a
    // ray starts or a closer hit is accepted and read when the ray is handed in (or, the world ray, when an instance is
    // entered / left), never inside the node and triangle tests.  Rows: tag, aux, hit u, v, det, inst, prim, leaf, and for
    // two-level scenes the world ray's origin and direction.
    constexpr int kColdWords = HASINST ? 14 : 8;
    __shared__ uint32_t s_cold[kColdWords][kTraceBlock];
    enum { C_TAG = 0, C_AUX, C_U, C_V, C_DET, C_INST, C_PRIM, C_LEAF, C_OX, C_OY, C_OZ, C_DX, C_DY, C_DZ };
#define GSB_COLD(w) s_cold[w][tid]
#define GSB_COLDF(w) __uint_as_float(s_cold[w][tid])
    uint2 lstack[kLocalStack];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    float4 (*pool)[POOL ? kPoolSize : 1] = s_pool[POOL ? tid >> 5 : 0];
    uint32_t poolHead = 0, poolCount = 0;   // warp-uniform ring state

    // the lane's ray
    bool hasRay = false;       // the lane holds a ray (in flight, or ended and not yet handed in)
    float tmin = 0.0f, tmax = 0.0f;
    float hitT = 0.0f;         // closest accepted t so far (tmax while there is none); the rest of the hit record is cold
    bool active = false;
    int sp = 0;
    // current level: origin, reciprocal direction and octant used by the node tests; shear constants used by the triangle tests
    float cox = 0.0f, coy = 0.0f, coz = 0.0f, cidx = 0.0f, cidy = 0.0f, cidz = 0.0f;
    uint32_t oct7 = 0;
    ObjRay oray;
    uint2 G = make_uint2(0u, 0u), T = make_uint2(0u, 0u);
    // two-level state (HASINST only)
    int spMark = 0;            // >= 0: inside a bottom-level structure whose stack starts here; -1: at the top level
    bool inWorld = true;       // the bottom-level structure is the world level
    uint32_t curInst = 0;
    const uint4* nodes = sc.worldNodes;
    const float4* tris = sc.worldTris;

#define GSB_PUSH(entry)                                                          \
    do {                                                                         \
        if (sp < kSmemStack) s_stack[sp++][tid] = (entry);                       \
        else if (sp < kSmemStack + kLocalStack) lstack[sp++ - kSmemStack] = (entry); \
        else if (overflowFlag) *overflowFlag = 1u;                               \
    } while (0)

    for (;;) {
        if (POOL) {
        // ---- hand in the rays that ended in the last pass
        const uint32_t idle = __ballot_sync(0xffffffffu, !active);
        if (idle) {
            const bool done = !active && hasRay;
            if (__any_sync(0xffffffffu, done)) {
                HitRecord hit;
                hit.t = hitT; hit.u = 0.0f; hit.v = 0.0f; hit.inst = GSB_NO_HIT; hit.prim = GSB_NO_HIT; hit.leaf = kNoLeaf;
                uint32_t tag = kNoTag, aux = 0u;
                if (done) {
                    tag = GSB_COLD(C_TAG); aux = GSB_COLD(C_AUX);
                    hit.inst = GSB_COLD(C_INST); hit.prim = GSB_COLD(C_PRIM); hit.leaf = GSB_COLD(C_LEAF);
                    if (!ANYHIT && hit.inst != GSB_NO_HIT) {
                        const float det = GSB_COLDF(C_DET);
                        hit.u = __fdiv_rn(GSB_COLDF(C_U), det);
                        hit.v = __fdiv_rn(GSB_COLDF(C_V), det);
                    }
                }
                src.commit(done, hit, tag, aux);
                if (done) hasRay = false;
            }
        }
        // ---- keep the pool stocked: 32 rays from the source with all lanes, prepared, appended to the ring
        if (poolCount <= kPoolSize - 32u && !src.drained()) {
            RayJob job;
            const bool got = src.fetch(true, job);
            const uint32_t m = __ballot_sync(0xffffffffu, got);
            if (got) {
                const uint32_t e = wrap(wrap(poolHead + poolCount) + (uint32_t)__popc(m & ((1u << lane) - 1u)));
                ObjRay r;
                makeObjRay(job.ox, job.oy, job.oz, job.dx, job.dy, job.dz, r);
                const uint32_t oct = (job.dx < 0.0f ? 0u : 4u) | (job.dy < 0.0f ? 0u : 2u) | (job.dz < 0.0f ? 0u : 1u);
                pool[0][e] = make_float4(job.ox, job.oy, job.oz, job.tmax);
                pool[1][e] = make_float4(slabInv(job.dx), slabInv(job.dy), slabInv(job.dz), __uint_as_float(oct | (r.perm << 3)));
                pool[2][e] = make_float4(r.Sx, r.Sy, r.Sz, __uint_as_float(job.tag));
                pool[3][e] = make_float4(job.tmin, __uint_as_float(job.aux), 0.0f, 0.0f);
                if (HASINST) pool[4][e] = make_float4(job.dx, job.dy, job.dz, 0.0f);
            }
            poolCount += (uint32_t)__popc(m);
            __syncwarp();
        }
        // ---- idle lanes take their next ray out of the pool
        if (idle && poolCount && ((uint32_t)__popc(idle) >= refillMin || idle == 0xffffffffu || src.drained())) {
            const uint32_t rank = (uint32_t)__popc(idle & ((1u << lane) - 1u));
            const uint32_t n = min((uint32_t)__popc(idle), poolCount);
            if (!active && rank < n) {
                const uint32_t e = wrap(poolHead + rank);
                const float4 a = pool[0][e], b = pool[1][e], c = pool[2][e], d = pool[3][e];
                active = true;
                cox = a.x; coy = a.y; coz = a.z; tmax = a.w;
                cidx = b.x; cidy = b.y; cidz = b.z;
                const uint32_t bits = __float_as_uint(b.w);
                oct7 = bits & 7u;
                oray.perm = bits >> 3;
                permute(oray.perm, a.x, a.y, a.z, oray.ox, oray.oy, oray.oz);
                oray.Sx = c.x; oray.Sy = c.y; oray.Sz = c.z;
                hasRay = true;
                GSB_COLD(C_TAG) = __float_as_uint(c.w);
                tmin = d.x; GSB_COLD(C_AUX) = __float_as_uint(d.y);
                if (HASINST) {
                    const float4 dd = pool[4][e];
                    GSB_COLD(C_OX) = __float_as_uint(a.x); GSB_COLD(C_OY) = __float_as_uint(a.y); GSB_COLD(C_OZ) = __float_as_uint(a.z);
                    GSB_COLD(C_DX) = __float_as_uint(dd.x); GSB_COLD(C_DY) = __float_as_uint(dd.y); GSB_COLD(C_DZ) = __float_as_uint(dd.z);
                    spMark = 0; inWorld = true; nodes = sc.worldNodes; tris = sc.worldTris;
                }
                hitT = tmax; GSB_COLD(C_INST) = GSB_NO_HIT; GSB_COLD(C_PRIM) = GSB_NO_HIT;
                sp = 0;
                G = make_uint2(0u, 0x80000000u);  // root group: bit 31, imask 0 -> node childBase + 0
                T = make_uint2(0u, 0u);
            }
            poolHead = wrap(poolHead + n);
            poolCount -= n;
            __syncwarp();   // the entries just read may be overwritten by the next fetch
        }
        } else {
        // ---- refill straight from the source: hand in finished rays, take new ones (warp-uniform branch).  While the
        // stream still has rays a refill waits until refillMin lanes are idle, so that its fixed cost (and the latency
        // of the state loads) is shared by several rays.
        const uint32_t idle = __ballot_sync(0xffffffffu, !active);
        const uint32_t pending = __ballot_sync(0xffffffffu, !active && hasRay);
        if (idle == 0xffffffffu || (src.drained() ? pending != 0u : (uint32_t)__popc(idle) >= refillMin)) {
            const bool done = !active && hasRay;
            HitRecord hit;
            hit.t = hitT; hit.u = 0.0f; hit.v = 0.0f; hit.inst = GSB_NO_HIT; hit.prim = GSB_NO_HIT; hit.leaf = kNoLeaf;
            uint32_t tag = kNoTag, aux = 0u;
            if (done) {
                tag = GSB_COLD(C_TAG); aux = GSB_COLD(C_AUX);
                hit.inst = GSB_COLD(C_INST); hit.prim = GSB_COLD(C_PRIM); hit.leaf = GSB_COLD(C_LEAF);
                if (!ANYHIT && hit.inst != GSB_NO_HIT) {
                    const float det = GSB_COLDF(C_DET);
                    hit.u = __fdiv_rn(GSB_COLDF(C_U), det);
                    hit.v = __fdiv_rn(GSB_COLDF(C_V), det);
                }
            }
            src.commit(done, hit, tag, aux);
            if (!active) hasRay = false;
            RayJob job;
            const bool got = src.fetch(!active, job);
            if (got) {
                active = true;
                hasRay = true;
                GSB_COLD(C_TAG) = job.tag; GSB_COLD(C_AUX) = job.aux; tmin = job.tmin; tmax = job.tmax;
                hitT = tmax; GSB_COLD(C_INST) = GSB_NO_HIT; GSB_COLD(C_PRIM) = GSB_NO_HIT;
                cox = job.ox; coy = job.oy; coz = job.oz;
                cidx = slabInv(job.dx); cidy = slabInv(job.dy); cidz = slabInv(job.dz);
                oct7 = (job.dx < 0.0f ? 0u : 4u) | (job.dy < 0.0f ? 0u : 2u) | (job.dz < 0.0f ? 0u : 1u);
                makeObjRay(job.ox, job.oy, job.oz, job.dx, job.dy, job.dz, oray);
                if (HASINST) {
                    GSB_COLD(C_OX) = __float_as_uint(job.ox); GSB_COLD(C_OY) = __float_as_uint(job.oy); GSB_COLD(C_OZ) = __float_as_uint(job.oz);
                    GSB_COLD(C_DX) = __float_as_uint(job.dx); GSB_COLD(C_DY) = __float_as_uint(job.dy); GSB_COLD(C_DZ) = __float_as_uint(job.dz);
                    spMark = 0; inWorld = true; nodes = sc.worldNodes; tris = sc.worldTris;
                }
                sp = 0;
                G = make_uint2(0u, 0x80000000u);  // root group: bit 31, imask 0 -> node childBase + 0
                T = make_uint2(0u, 0u);
            }
        }
        }
        if (!__any_sync(0xffffffffu, active)) {
            if (poolCount == 0u && src.drained()) break;
            continue;
        }
        // ---- vote: which phases run in this pass
        bool run1, run2a = false, run2b;
        {
            const bool hasT = active && T.y != 0u;
            const int n1 = __popc(__ballot_sync(0xffffffffu, active && !hasT && (G.y & 0xff000000u) != 0u));
            const int n2a = HASINST ? __popc(__ballot_sync(0xffffffffu, hasT && spMark < 0)) : 0;
            const int n2b = __popc(__ballot_sync(0xffffffffu, hasT && (!HASINST || spMark >= 0)));
            run1 = n1 >= vote; run2b = n2b >= vote;
            if (HASINST) run2a = n2a >= vote;
            if (!(run1 || run2a || run2b)) {
                if (n1 >= n2a && n1 >= n2b) run1 = true;
                else if (n2a >= n2b) run2a = true;
                else run2b = true;
            }
        }
        // ---- phase 1: descend into the pending internal child that comes first in octant order
        if (run1 && active && !T.y && (G.y & 0xff000000u)) {
            const uint32_t bit = 31u - (uint32_t)__clz(G.y);
            G.y &= ~(1u << bit);
            const uint32_t slot = (bit - 24u) ^ oct7;
            const uint32_t nodeIndex = G.x + __popc(G.y & ((1u << slot) - 1u) & 0xffu);
            if (G.y & 0xff000000u) GSB_PUSH(G);
            if (STATS) stats->nodes++;
            intersectNode(HASINST ? nodes : sc.worldNodes, nodeIndex, cox, coy, coz, cidx, cidy, cidz, tmin, hitT, oct7, G, T);
        }
        // ---- phase 2a (top level): enter one instance, park the rest of this node
        if (HASINST && run2a && active && spMark < 0 && T.y) {
            const uint32_t bit = (uint32_t)__ffs(T.y) - 1u;
            T.y &= ~(1u << bit);
            const TraceInstance* ti = sc.tlasInst + (T.x + bit);
            if (G.y & 0xff000000u) GSB_PUSH(G);
            if (T.y) GSB_PUSH(T);
            spMark = sp;
            const float4 r0 = __ldg(&ti->w2o[0]), r1 = __ldg(&ti->w2o[1]), r2 = __ldg(&ti->w2o[2]);
            const uint4 tail = __ldg(reinterpret_cast<const uint4*>(ti) + 3);
            if (STATS) stats->insts++;
            // o' = M^-1 (o,1), d' = M^-1 (d,0); rounded op by op, same order as the checker
            const float ox = GSB_COLDF(C_OX), oy = GSB_COLDF(C_OY), oz = GSB_COLDF(C_OZ), dx = GSB_COLDF(C_DX), dy = GSB_COLDF(C_DY), dz = GSB_COLDF(C_DZ);
            const float tox = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r0.x, ox), __fmul_rn(r0.y, oy)), __fmul_rn(r0.z, oz)), r0.w);
            const float toy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r1.x, ox), __fmul_rn(r1.y, oy)), __fmul_rn(r1.z, oz)), r1.w);
            const float toz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r2.x, ox), __fmul_rn(r2.y, oy)), __fmul_rn(r2.z, oz)), r2.w);
            const float tdx = __fadd_rn(__fadd_rn(__fmul_rn(r0.x, dx), __fmul_rn(r0.y, dy)), __fmul_rn(r0.z, dz));
            const float tdy = __fadd_rn(__fadd_rn(__fmul_rn(r1.x, dx), __fmul_rn(r1.y, dy)), __fmul_rn(r1.z, dz));
            const float tdz = __fadd_rn(__fadd_rn(__fmul_rn(r2.x, dx), __fmul_rn(r2.y, dy)), __fmul_rn(r2.z, dz));
            makeObjRay(tox, toy, toz, tdx, tdy, tdz, oray);
            const uint64_t base = ((uint64_t)tail.y << 32) | (uint64_t)tail.x;
            nodes = reinterpret_cast<const uint4*>(base);
            tris = reinterpret_cast<const float4*>(nodes + (tail.z & ((1u << kDirectShift) - 1u)));
            curInst = tail.w;
            const uint32_t direct = tail.z >> kDirectShift;
            if (direct) {
                // node-less mesh (<= 4 triangles): its world box was already tested in the TLAS node
                G = make_uint2(0u, 0u);
                T = make_uint2(0u, (1u << direct) - 1u);
            } else {
                cox = tox; coy = toy; coz = toz;
                cidx = slabInv(tdx); cidy = slabInv(tdy); cidz = slabInv(tdz);
                oct7 = (tdx < 0.0f ? 0u : 4u) | (tdy < 0.0f ? 0u : 2u) | (tdz < 0.0f ? 0u : 1u);
                G = make_uint2(0u, 0x80000000u);
                T = make_uint2(0u, 0u);
            }
        }
        // ---- phase 2b (bottom level): test the pending triangles, one per lane per round; they can only shrink
        // the interval.  The loop keeps spinning while at least triMin lanes still have triangles.
        if (run2b) {
            for (bool first = true;; first = false) {
                const bool w = active && (!HASINST || spMark >= 0) && T.y != 0u;
                const uint32_t m = __ballot_sync(0xffffffffu, w);
                if (!m || (!first && __popc(m) < triMin)) break;
                if (w) {
                    const uint32_t bit = (uint32_t)__ffs(T.y) - 1u;
                    T.y &= ~(1u << bit);
                    const float4* tp = (HASINST ? tris : sc.worldTris) + ((size_t)(T.x + bit) * 9 + oray.perm * 3u);
                    const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                    if (STATS) stats->tris++;
                    float Tn, V, W, det;
                    if (woopEdges(oray, v0, v1, v2, Tn, V, W, det)) {
                        // The rule is on the correctly rounded quotient t = T / det.  An occlusion ray only needs to know
                        // whether t lies in (tmin, tmax): an approximate quotient (one MUFU.RCP and a product, relative error
                        // below 2^-21) decides every candidate that is not within 2^-16 of a bound, and only those pay for
                        // the IEEE division — same decisions (r2k: k_shadow 77.3 -> 74.8 ms per Cornell step).  A closest-hit
                        // candidate that survives needs the exact t anyway; there the filter measured as a small loss.
                        bool inside;
                        float t = 0.0f;
                        if (ANYHIT) {
                            const float ta = Tn * slabInvRaw(det);
                            const float slack = 1.52587890625e-05f * fabsf(ta);
                            if (ta - slack > tmax || ta + slack < tmin) inside = false;
                            else if (ta - slack > tmin && ta + slack < tmax) inside = true;
                            else {
                                t = __fdiv_rn(Tn, det);
                                inside = t > tmin && t < tmax;
                            }
                        } else {
                            t = __fdiv_rn(Tn, det);
                            inside = t > tmin && t < tmax;
                        }
                        if (inside) {
                            if (ANYHIT) {
                                // occluded: drop everything still pending, phase 3 ends the ray
                                GSB_COLD(C_INST) = 0u; GSB_COLD(C_PRIM) = 0u; hitT = 0.0f;
                                T.y = 0u; G.y = 0u; sp = 0;
                                if (HASINST) spMark = -1;
                            } else {
                                const uint32_t prim = __float_as_uint(v0.w);
                                const uint32_t inst = (!HASINST || inWorld) ? __float_as_uint(v1.w) : curInst;
                                bool closer = t < hitT;
                                if (t == hitT) {   // a tie: the smaller (gl_InstanceID, gl_PrimitiveID) wins
                                    const uint32_t hw = GSB_COLD(C_INST), hp = GSB_COLD(C_PRIM);
                                    const uint32_t hi = hw == GSB_NO_HIT ? GSB_NO_HIT : (hw & kInstMask), ci = inst & kInstMask;
                                    closer = ci < hi || (ci == hi && prim < hp);
                                }
                                if (closer) {
                                    hitT = t;
                                    GSB_COLD(C_U) = __float_as_uint(V); GSB_COLD(C_V) = __float_as_uint(W); GSB_COLD(C_DET) = __float_as_uint(det);
                                    GSB_COLD(C_INST) = inst; GSB_COLD(C_PRIM) = prim;
                                    GSB_COLD(C_LEAF) = (!HASINST || inWorld) ? T.x + bit : kNoLeaf;
                                }
                            }
                        }
                    }
                }
            }
        }
        // ---- phase 3: nothing pending in this node -> leave the level if its stack is drained, then pop
        if (active && !T.y && !(G.y & 0xff000000u)) {
            bool pop = true;
            if (HASINST && sp == spMark) {
                // back to the top level: the world ray again
                spMark = -1;
                nodes = sc.tlasNodes;
                cox = GSB_COLDF(C_OX); coy = GSB_COLDF(C_OY); coz = GSB_COLDF(C_OZ);
                const float wdx = GSB_COLDF(C_DX), wdy = GSB_COLDF(C_DY), wdz = GSB_COLDF(C_DZ);
                cidx = slabInv(wdx); cidy = slabInv(wdy); cidz = slabInv(wdz);
                oct7 = (wdx < 0.0f ? 0u : 4u) | (wdy < 0.0f ? 0u : 2u) | (wdz < 0.0f ? 0u : 1u);
                if (inWorld) {
                    // the world level is done: start the top level at its root
                    inWorld = false;
                    G = make_uint2(0u, 0x80000000u);
                    pop = false;
                }
            }
            if (pop) {
                if (sp == 0) {
                    active = false;  // this ray is done; its result is handed in at the top of the next pass
                } else {
                    --sp;
                    const uint2 e = sp < kSmemStack ? s_stack[sp][tid] : lstack[sp - kSmemStack];
                    if (e.y & 0xff000000u) { G = e; T = make_uint2(0u, 0u); }
                    else { T = e; G = make_uint2(0u, 0u); }
                }
            }
        }
    }
#undef GSB_PUSH
#undef GSB_COLD
#undef GSB_COLDF
}

}  // namespace gsb

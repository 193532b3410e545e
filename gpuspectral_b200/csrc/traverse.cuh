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
// Bottom level primitive = triangle (3 x float4: v0.xyz|primId, v1.xyz, v2.xyz), 48 B.
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

constexpr int kStackSize = 48;
constexpr uint32_t kSentinel = 0u;  // stack entry with .y == 0 marks "leave the current instance"

struct TraceStats { uint32_t nodes, tris, insts; };

struct HitRecord {
    float t, u, v;
    uint32_t inst, prim;
};

// object-space ray with the Woop shear constants
struct ObjRay {
    float ox, oy, oz;     // origin, permuted so that z is the dominant axis
    float Sx, Sy, Sz;
    uint32_t perm;        // kz | swap << 2
};

GSB_D void permute(uint32_t perm, float x, float y, float z, float& px, float& py, float& pz) {
    const uint32_t kz = perm & 3u;
    pz = kz == 0 ? x : (kz == 1 ? y : z);
    float a = kz == 0 ? y : (kz == 1 ? z : x);
    float b = kz == 0 ? z : (kz == 1 ? x : y);
    const bool swap = (perm >> 2) != 0;
    px = swap ? b : a;
    py = swap ? a : b;
}

GSB_D void makeObjRay(float ox, float oy, float oz, float dx, float dy, float dz, ObjRay& r) {
    const float ax = fabsf(dx), ay = fabsf(dy), az = fabsf(dz);
    const uint32_t kz = (ax > ay) ? ((ax > az) ? 0u : 2u) : ((ay > az) ? 1u : 2u);
    const float dkz = kz == 0 ? dx : (kz == 1 ? dy : dz);
    r.perm = kz | ((dkz < 0.0f) ? 4u : 0u);
    float pdx, pdy, pdz;
    permute(r.perm, dx, dy, dz, pdx, pdy, pdz);
    r.Sx = __fdiv_rn(pdx, pdz);
    r.Sy = __fdiv_rn(pdy, pdz);
    r.Sz = __fdiv_rn(1.0f, pdz);
    permute(r.perm, ox, oy, oz, r.ox, r.oy, r.oz);
}

// Watertight ray/triangle test; every operation individually rounded (no FMA contraction).
GSB_D bool woopTest(const ObjRay& r, float4 v0, float4 v1, float4 v2, float& t, float& bu, float& bv) {
    float Akx, Aky, Akz, Bkx, Bky, Bkz, Ckx, Cky, Ckz;
    permute(r.perm, v0.x, v0.y, v0.z, Akx, Aky, Akz);
    permute(r.perm, v1.x, v1.y, v1.z, Bkx, Bky, Bkz);
    permute(r.perm, v2.x, v2.y, v2.z, Ckx, Cky, Ckz);
    Akx = __fsub_rn(Akx, r.ox); Aky = __fsub_rn(Aky, r.oy); Akz = __fsub_rn(Akz, r.oz);
    Bkx = __fsub_rn(Bkx, r.ox); Bky = __fsub_rn(Bky, r.oy); Bkz = __fsub_rn(Bkz, r.oz);
    Ckx = __fsub_rn(Ckx, r.ox); Cky = __fsub_rn(Cky, r.oy); Ckz = __fsub_rn(Ckz, r.oz);
    const float Ax = __fsub_rn(Akx, __fmul_rn(r.Sx, Akz)), Ay = __fsub_rn(Aky, __fmul_rn(r.Sy, Akz));
    const float Bx = __fsub_rn(Bkx, __fmul_rn(r.Sx, Bkz)), By = __fsub_rn(Bky, __fmul_rn(r.Sy, Bkz));
    const float Cx = __fsub_rn(Ckx, __fmul_rn(r.Sx, Ckz)), Cy = __fsub_rn(Cky, __fmul_rn(r.Sy, Ckz));
    float U = __fsub_rn(__fmul_rn(Cx, By), __fmul_rn(Cy, Bx));
    float V = __fsub_rn(__fmul_rn(Ax, Cy), __fmul_rn(Ay, Cx));
    float W = __fsub_rn(__fmul_rn(Bx, Ay), __fmul_rn(By, Ax));
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        // products of two floats are exact in double; one rounding in the subtraction, one to float
        U = (float)__dsub_rn(__dmul_rn((double)Cx, (double)By), __dmul_rn((double)Cy, (double)Bx));
        V = (float)__dsub_rn(__dmul_rn((double)Ax, (double)Cy), __dmul_rn((double)Ay, (double)Cx));
        W = (float)__dsub_rn(__dmul_rn((double)Bx, (double)Ay), __dmul_rn((double)By, (double)Ax));
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = __fadd_rn(__fadd_rn(U, V), W);
    if (det == 0.0f) return false;
    const float Az = __fmul_rn(r.Sz, Akz), Bz = __fmul_rn(r.Sz, Bkz), Cz = __fmul_rn(r.Sz, Ckz);
    const float T = __fadd_rn(__fadd_rn(__fmul_rn(U, Az), __fmul_rn(V, Bz)), __fmul_rn(W, Cz));
    t = __fdiv_rn(T, det);
    bu = __fdiv_rn(V, det);
    bv = __fdiv_rn(W, det);
    return true;
}

GSB_D float safeInv(float d) {
    // keeps 1/d finite for the box tests; 2^-80 as in the CWBVH paper
    const float eps = 8.271806125530277e-25f;
    return 1.0f / (fabsf(d) > eps ? d : copysignf(eps, d));
}

GSB_D uint32_t byteOf(uint32_t w, int j) { return (w >> (8 * j)) & 0xffu; }

// Intersects the 8 quantised child boxes of one node.  Returns the node group
// (childBase, internal hit bits 24..31 | imask) and the leaf group (leafBase, hit bits 0..23).
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
    uint32_t hits = 0;
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const uint32_t meta4 = half ? w1.w : w1.z;
        const uint32_t qlox = half ? w2.y : w2.x, qloy = half ? w2.w : w2.z, qloz = half ? w3.y : w3.x;
        const uint32_t qhix = half ? w3.w : w3.z, qhiy = half ? w4.y : w4.x, qhiz = half ? w4.w : w4.z;
        const uint32_t nearx = nx ? qhix : qlox, farx = nx ? qlox : qhix;
        const uint32_t neary = ny ? qhiy : qloy, fary = ny ? qloy : qhiy;
        const uint32_t nearz = nz ? qhiz : qloz, farz = nz ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const uint32_t m = byteOf(meta4, j);
            if (m == 0) continue;
            const float t0x = fmaf((float)byteOf(nearx, j), sx, bx), t1x = fmaf((float)byteOf(farx, j), sx, bx);
            const float t0y = fmaf((float)byteOf(neary, j), sy, by), t1y = fmaf((float)byteOf(fary, j), sy, by);
            const float t0z = fmaf((float)byteOf(nearz, j), sz, bz), t1z = fmaf((float)byteOf(farz, j), sz, bz);
            const float tn = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, tmin));
            const float tf = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            if (tn <= tf) {
                const int i = half * 4 + j;
                if ((imask >> i) & 1u) hits |= 1u << (24 + ((uint32_t)i ^ oct7));
                else hits |= (m >> 5) << (m & 31u);
            }
        }
    }
    nodeGroup = make_uint2(w1.x, (hits & 0xff000000u) | imask);
    leafGroup = make_uint2(w1.y, hits & 0x00ffffffu);
}

// Traces one ray.  ANYHIT: returns as soon as any triangle has tmin < t < tmax (hit.inst = 0).
// STATS: counts fetched nodes / triangles / instance records.
template <bool ANYHIT, bool STATS>
GSB_D void traceRay(const SceneDev& sc, float ox, float oy, float oz, float dx, float dy, float dz, float tmin, float tmax,
                    HitRecord& hit, TraceStats* stats, uint32_t* overflowFlag) {
    hit.t = tmax; hit.u = 0.0f; hit.v = 0.0f; hit.inst = GSB_NO_HIT; hit.prim = GSB_NO_HIT;
    uint2 stack[kStackSize];
    int sp = 0;

    // world-space box-test constants
    const float widx = safeInv(dx), widy = safeInv(dy), widz = safeInv(dz);
    const uint32_t woct = (dx < 0.0f ? 0u : 4u) | (dy < 0.0f ? 0u : 2u) | (dz < 0.0f ? 0u : 1u);

    // current level state (top level first)
    const uint4* nodes = sc.tlasNodes;
    const float4* tris = nullptr;
    float cox = ox, coy = oy, coz = oz, cidx = widx, cidy = widy, cidz = widz;
    uint32_t oct7 = woct;
    bool inBlas = false;
    uint32_t curInst = 0;
    ObjRay oray;

    uint2 G = make_uint2(0u, 0x80000000u);  // root group: bit 31, imask 0 -> node childBase + 0
    uint2 T = make_uint2(0u, 0u);

#define GSB_PUSH(entry)                                   \
    do {                                                  \
        if (sp < kStackSize) stack[sp++] = (entry);       \
        else if (overflowFlag) *overflowFlag = 1u;        \
    } while (0)

    for (;;) {
        if (T.y) {
            // leaf work first: it can only shrink the interval for the nodes still pending
            const uint32_t bit = (uint32_t)__ffs(T.y) - 1u;
            T.y &= ~(1u << bit);
            if (inBlas) {
                const float4* tp = tris + (size_t)(T.x + bit) * 3;
                const float4 v0 = __ldg(tp), v1 = __ldg(tp + 1), v2 = __ldg(tp + 2);
                if (STATS) stats->tris++;
                float t, bu, bv;
                if (woopTest(oray, v0, v1, v2, t, bu, bv)) {
                    if (ANYHIT) {
                        if (t > tmin && t < tmax) { hit.inst = 0u; hit.prim = 0u; hit.t = 0.0f; return; }
                    } else if (t > tmin && t < tmax) {
                        const uint32_t prim = __float_as_uint(v0.w);
                        if (t < hit.t || (t == hit.t && (curInst < hit.inst || (curInst == hit.inst && prim < hit.prim)))) {
                            hit.t = t; hit.u = bu; hit.v = bv; hit.inst = curInst; hit.prim = prim;
                        }
                    }
                }
            } else {
                // enter an instance: park what is left of this top-level node, then a sentinel
                const TraceInstance* ti = sc.tlasInst + (T.x + bit);
                if (G.y & 0xff000000u) GSB_PUSH(G);
                if (T.y) GSB_PUSH(T);
                GSB_PUSH(make_uint2(0u, kSentinel));
                const float4 r0 = __ldg(&ti->w2o[0]), r1 = __ldg(&ti->w2o[1]), r2 = __ldg(&ti->w2o[2]);
                const uint4 tail = __ldg(reinterpret_cast<const uint4*>(ti) + 3);
                if (STATS) stats->insts++;
                // o' = M^-1 (o,1), d' = M^-1 (d,0); rounded op by op, same order as the checker
                const float tox = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r0.x, ox), __fmul_rn(r0.y, oy)), __fmul_rn(r0.z, oz)), r0.w);
                const float toy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r1.x, ox), __fmul_rn(r1.y, oy)), __fmul_rn(r1.z, oz)), r1.w);
                const float toz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r2.x, ox), __fmul_rn(r2.y, oy)), __fmul_rn(r2.z, oz)), r2.w);
                const float tdx = __fadd_rn(__fadd_rn(__fmul_rn(r0.x, dx), __fmul_rn(r0.y, dy)), __fmul_rn(r0.z, dz));
                const float tdy = __fadd_rn(__fadd_rn(__fmul_rn(r1.x, dx), __fmul_rn(r1.y, dy)), __fmul_rn(r1.z, dz));
                const float tdz = __fadd_rn(__fadd_rn(__fmul_rn(r2.x, dx), __fmul_rn(r2.y, dy)), __fmul_rn(r2.z, dz));
                makeObjRay(tox, toy, toz, tdx, tdy, tdz, oray);
                cox = tox; coy = toy; coz = toz;
                cidx = safeInv(tdx); cidy = safeInv(tdy); cidz = safeInv(tdz);
                oct7 = (tdx < 0.0f ? 0u : 4u) | (tdy < 0.0f ? 0u : 2u) | (tdz < 0.0f ? 0u : 1u);
                const uint64_t base = ((uint64_t)tail.y << 32) | (uint64_t)tail.x;
                nodes = reinterpret_cast<const uint4*>(base);
                tris = reinterpret_cast<const float4*>(nodes + tail.z);
                curInst = tail.w;
                inBlas = true;
                G = make_uint2(0u, 0x80000000u);
                T = make_uint2(0u, 0u);
            }
        } else if (G.y & 0xff000000u) {
            // descend into the pending internal child that comes first in octant order
            const uint32_t bit = 31u - (uint32_t)__clz(G.y);
            G.y &= ~(1u << bit);
            const uint32_t slot = (bit - 24u) ^ oct7;
            const uint32_t rel = __popc(G.y & 0xffu & ~(0xffffffffu << slot));
            const uint32_t nodeIndex = G.x + rel;
            if (G.y & 0xff000000u) GSB_PUSH(G);
            if (STATS) stats->nodes++;
            intersectNode(nodes, nodeIndex, cox, coy, coz, cidx, cidy, cidz, tmin, hit.t, oct7, G, T);
        } else {
            if (sp == 0) break;
            const uint2 e = stack[--sp];
            if (e.y == kSentinel) {
                inBlas = false;
                nodes = sc.tlasNodes;
                cox = ox; coy = oy; coz = oz; cidx = widx; cidy = widy; cidz = widz;
                oct7 = woct;
                G = make_uint2(0u, 0u);
                T = make_uint2(0u, 0u);
            } else if (e.y & 0xff000000u) {
                G = e;
            } else {
                T = e;
            }
        }
    }
#undef GSB_PUSH
    if (ANYHIT) { hit.inst = GSB_NO_HIT; hit.prim = GSB_NO_HIT; }
}

}  // namespace gsb

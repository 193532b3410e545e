// wavefront.cu — the wavefront path tracer.
//
// The reference runs one megakernel-by-recursion per pixel (raygen.rgen:29-109 calling
// rayhit.rchit:666-797 through traceRayEXT).  Here the same per-path state machine is cut
// into stages that each run as one wide kernel over many paths:
//
//   k_extend   first, for slots whose path has ended: adds the sample to the image (raygen.rgen:84-108) and
//              starts the slot's next pixel-sample (raygen.rgen:31-48); then closest-hit traversal of every
//              live path's ray (raygen.rgen:54-58), bins the hits into one queue per BSDF type; a miss
//              (miss.rmiss) ends the path
//   k_shade<T> material-sorted closest-hit shading             (rayhit.rchit:666-797) + the
//              per-bounce bookkeeping of the raygen loop       (raygen.rgen:59-80)
//   k_shadow   any-hit traversal of the NEE segments           (rayhit.rchit:737-757)
// Three launches per wavefront iteration (for one BSDF type in use); queue counters are double-banked by
// iteration parity so that no separate reset kernel is needed.
//
// Slots are not tied to pixels: a slot whose path ends takes the next pixel-sample from a global work counter
// (work item w = sample w / npix of pixel w % npix), so every slot stays busy until the call's work is exhausted and
// only the last few iterations of a call run with a thinning wavefront.  Finished samples are added to the image with
// one 128-bit atomic add (rgb, count).  Every path carries its own RNG word; each stage consumes draws in the order
// the shader does, which makes per-path results match the checker's.
#include <algorithm>
#include <cstdio>
#include <vector>

#include "bsdf.cuh"
#include "context.h"
#include "traverse.cuh"

namespace gsb {

namespace {

constexpr int kBlock = kTraceBlock;
constexpr uint32_t kRetired = 0xffffffffu;
constexpr uint32_t kFinished = 0x80000000u;  // sampleK bit: the path has ended, its sample waits to be added to the image
constexpr uint32_t kFlagWasDelta = 1u << 16;
constexpr uint32_t kFlagCountEmitted = 1u << 17;

struct RenderArgs {
    uint32_t width, height, npix;
    uint32_t first, count, stride;  // timestamps first + k*stride, k < count
    uint32_t slots;                 // paths in flight
    float camZ;                     // (max(W,H)/2) / tan(fov/2) of rayDir(), raygen.rgen:20-25 (evaluated once on the host)
    unsigned long long totalWork;   // npix * count pixel-samples
    float4* accum;                  // the image: rgb = sum of samples, a = samples added
};

// Warp-aggregated append: every lane of the warp must call this (pred false = no push).
// key selects one of the queue counters; returns the position for lanes with pred.
GSB_D uint32_t queueSlot(uint32_t* counters, uint32_t key, bool pred) {
    const uint32_t k = pred ? key : 0xffffu;
    const uint32_t peers = __match_any_sync(0xffffffffu, k);
    uint32_t base = 0;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t leader = (uint32_t)__ffs(peers) - 1u;
    if (pred && lane == leader) base = atomicAdd(&counters[key], (uint32_t)__popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t)__popc(peers & ((1u << lane) - 1u));
}

// raygen.rgen:20-48: seed + primary ray + payload initialisation for work item `work` (sample work / npix of
// pixel work % npix); the new ray is also returned in registers
GSB_D void startPath(const SceneDev& sc, const WavefrontState& wf, const RenderArgs& ra, uint32_t slot, unsigned long long work,
                     float4& o, float4& dOut, uint32_t& pixelOut) {
    const uint32_t k = (uint32_t)(work / ra.npix);
    const uint32_t pixel = (uint32_t)(work - (unsigned long long)k * ra.npix);
    const uint32_t x = pixel % ra.width, y = pixel / ra.width;
    const uint32_t timestamp = ra.first + k * ra.stride;
    // rayDir(size, fragCoord, fov)
    const float sx = (float)ra.width, sy = (float)ra.height;
    const float fx = (float)x - sx / 2.0f;
    const float fy = (float)y - sy / 2.0f;
    const float3 dir = normalize(f3(-fx, fy, ra.camZ));
    const float* V = sc.view;
    float3 d = f3(V[0] * dir.x + V[4] * dir.y + V[8] * dir.z, V[1] * dir.x + V[5] * dir.y + V[9] * dir.z,
                  V[2] * dir.x + V[6] * dir.y + V[10] * dir.z);
    d.y *= -1.0f;
    o = make_float4(sc.eye[0], sc.eye[1], sc.eye[2], 0.0f);
    dOut = make_float4(d.x, d.y, d.z, 0.0f);
    wf.seed[slot] = pcgHash(tea(ra.width * y + x, timestamp));
    wf.rayO[slot] = o;
    wf.rayD[slot] = dOut;
    wf.weightDW[slot] = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
    wf.result[slot] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kFlagCountEmitted));  // depth 0, wasDelta 0, countEmitted 1
    wf.sampleK[slot] = pixel;
    pixelOut = pixel;
}

__global__ void k_generate(SceneDev sc, WavefrontState wf, RenderArgs ra) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= ra.slots) return;
    float4 o, d;
    uint32_t pixel;
    if (slot < ra.totalWork) startPath(sc, wf, ra, slot, slot, o, d, pixel);
    else wf.sampleK[slot] = kRetired;
}

// ---- extend ---------------------------------------------------------------------------------------
// instrumented builds: fold one thread's fetch counts into the context counters
GSB_D void flushTraceStats(unsigned long long* stats, TraceStats ts, uint32_t traced) {
    uint32_t nn = ts.nodes, nt = ts.tris, ni = ts.insts, nr = traced;
    for (int o = 16; o > 0; o >>= 1) {
        nn += __shfl_xor_sync(0xffffffffu, nn, o);
        nt += __shfl_xor_sync(0xffffffffu, nt, o);
        ni += __shfl_xor_sync(0xffffffffu, ni, o);
        nr += __shfl_xor_sync(0xffffffffu, nr, o);
    }
    if ((threadIdx.x & 31) == 0 && nr) {
        atomicAdd(&stats[ST_NODES], (unsigned long long)nn);
        atomicAdd(&stats[ST_TRIS], (unsigned long long)nt);
        atomicAdd(&stats[ST_INSTS], (unsigned long long)ni);
        atomicAdd(&stats[ST_RAYS], (unsigned long long)nr);
    }
}

template <bool STATS>
__global__ void __launch_bounds__(kBlock) k_extend(SceneDev sc, WavefrontState wf, RenderArgs ra, uint32_t parity) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t* qc = wf.qCount + parity * Q_COUNT;
    if (slot == 0) {
        // roll the previous iteration's queue sizes into the totals and clear that bank for the next iteration
        uint32_t* prev = wf.qCount + (parity ^ 1u) * Q_COUNT;
        unsigned long long ext = prev[Q_MISS];
        for (int q = 0; q < GSB_BSDF_TYPE_COUNT; q++) ext += prev[q];
        wf.stats[ST_EXT] += ext;
        wf.stats[ST_SHADOW] += prev[Q_SHADOW];
        wf.stats[ST_ITERS] += 1;
        for (int q = 0; q < Q_COUNT; q++) prev[q] = 0;
    }
    uint32_t pix = kRetired;
    if (slot < ra.slots) pix = wf.sampleK[slot];
    float4 o, d;
    bool live = pix != kRetired;
    const bool finished = live && (pix & kFinished);

    // ---- the raygen epilogue of paths that ended last iteration, and the prologue of the slot's next path
    if (finished) {
        const float4 res = wf.result[slot];
        // raygen.rgen:106-108: a NaN sample leaves the image untouched
        if (!(isnan(res.x) || isnan(res.y) || isnan(res.z)))
            atomicAdd(&ra.accum[pix & ~kFinished], make_float4(res.x, res.y, res.z, 1.0f));
    }
    const uint32_t fmask = __ballot_sync(0xffffffffu, finished);
    if (fmask) {
        // next pixel-sample for every finished slot of the warp: one atomic per warp
        const uint32_t lane = threadIdx.x & 31u;
        const uint32_t leader = (uint32_t)__ffs(fmask) - 1u;
        unsigned long long base = 0;
        if (lane == leader) base = atomicAdd(&wf.stats[ST_NEXTWORK], (unsigned long long)__popc(fmask));
        base = __shfl_sync(0xffffffffu, base, leader);
        bool retire = false;
        if (finished) {
            const unsigned long long work = base + (unsigned long long)__popc(fmask & ((1u << lane) - 1u));
            if (work < ra.totalWork) startPath(sc, wf, ra, slot, work, o, d, pix);
            else { wf.sampleK[slot] = kRetired; retire = true; live = false; }
        }
        const uint32_t rmask = __ballot_sync(0xffffffffu, retire);
        if (rmask && lane == (uint32_t)__ffs(rmask) - 1u) atomicAdd(&wf.stats[ST_LIVE], (unsigned long long)(-(long long)__popc(rmask)));
    }
    if (live && !finished) { o = wf.rayO[slot]; d = wf.rayD[slot]; }

    // ---- closest hit
    uint32_t key = 0;
    TraceStats ts = {0u, 0u, 0u};
    if (live) {
        HitRecord hit;
        traceRay<false, STATS>(sc, o.x, o.y, o.z, d.x, d.y, d.z, 0.0f, 1e10f, hit, &ts, (uint32_t*)&wf.stats[ST_OVERFLOW]);
        if (hit.inst == GSB_NO_HIT) {
            key = Q_MISS;                                           // miss.rmiss: done = 1, nothing is added
            wf.sampleK[slot] = pix | kFinished;
        } else {
            wf.hitTUVP[slot] = make_float4(hit.t, hit.u, hit.v, __uint_as_float(hit.prim));
            wf.hitInst[slot] = hit.inst;
            key = __ldg(sc.instBsdf + hit.inst) >> 16;
        }
    }
    const uint32_t pos = queueSlot(qc, key, live);
    if (live && key != Q_MISS) wf.matQueue[(size_t)key * wf.capacity + pos] = slot;
    if (STATS) flushTraceStats(wf.stats, ts, live ? 1u : 0u);
}

// ---- shade ----------------------------------------------------------------------------------------
template <int TYPE>
__global__ void __launch_bounds__(kBlock) k_shade(SceneDev sc, WavefrontState wf, uint32_t parity) {
    uint32_t* qc = wf.qCount + parity * Q_COUNT;
    const uint32_t count = qc[TYPE];
    const uint32_t* queue = wf.matQueue + (size_t)TYPE * wf.capacity;
    const uint32_t rounds = (count + gridDim.x * blockDim.x - 1) / (gridDim.x * blockDim.x);
    for (uint32_t r = 0; r < rounds; r++) {
        const uint32_t i = (r * gridDim.x + blockIdx.x) * blockDim.x + threadIdx.x;
        const bool active = i < count;
        bool terminate = false, needShadow = false;
        uint32_t slot = 0;
        float4 shO, shD, shBase, shNee;
        if (active) {
            slot = queue[i];
            const float4 ro = wf.rayO[slot], rd = wf.rayD[slot];
            const float4 hit = wf.hitTUVP[slot];
            const uint32_t inst = wf.hitInst[slot];
            uint32_t rng = wf.seed[slot];                                         // rayhit.rchit:668
            float4 wdw = wf.weightDW[slot];
            float4 res = wf.result[slot];
            uint32_t flags = __float_as_uint(res.w);
            float3 weight = xyz(wdw);
            const float directWeight = wdw.w;
            const uint32_t prim = __float_as_uint(hit.w);

            // :671-688 instance record, positions through object->world, normals through transformInvT
            const float4* si = reinterpret_cast<const float4*>(sc.shadeInst + inst);
            const float4 m0 = __ldg(si), m1 = __ldg(si + 1), m2 = __ldg(si + 2);
            const float4 n0 = __ldg(si + 3), n1 = __ldg(si + 4), n2 = __ldg(si + 5);
            const uint4 p0 = __ldg(reinterpret_cast<const uint4*>(si) + 6);
            const uint4 p1 = __ldg(reinterpret_cast<const uint4*>(si) + 7);
            const float4* vpos = reinterpret_cast<const float4*>(((uint64_t)p0.y << 32) | p0.x);
            const float4* vnrm = reinterpret_cast<const float4*>(((uint64_t)p0.w << 32) | p0.z);
            const uint32_t handle = p1.x, twofaced = p1.y;
            const float3 emission = f3(n0.w, n1.w, n2.w);
            float3 pos[3], nrm[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                const float4 p = __ldg(vpos + 3 * (size_t)prim + k);
                const float4 n = __ldg(vnrm + 3 * (size_t)prim + k);
                pos[k] = f3(m0.x * p.x + m0.y * p.y + m0.z * p.z + m0.w, m1.x * p.x + m1.y * p.y + m1.z * p.z + m1.w,
                            m2.x * p.x + m2.y * p.y + m2.z * p.z + m2.w);
                nrm[k] = f3(n0.x * n.x + n0.y * n.y + n0.z * n.z, n1.x * n.x + n1.y * n.y + n1.z * n.z,
                            n2.x * n.x + n2.y * n.y + n2.z * n.z);
            }
            const float3 bary = f3(1.0f - hit.y - hit.z, hit.y, hit.z);               // :690
            const float3 rayDir = xyz(rd);                                          // :696
            const float3 position = xyz(ro) + rayDir * hit.x;                        // :692
            float3 SN = normalize(bary.x * nrm[0] + bary.y * nrm[1] + bary.z * nrm[2]);  // :693
            float3 N = normalize(cross(pos[1] - pos[0], pos[2] - pos[0]));           // :694
            if (dot(N, -rayDir) < 0.0f) {                                            // :698-707
                if (twofaced == 1u && emission.x == 0.0f && emission.y == 0.0f && emission.z == 0.0f) {
                    N = N * -1.0f;
                    SN = SN * -1.0f;
                }
            }
            const Onb onb = onbCreate(SN);                                           // :712
            const float3 wo = normalize(onbTransform(onb, -rayDir));                 // :713
            BSDFOutput bsdfRes;
            bsdfRes.bsdf = f3(0.0f); bsdfRes.pdf = 0.0f; bsdfRes.isDelta = false;
            float3 wi = f3(0.0f);
            sampleBSDFT<TYPE>(sc, handle & 0xffffu, rng, wo, wi, bsdfRes);            // :716
            const float NoW = fabsf(wi.z);                                           // :717
            wi = onbUntransform(onb, wi);                                            // :718

            bool haveLight;
            const LightOutput lightRes = sampleLight(sc, rng, position, haveLight);  // :720
            const float3 L = normalize(lightRes.position - position);                // :722
            const float3 wL = onbTransform(onb, L);                                  // :723
            const float Ldist = length(lightRes.position - position);                // :724
            const float NoL = fabsf(dot(SN, L));                                     // :725
            const float lightPdf = lightRes.pdf;                                     // :726
            BSDFOutput lightBsdfRes;
            lightBsdfRes.bsdf = f3(0.0f); lightBsdfRes.pdf = 0.0f; lightBsdfRes.isDelta = false;
            evalBSDFT<TYPE>(sc, handle & 0xffffu, wo, wL, lightBsdfRes);             // :729

            const bool transmissive = isTransmissionBSDF(TYPE);
            const float NdotV = dot(N, -rayDir);
            float3 nee = f3(0.0f);
            float dwVisible = 1.0f;
            if (sc.params.nee && haveLight && !bsdfRes.isDelta &&
                ((NdotV > 0.0f && dot(N, L) > 0.0f) || transmissive)) {              // :734-736
                needShadow = true;
                if (lightPdf != 0.0f) {                                              // :750-753, applied if the segment is free
                    const float w = powerHeuristic(lightPdf, bsdfRes.pdf);
                    nee = w * NoL * lightBsdfRes.bsdf * weight * lightRes.emission / lightPdf;
                    dwVisible = powerHeuristic(bsdfRes.pdf, lightPdf);               // :786
                }
            }
            // :760-768 surface emission
            const float lightFlag = NdotV > 0.0f ? 1.0f : 0.0f;
            const bool NEE = sc.params.nee != 0;
            const bool countEmitted = (flags & kFlagCountEmitted) != 0, wasDelta = (flags & kFlagWasDelta) != 0;
            float3 emitted = f3(0.0f);
            if (NEE && !countEmitted && !wasDelta) emitted = emitted + directWeight * emission * lightFlag * weight;
            if (!NEE || countEmitted || wasDelta) emitted = emitted + emission * lightFlag * weight;

            // :770-796 termination tests and next-ray setup
            bool done = false;
            if (dot(wi, N) <= 0.0f && !transmissive) done = true;
            else if (NdotV <= 0.0f && !transmissive) done = true;
            else if (!isvalidf(bsdfRes.pdf) || !isvalid3(bsdfRes.bsdf) || bsdfRes.pdf == 0.0f) done = true;
            float3 newO = xyz(ro), newD = rayDir;
            if (!done) {
                flags &= ~(kFlagCountEmitted | kFlagWasDelta);
                newO = position + 0.0001f * faceforward(N, -wi, N);
                newD = wi;
                weight = weight * (bsdfRes.bsdf * NoW / bsdfRes.pdf);
                if (bsdfRes.isDelta) flags |= kFlagWasDelta;
            }
            // raygen.rgen:59-80.  The Russian-roulette draw reads a COPY of the payload seed
            // (rngState = prd.seed is never written back), so the path's stream does not advance.
            uint32_t depth = flags & 0xffffu;
            if ((int)depth > sc.params.rr_start) {
                const float q = fminf(fmaxf(fmaxf(fmaxf(weight.x, weight.y), weight.z), 0.05f), 1.0f);
                uint32_t rr = rng;
                if (randUniform(rr) > q) terminate = true;
                weight = weight / q;
            }
            if ((int)depth > sc.params.max_depth) terminate = true;
            if (done) terminate = true;
            if (!terminate) flags = (flags & ~0xffffu) | (depth + 1u);

            if (needShadow) {
                shO = make_float4(position.x, position.y, position.z, Ldist - 0.01f);
                shD = make_float4(L.x, L.y, L.z, __uint_as_float(slot));
                shBase = make_float4(emitted.x, emitted.y, emitted.z, dwVisible);
                shNee = make_float4(nee.x, nee.y, nee.z, 0.0f);
            } else {
                const float cutoff = sc.params.cutoff;                               // raygen.rgen:60-63
                if (emitted.x < cutoff && emitted.y < cutoff && emitted.z < cutoff) {
                    res.x += emitted.x; res.y += emitted.y; res.z += emitted.z;
                }
            }
            res.w = __uint_as_float(flags);
            wf.result[slot] = res;
            wf.seed[slot] = rng;                                                     // rayhit.rchit:759
            if (!terminate) {
                wf.rayO[slot] = make_float4(newO.x, newO.y, newO.z, 0.0f);
                wf.rayD[slot] = make_float4(newD.x, newD.y, newD.z, 0.0f);
                wf.weightDW[slot] = make_float4(weight.x, weight.y, weight.z, 1.0f);  // directWeight 1 unless the shadow stage finds the light
            }
        }
        if (terminate) wf.sampleK[slot] |= kFinished;   // k_extend adds the sample to the image next iteration
        const uint32_t sp = queueSlot(qc, Q_SHADOW, needShadow);
        if (needShadow) {
            wf.shadowO[sp] = shO;
            wf.shadowD[sp] = shD;
            wf.shadowBase[sp] = shBase;
            wf.shadowNee[sp] = shNee;
        }
    }
}

// ---- shadow ---------------------------------------------------------------------------------------
template <bool STATS>
__global__ void __launch_bounds__(kBlock) k_shadow(SceneDev sc, WavefrontState wf, uint32_t parity) {
    const uint32_t count = wf.qCount[parity * Q_COUNT + Q_SHADOW];
    TraceStats ts = {0u, 0u, 0u};
    uint32_t traced = 0;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const float4 o = wf.shadowO[i], d = wf.shadowD[i];
        HitRecord hit;
        traceRay<true, STATS>(sc, o.x, o.y, o.z, d.x, d.y, d.z, 0.01f, o.w, hit, &ts, (uint32_t*)&wf.stats[ST_OVERFLOW]);
        traced++;
        const bool visible = hit.inst == GSB_NO_HIT;
        const uint32_t slot = __float_as_uint(d.w);
        const float4 base = wf.shadowBase[i];
        float3 emitted = f3(0.0f);
        if (visible) {
            const float4 nee = wf.shadowNee[i];
            emitted = emitted + xyz(nee);                                            // rayhit.rchit:752
            wf.weightDW[slot].w = base.w;                                            // :786 (neeDone)
        }
        emitted = emitted + xyz(base);                                               // :763-768
        const float cutoff = sc.params.cutoff;                                       // raygen.rgen:60-63
        if (emitted.x < cutoff && emitted.y < cutoff && emitted.z < cutoff) {
            float4 res = wf.result[slot];
            res.x += emitted.x; res.y += emitted.y; res.z += emitted.z;
            wf.result[slot] = res;
        }
    }
    if (STATS) flushTraceStats(wf.stats, ts, traced);
}

// one thread, once per render call: roll the last iteration's queue sizes into the totals
__global__ void k_foldCounters(WavefrontState wf, uint32_t parity, unsigned long long paths) {
    uint32_t* bank = wf.qCount + parity * Q_COUNT;
    unsigned long long ext = bank[Q_MISS];
    for (int q = 0; q < GSB_BSDF_TYPE_COUNT; q++) ext += bank[q];
    wf.stats[ST_EXT] += ext;
    wf.stats[ST_SHADOW] += bank[Q_SHADOW];
    wf.stats[ST_PATHS] += paths;
    for (int q = 0; q < 2 * Q_COUNT; q++) wf.qCount[q] = 0;
}

template <class T>
void ensure(T*& p, size_t count) {
    if (p) cudaFree(p);
    p = nullptr;
    void* q = nullptr;
    GSB_CUDA(cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
    p = (T*)q;
}

void allocWavefront(gsb_context* ctx, uint32_t slots) {
    WavefrontState& wf = ctx->wf;
    if (wf.capacity >= slots && wf.qCount) return;
    const size_t n = slots;
    ensure(wf.rayO, n); ensure(wf.rayD, n); ensure(wf.weightDW, n); ensure(wf.result, n);
    ensure(wf.seed, n); ensure(wf.sampleK, n); ensure(wf.hitTUVP, n); ensure(wf.hitInst, n);
    ensure(wf.matQueue, n * GSB_BSDF_TYPE_COUNT);
    ensure(wf.shadowO, n); ensure(wf.shadowD, n); ensure(wf.shadowBase, n); ensure(wf.shadowNee, n);
    if (!wf.qCount) ensure(wf.qCount, (size_t)2 * Q_COUNT);
    if (!wf.stats) {
        ensure(wf.stats, (size_t)ST_COUNT);
        GSB_CUDA(cudaMemsetAsync(wf.stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
    }
    wf.capacity = slots;
}

template <int TYPE>
void launchShade(gsb_context* ctx, const SceneDev& sc, int grid, uint32_t parity) {
    k_shade<TYPE><<<grid, kBlock, 0, ctx->stream>>>(sc, ctx->wf, parity);
}

}  // namespace

void freeWavefront(gsb_context* ctx) {
    WavefrontState& wf = ctx->wf;
    void* ptrs[] = {wf.rayO, wf.rayD, wf.weightDW, wf.result, wf.seed, wf.sampleK, wf.hitTUVP, wf.hitInst, wf.matQueue,
                    wf.shadowO, wf.shadowD, wf.shadowBase, wf.shadowNee, wf.qCount, wf.stats};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    wf = WavefrontState();
}

void renderImpl(gsb_context* ctx, uint32_t width, uint32_t height, uint32_t first, uint32_t count, uint32_t stride) {
    cudaStream_t s = ctx->stream;
    const uint32_t npix = width * height;
    if (ctx->width != width || ctx->height != height || !ctx->dAccum) {
        ensure(ctx->dAccum, (size_t)npix);
        GSB_CUDA(cudaMemsetAsync(ctx->dAccum, 0, (size_t)npix * sizeof(float4), s));
        ctx->width = width;
        ctx->height = height;
    }
    if (count == 0) return;
    // paths in flight: enough to fill the machine several times over, small enough that the path state stays near L2
    const unsigned long long totalWork = (unsigned long long)npix * count;
    const uint32_t slots = (uint32_t)std::min<unsigned long long>(totalWork, ctx->targetSlots);
    allocWavefront(ctx, slots);
    WavefrontState& wf = ctx->wf;
    RenderArgs ra;
    ra.width = width; ra.height = height; ra.npix = npix;
    ra.first = first; ra.count = count; ra.stride = stride;
    ra.slots = slots;
    ra.totalWork = totalWork;
    ra.accum = ctx->dAccum;
    ra.camZ = (std::fmax((float)width, (float)height) / 2.0f) / std::tan(ctx->fov / 2.0f);
    const SceneDev sc = ctx->sceneDev();
    GSB_CUDA(cudaMemsetAsync(wf.qCount, 0, 2 * Q_COUNT * sizeof(uint32_t), s));
    const unsigned long long liveInit[2] = {(unsigned long long)slots, 0ull};
    const unsigned long long nextInit = (unsigned long long)slots;
    GSB_CUDA(cudaMemcpyAsync(wf.stats, liveInit, sizeof(liveInit), cudaMemcpyHostToDevice, s));
    GSB_CUDA(cudaMemcpyAsync(&wf.stats[ST_NEXTWORK], &nextInit, sizeof(nextInit), cudaMemcpyHostToDevice, s));

    cudaEvent_t evStart = nullptr, evStop = nullptr;
    GSB_CUDA(cudaEventCreate(&evStart));
    GSB_CUDA(cudaEventCreate(&evStop));
    GSB_CUDA(cudaEventRecord(evStart, s));

    const int fullGrid = (int)((slots + kBlock - 1) / kBlock);
    int dev = ctx->device, sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int queueGrid = std::min(fullGrid, sms * 16);
    k_generate<<<fullGrid, kBlock, 0, s>>>(sc, wf, ra);
    ctx->launches += 1;
    ctx->lastRenderIters = 0;
    int typesUsed = 0;
    for (bool u : ctx->typeUsed) typesUsed += u ? 1 : 0;

    // stage timing (profiling mode): events around every stage of every iteration
    const bool prof = ctx->profiling;
    std::vector<cudaEvent_t> pev;
    auto mark = [&]() {
        if (!prof) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, s);
        pev.push_back(e);
    };

    uint32_t parity = 0;
    auto iteration = [&]() {
        mark();
        if (ctx->instrumented & 1) k_extend<true><<<fullGrid, kBlock, 0, s>>>(sc, wf, ra, parity);
        else k_extend<false><<<fullGrid, kBlock, 0, s>>>(sc, wf, ra, parity);
        mark();
        if (ctx->typeUsed[0]) launchShade<0>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[1]) launchShade<1>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[2]) launchShade<2>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[3]) launchShade<3>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[4]) launchShade<4>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[5]) launchShade<5>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[6]) launchShade<6>(ctx, sc, queueGrid, parity);
        if (ctx->typeUsed[7]) launchShade<7>(ctx, sc, queueGrid, parity);
        mark();
        if (ctx->instrumented & 2) k_shadow<true><<<queueGrid, kBlock, 0, s>>>(sc, wf, parity);
        else k_shadow<false><<<queueGrid, kBlock, 0, s>>>(sc, wf, parity);
        mark();
        mark();
        parity ^= 1u;
        ctx->launches += 2 + (uint64_t)typesUsed;
        ctx->lastRenderIters++;
    };

    // Host loop: batches of iterations; the live-slot count of batch b is read back while batch
    // b+1 is already queued, so the device never waits for the host.
    const int batch = 8;
    unsigned long long* hLive = nullptr;
    GSB_CUDA(cudaMallocHost(&hLive, 2 * sizeof(unsigned long long)));
    cudaEvent_t evBatch[2];
    GSB_CUDA(cudaEventCreate(&evBatch[0]));
    GSB_CUDA(cudaEventCreate(&evBatch[1]));
    int b = 0;
    bool finished = false;
    while (!finished) {
        for (int i = 0; i < batch; i++) iteration();
        GSB_CUDA(cudaMemcpyAsync(&hLive[b & 1], &wf.stats[ST_LIVE], sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
        GSB_CUDA(cudaEventRecord(evBatch[b & 1], s));
        if (b > 0) {
            GSB_CUDA(cudaEventSynchronize(evBatch[(b - 1) & 1]));
            if (hLive[(b - 1) & 1] == 0) finished = true;
        }
        b++;
        if (prof && pev.size() > 200000) break;
    }
    k_foldCounters<<<1, 1, 0, s>>>(wf, parity ^ 1u, totalWork);
    ctx->launches += 1;
    GSB_CUDA(cudaEventRecord(evStop, s));
    GSB_CUDA(cudaEventSynchronize(evStop));
    GSB_CUDA(cudaGetLastError());
    float ms = 0.0f;
    GSB_CUDA(cudaEventElapsedTime(&ms, evStart, evStop));
    ctx->lastRenderMs[0] = ms;
    for (int k = 1; k < 5; k++) ctx->lastRenderMs[k] = 0.0;
    if (prof) {
        for (size_t i = 0; i + 4 < pev.size(); i += 5) {
            for (int k = 0; k < 4; k++) {
                float t = 0.0f;
                cudaEventElapsedTime(&t, pev[i + k], pev[i + k + 1]);
                ctx->lastRenderMs[1 + k] += t;
            }
        }
        for (cudaEvent_t e : pev) cudaEventDestroy(e);
    }
    cudaEventDestroy(evBatch[0]);
    cudaEventDestroy(evBatch[1]);
    cudaEventDestroy(evStart);
    cudaEventDestroy(evStop);
    cudaFreeHost(hLive);
}

}  // namespace gsb

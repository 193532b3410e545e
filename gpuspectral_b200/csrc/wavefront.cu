// wavefront.cu — the wavefront path tracer.
//
// The reference runs one megakernel-by-recursion per pixel (raygen.rgen:29-109 calling
// rayhit.rchit:666-797 through traceRayEXT).  Here the same per-path state machine is cut
// into stages that each run as one wide kernel over many paths:
//
//   k_regen    for slots whose path has ended: adds the sample to the image (raygen.rgen:84-108) and starts the slot's
//              next pixel-sample (raygen.rgen:31-48), all lanes busy (ended slots are compacted per block)
//   [sort]     optional: slot indices ordered by (origin cell, direction octant) so that a warp's rays walk the same
//              part of the BVH (cub radix sort over keys written by k_regen / k_shade)
//   k_extend   closest-hit traversal of every live path's ray (raygen.rgen:54-58); a miss (miss.rmiss) ends the path
//   k_shade<T> material-sorted closest-hit shading             (rayhit.rchit:666-797) + the
//              per-bounce bookkeeping of the raygen loop       (raygen.rgen:59-80)
//   k_shadow   any-hit traversal of the NEE segments           (rayhit.rchit:737-757)
// Four launches per wavefront iteration (for one BSDF type in use); queue counters are double-banked by
// iteration parity so that no separate reset kernel is needed.
//
// Slots are not tied to pixels: a slot whose path ends takes the next pixel-sample from a global work counter
// (work item w = sample w / npix of pixel w % npix), so every slot stays busy until the call's work is exhausted and
// only the last few iterations of a call run with a thinning wavefront.  Finished samples are added to the image with
// one 128-bit atomic add (rgb, count).  Every path carries its own RNG word; each stage consumes draws in the order
// the shader does, which makes per-path results match the checker's.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <cstdio>
#include <vector>

#include "bsdf.cuh"
#include "context.h"
#include "traverse.cuh"

namespace gsb {

namespace {

constexpr int kBlock = kTraceBlock;
#ifndef GSB_TRACE_MIN_BLOCKS
#define GSB_TRACE_MIN_BLOCKS 8   // resident blocks per SM the traversal kernels are compiled for
#endif
#ifndef GSB_SHADE_MIN_BLOCKS
#define GSB_SHADE_MIN_BLOCKS 8   // resident blocks per SM the shading kernels are compiled for (register cap 64)
#endif
#ifndef GSB_SHADE_MIN_BLOCKS_DIFFUSE
#define GSB_SHADE_MIN_BLOCKS_DIFFUSE 6   // ... the diffuse kernel: 80 registers, no spills (r2i: 66.6 -> 61.5 ms per Cornell step with the state pipeline)
#endif
constexpr uint32_t kRetired = 0xffffffffu;   // result.w (where the pixel index lives): the slot has no path and will get none (the call's work is used up)
// hitInst of a slot whose path has not been traced yet (k_generate): any word that is neither GSB_NO_HIT (= path ended)
// nor carries a valid BSDF type in bits 28..31
constexpr uint32_t kInstFresh = 0xfffffffeu;
// hitInst of a slot whose path left the scene and whose sample k_extend has already added to the image (GSB_MISS_DELIVERS):
// k_regen only has to start the slot's next path.  Like kInstFresh it decodes to no BSDF type.
constexpr uint32_t kInstDelivered = 0xfffffffdu;
// hitInst of a slot that has no path and will get none (the call's work is used up): k_regen tells it from an ended path
// by this word alone (it used to read result.w of every GSB_NO_HIT slot for that: a second, dependent DRAM round trip)
constexpr uint32_t kInstRetired = 0xfffffffcu;
#ifndef GSB_MISS_DELIVERS
#define GSB_MISS_DELIVERS 1
#endif
#ifndef GSB_SHADE_PREFETCH
#define GSB_SHADE_PREFETCH 0  // k_shade prefetches its next tile's state: 1 into L1, 2 into L2
#endif
#ifndef GSB_SHADE_STREAM
#define GSB_SHADE_STREAM 0    // 1: k_shade reads the per-slot state with ld.global.cs
#endif
#ifndef GSB_LAZY_RESULT
#define GSB_LAZY_RESULT 0     // 1: k_extend loads a slot's radiance only when its path has ended (a dependent load)
#endif
constexpr uint32_t kFlagWasDelta = 1u << 16;
constexpr uint32_t kFlagCountEmitted = 1u << 17;
constexpr uint32_t kFlagDead = 1u << 18;        // direction.w of a retired slot: k_extend skips it
#ifndef GSB_RAY_AOS
#define GSB_RAY_AOS 1
#endif
GSB_D float4* rayOrigin(const WavefrontState& wf, uint32_t slot) { return GSB_RAY_AOS ? wf.ray + 2 * (size_t)slot : wf.ray + slot; }
GSB_D float4* rayDirection(const WavefrontState& wf, uint32_t slot) { return GSB_RAY_AOS ? wf.ray + 2 * (size_t)slot + 1 : wf.ray + (size_t)wf.capacity + slot; }
constexpr uint32_t kShadowHasBase = 1u << 31;   // shadowD.w = path slot | this bit when shadowBase holds surface emission

struct RenderArgs {
    uint32_t width, height, npix;
    uint32_t first, count, stride;  // timestamps first + k*stride, k < count
    uint32_t slots;                 // paths in flight
    double invNpix;                 // 1 / npix
    uint32_t refillMin;             // idle lanes that trigger a refill of a persistent traversal warp
    float camZ;                     // (max(W,H)/2) / tan(fov/2) of rayDir(), raygen.rgen:20-25 (evaluated once on the host)
    unsigned long long totalWork;   // npix * count pixel-samples
    float4* accum;                  // the image: rgb = sum of samples, a = samples added
    const uint32_t* order;          // ray sorting: slot of the i-th ray k_extend takes (nullptr: slot order)
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

// Sort key of a ray: Morton code of the origin's cell in a (2^sortBits)^3 grid over the scene bounds, then the direction
// octant — rays that start in the same region and head the same way are neighbours in the sorted order, so the lanes of a
// traversal warp fetch the same nodes and stay in the same phase for longer.
GSB_D uint32_t spreadBits3(uint32_t v) {   // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000ffu;
    v = (v | (v << 8)) & 0x0300f00fu;
    v = (v | (v << 4)) & 0x030c30c3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
GSB_D uint32_t sortKey(const SceneDev& sc, float3 o, float3 d) {
    const float top = (float)((1u << sc.sortBits) - 1u);
    const uint32_t cx = (uint32_t)fminf(fmaxf((o.x - sc.sortLo[0]) * sc.sortScale[0], 0.0f), top);
    const uint32_t cy = (uint32_t)fminf(fmaxf((o.y - sc.sortLo[1]) * sc.sortScale[1], 0.0f), top);
    const uint32_t cz = (uint32_t)fminf(fmaxf((o.z - sc.sortLo[2]) * sc.sortScale[2], 0.0f), top);
    const uint32_t oct = (d.x < 0.0f ? 0u : 4u) | (d.y < 0.0f ? 0u : 2u) | (d.z < 0.0f ? 0u : 1u);
    const uint32_t cell = spreadBits3(cx) | (spreadBits3(cy) << 1) | (spreadBits3(cz) << 2);
    if (sc.sortMode == 1u) return (oct << (3u * sc.sortBits)) | cell;      // octant major
    if (sc.sortMode == 2u) {
        // 2 bits of direction per axis below the cell
        const uint32_t qx = (uint32_t)fminf(fmaxf((d.x * 0.5f + 0.5f) * 4.0f, 0.0f), 3.0f), qy = (uint32_t)fminf(fmaxf((d.y * 0.5f + 0.5f) * 4.0f, 0.0f), 3.0f),
                       qz = (uint32_t)fminf(fmaxf((d.z * 0.5f + 0.5f) * 4.0f, 0.0f), 3.0f);
        return (cell << 6) | spreadBits3(qx) | (spreadBits3(qy) << 1) | (spreadBits3(qz) << 2);
    }
    return (cell << 3) | oct;
}

// Offset of the primary ray from the pixel's integer coordinate for one uniform draw u (gsb_params::pixel_filter):
// box = u - 1/2; tent of radius 1 by inverting its CDF.  Individually rounded operations: the checker computes the same bits.
GSB_D float filterOffset(int filter, float u) {
    if (filter == 1) return __fsub_rn(u, 0.5f);
    return u < 0.5f ? __fsub_rn(__fsqrt_rn(__fmul_rn(2.0f, u)), 1.0f) : __fsub_rn(1.0f, __fsqrt_rn(__fsub_rn(2.0f, __fmul_rn(2.0f, u))));
}

// raygen.rgen:20-48: seed + primary ray + payload initialisation for work item `work` (sample work / npix of
// pixel work % npix); the new ray is also returned in registers
GSB_D void startPath(const SceneDev& sc, const WavefrontState& wf, const RenderArgs& ra, uint32_t slot, unsigned long long work,
                     float4& o, float4& dOut, uint32_t& pixelOut) {
    // k = work / npix without a 64-bit integer division: double quotient (exact to 2^53) and one correction step
    uint32_t k = (uint32_t)__double2ull_rz(__ull2double_rn(work) * ra.invNpix);
    long long rem = (long long)(work - (unsigned long long)k * ra.npix);
    if (rem < 0) { k--; rem += ra.npix; }
    else if (rem >= (long long)ra.npix) { k++; rem -= ra.npix; }
    const uint32_t pixel = (uint32_t)rem;
    const uint32_t x = pixel % ra.width, y = pixel / ra.width;
    const uint32_t timestamp = ra.first + k * ra.stride;
    // rayDir(size, fragCoord, fov)
    const float sx = (float)ra.width, sy = (float)ra.height;
    uint32_t rng = pcgHash(tea(ra.width * y + x, timestamp));                        // raygen.rgen:37
    float px = (float)x, py = (float)y;
    if (sc.params.pixel_filter) {
        // extension (gsb_params::pixel_filter): the sample's first two draws place the ray inside the pixel's footprint
        const float u0 = randUniform(rng), u1 = randUniform(rng);
        px += filterOffset(sc.params.pixel_filter, u0);
        py += filterOffset(sc.params.pixel_filter, u1);
    }
    const float fx = px - sx / 2.0f;
    const float fy = py - sy / 2.0f;
    const float3 dir = normalize(f3(-fx, fy, ra.camZ));
    const float* V = sc.view;
    float3 d = f3(V[0] * dir.x + V[4] * dir.y + V[8] * dir.z, V[1] * dir.x + V[5] * dir.y + V[9] * dir.z,
                  V[2] * dir.x + V[6] * dir.y + V[10] * dir.z);
    d.y *= -1.0f;
    // .w: the path's RNG state (raygen.rgen:37)
    o = make_float4(sc.eye[0], sc.eye[1], sc.eye[2], __uint_as_float(rng));
    dOut = make_float4(d.x, d.y, d.z, __uint_as_float(kFlagCountEmitted));   // .w: depth 0, wasDelta 0, countEmitted 1
    *rayOrigin(wf, slot) = o;
    *rayDirection(wf, slot) = dOut;
    // weight and directWeight start out as 1: k_shade knows that from depth 0 and does not read them (nothing is written here)
    wf.result[slot] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(pixel));   // .w: the pixel the sample belongs to
    if (wf.sortKey) wf.sortKey[slot] = sortKey(sc, xyz(o), d);
    pixelOut = pixel;
}

__global__ void k_generate(SceneDev sc, WavefrontState wf, RenderArgs ra) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= ra.slots) return;
    float4 o, d;
    uint32_t pixel;
    startPath(sc, wf, ra, slot, slot, o, d, pixel);   // slots <= totalWork: the first `slots` items start out in slot order
    wf.hitInst[slot] = kInstFresh;   // not traced yet: k_regen of the first iteration must not take it for an ended path
}

// ---- regenerate ------------------------------------------------------------------------------------
// raygen.rgen:84-108 + :31-48 for every slot whose path ended in the last iteration (hitInst == GSB_NO_HIT: a miss,
// or k_shade ended it).  One block owns kRegenSpan consecutive slots: it counts its ended slots, reserves that many
// consecutive work items with ONE atomic (work item w = sample w / npix of pixel w % npix, so a block's new paths are
// neighbouring pixels of one sample), and starts them.  A slot that finds no work left is retired.
#ifndef GSB_REGEN_PER
#define GSB_REGEN_PER 8
#endif
constexpr int kRegenBlock = 256, kRegenPer = GSB_REGEN_PER, kRegenSpan = kRegenBlock * kRegenPer;
static_assert(kRegenPer % 4 == 0, "k_regen reads the instance words four at a time");
// slot k (0 .. kRegenPer - 1) of thread tid in the block that starts at `base`
GSB_D uint32_t regenSlot(uint32_t base, uint32_t tid, int k) { return base + (uint32_t)(k >> 2) * (4u * kRegenBlock) + 4u * tid + (uint32_t)(k & 3); }
__global__ void __launch_bounds__(kRegenBlock, 8) k_regen(SceneDev sc, WavefrontState wf, RenderArgs ra) {
    __shared__ uint32_t s_warp[kRegenBlock / 32];
    __shared__ unsigned long long s_base;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t base = blockIdx.x * (uint32_t)kRegenSpan;
    uint32_t ended = 0, delivered = 0;   // bit k: slot base + k * kRegenBlock + tid holds an ended path / one whose sample k_extend delivered
    // The scan is the kernel's floor (64 MB of instance words per launch at 16 M slots): 16-byte loads, all of a thread's
    // in flight before the first word is looked at.  Thread tid owns slots base + 4 tid + 1024 j + {0..3}; slot k of the
    // masks is j * 4 + c.
    uint32_t hi[kRegenPer];
#pragma unroll
    for (int j = 0; j < kRegenPer / 4; j++) {
        const uint32_t slot = regenSlot(base, tid, 4 * j);
        uint4 w = make_uint4(kInstRetired, kInstRetired, kInstRetired, kInstRetired);
        if (slot + 3u < ra.slots) w = __ldcs(reinterpret_cast<const uint4*>(wf.hitInst + slot));
        else {
            if (slot < ra.slots) w.x = wf.hitInst[slot];
            if (slot + 1u < ra.slots) w.y = wf.hitInst[slot + 1u];
            if (slot + 2u < ra.slots) w.z = wf.hitInst[slot + 2u];
        }
        hi[4 * j] = w.x; hi[4 * j + 1] = w.y; hi[4 * j + 2] = w.z; hi[4 * j + 3] = w.w;
    }
#pragma unroll
    for (int k = 0; k < kRegenPer; k++) {
        if (hi[k] == kInstDelivered) { ended |= 1u << k; delivered |= 1u << k; }
        else if (hi[k] == GSB_NO_HIT) ended |= 1u << k;
    }
    // exclusive prefix of the ended counts over the block
    const uint32_t n = (uint32_t)__popc(ended);
    uint32_t incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += v;
    }
    if (lane == 31u) s_warp[warp] = incl;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kRegenBlock / 32; w++) {
        const uint32_t c = s_warp[w];
        if ((uint32_t)w < warp) before += c;
        total += c;
    }
    if (total == 0u) return;
    // the block's work items are reserved now and the answer is looked at after the list is built and the first samples
    // are requested: the atomic's round trip overlaps both
    unsigned long long reserved = 0ull;
    if (tid == 0u) reserved = atomicAdd(&wf.stats[ST_NEXTWORK], (unsigned long long)total);
    // the block's ended slots, compacted: entry i of the list gets work item s_base + i, and the list is then processed by
    // all threads side by side (r2q: with every thread working through its own eight slots the path start — hash, camera
    // ray — ran with 5.6 of 32 lanes, 1572 instructions per warp; it was 103 M of the kernel's instructions)
    __shared__ uint32_t s_list[kRegenSpan];
    {
        uint32_t at = before + (incl - n);
#pragma unroll
        for (int k = 0; k < kRegenPer; k++)
            if (ended & (1u << k)) s_list[at++] = regenSlot(base, tid, k) | (delivered & (1u << k) ? 0x80000000u : 0u);
    }
    __syncthreads();
    // A thread's first entry and that path's sample are requested BEFORE the reservation's answer is waited for: the
    // atomic's round trip (thread 0, then the barrier everybody waits at) and the sample's trip from DRAM overlap instead
    // of following one another (r2w ncu: 16 % of the stall samples on the sample load, 10 per issue at the barrier).
    uint32_t firstEntry = 0x80000000u;
    float4 firstRes = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (tid < total) {
        firstEntry = s_list[tid];
        if (!(firstEntry & 0x80000000u)) firstRes = wf.result[firstEntry & 0x7fffffffu];
    }
    asm volatile("" ::: "memory");
    if (tid == 0u) s_base = reserved;
    __syncthreads();
    uint32_t retired = 0;
    for (uint32_t i = tid; i < total; i += kRegenBlock) {
        const uint32_t entry = i == tid ? firstEntry : s_list[i], slot = entry & 0x7fffffffu;
        const unsigned long long work = s_base + i;
        if (!(entry & 0x80000000u)) {      // (a path that left the scene was delivered by k_extend)
            const float4 res = i == tid ? firstRes : wf.result[slot];
            const uint32_t pixel = __float_as_uint(res.w);
            // raygen.rgen:106-108: a NaN sample leaves the image untouched
            if (!(isnan(res.x) || isnan(res.y) || isnan(res.z))) atomicAdd(&ra.accum[pixel], make_float4(res.x, res.y, res.z, 1.0f));
        }
        if (work < ra.totalWork) {
            float4 o, d;
            uint32_t p;
            startPath(sc, wf, ra, slot, work, o, d, p);
        } else {
            wf.result[slot] = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kRetired));
            wf.hitInst[slot] = kInstRetired;   // neither an ended path nor one with a BSDF type: nobody picks the slot up again
            *rayDirection(wf, slot) = make_float4(0.0f, 0.0f, 0.0f, __uint_as_float(kFlagDead));
            if (wf.sortKey) wf.sortKey[slot] = 0xffffffffu;   // sorts behind every live ray
            retired++;
        }
    }
    // live-slot count: one atomic per warp that retired something
    for (int o = 16; o > 0; o >>= 1) retired += __shfl_xor_sync(0xffffffffu, retired, o);
    if (lane == 0u && retired) atomicAdd(&wf.stats[ST_LIVE], (unsigned long long)(-(long long)retired));
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

// The ray stream of k_extend: the i-th ray is the one of slot order[i] (ray sorting) or of slot i.  Two 128-bit loads per
// ray in, the instance word (and the hit record on a hit) out; a retired slot carries kFlagDead and is skipped.
struct ExtendSource {
    const SceneDev& sc;
    const WavefrontState& wf;
    const RenderArgs& ra;
    uint32_t* qc;       // this iteration's counter bank
    WorkRange range;
    uint32_t traced = 0;

    GSB_D ExtendSource(const SceneDev& s, const WavefrontState& w, const RenderArgs& r, uint32_t* q) : sc(s), wf(w), ra(r), qc(q) {}
    GSB_D bool drained() const { return range.exhausted; }

    GSB_D bool fetch(bool want, RayJob& job) {
        const uint32_t idx = range.take(&qc[Q_NEXT_EXTEND], ra.slots, want);
        const uint32_t ahead = range.peek(ra.slots);
        uint32_t slot = idx;
        if (ra.order) {
            if (idx != kNoTag) slot = ra.order[idx];
            if (ahead != kNoTag) prefetchL1(ra.order + ahead);
        } else if (ahead != kNoTag) {
            // the state words of the slots this warp will take next: into L1 while the rays just fetched are traced
            prefetchL1(rayOrigin(wf, ahead));
            if (!GSB_RAY_AOS) prefetchL1(rayDirection(wf, ahead));
        }
        if (idx == kNoTag) return false;
        const float4 o = *rayOrigin(wf, slot), d = *rayDirection(wf, slot);
        if (__float_as_uint(d.w) & kFlagDead) return false;
#if GSB_MISS_DELIVERS
        prefetchL2(wf.result + slot);   // should the ray leave the scene, its sample is read at hand-in: have it in L2 by then
#endif
        job.ox = o.x; job.oy = o.y; job.oz = o.z;
        job.dx = d.x; job.dy = d.y; job.dz = d.z;
        job.tmin = 0.0f; job.tmax = 1e10f;                              // raygen.rgen:54-58
        job.tag = slot;
        job.aux = 0u;
        traced++;
        return true;
    }

    // stores only: k_shade finds its work by scanning hitInst in slot order; GSB_NO_HIT (miss.rmiss: done = 1, nothing is
    // added) is also what k_regen reads as "this path has ended"
    GSB_D void commit(bool have, const HitRecord& hit, uint32_t slot, uint32_t) {
        if (!have) return;
#if GSB_MISS_DELIVERS
        if (hit.inst == GSB_NO_HIT) {
            // miss.rmiss: the path has ended, and nothing is pending for it (its last shadow ray was resolved before this
            // launch): the sample goes to the image right here (raygen.rgen:84-108) — k_regen then only starts the slot's
            // next path and touches neither the sample nor the image.  With an environment emitter its radiance joins first.
            float4 res = wf.result[slot];
            if (sc.env.type) {
                const float4 rd = *rayDirection(wf, slot);
                const uint32_t flags = __float_as_uint(rd.w);
                float4 wdw = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
                if ((flags & 0xffffu) != 0u) wdw = wf.weightDW[slot];
                const bool plain = !sc.params.nee || (flags & kFlagCountEmitted) || (flags & kFlagWasDelta);
                const float3 e = envRadiance(sc.env, xyz(rd)) * xyz(wdw) * (plain ? 1.0f : wdw.w);
                const float cutoff = sc.params.cutoff;
                if ((e.x != 0.0f || e.y != 0.0f || e.z != 0.0f) && e.x < cutoff && e.y < cutoff && e.z < cutoff) { res.x += e.x; res.y += e.y; res.z += e.z; }
            }
            if (!(isnan(res.x) || isnan(res.y) || isnan(res.z))) atomicAdd(&ra.accum[__float_as_uint(res.w)], make_float4(res.x, res.y, res.z, 1.0f));
            wf.hitInst[slot] = kInstDelivered;
            return;
        }
#endif
        wf.hitInst[slot] = hit.inst;
        if (hit.inst == GSB_NO_HIT && sc.env.type) {
            // the ray left the scene: radiance of the environment emitter (extension), weighted like a surface emitter
            // that a BSDF sample reaches (rayhit.rchit:760-768) and subject to the per-bounce cutoff (raygen.rgen:60-63)
            const float4 rd = *rayDirection(wf, slot);
            const uint32_t flags = __float_as_uint(rd.w);
            float4 wdw = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
            if ((flags & 0xffffu) != 0u) wdw = wf.weightDW[slot];
            const bool plain = !sc.params.nee || (flags & kFlagCountEmitted) || (flags & kFlagWasDelta);
            const float3 e = envRadiance(sc.env, xyz(rd)) * xyz(wdw) * (plain ? 1.0f : wdw.w);
            const float cutoff = sc.params.cutoff;
            if ((e.x != 0.0f || e.y != 0.0f || e.z != 0.0f) && e.x < cutoff && e.y < cutoff && e.z < cutoff) {
                float4 res = wf.result[slot];
                res.x += e.x; res.y += e.y; res.z += e.z;
                wf.result[slot] = res;
            }
        }
        // a baked hit is shaded from the world level's own records: it carries its leaf position instead of gl_PrimitiveID
        if (hit.inst != GSB_NO_HIT) wf.hitTUVP[slot] = make_float4(hit.t, hit.u, hit.v, __uint_as_float((hit.inst & kInstBaked) ? hit.leaf : hit.prim));
    }
};

template <bool STATS, bool HASINST, bool POOL>
__global__ void __launch_bounds__(kBlock, GSB_TRACE_MIN_BLOCKS) k_extend(SceneDev sc, WavefrontState wf, RenderArgs ra, uint32_t parity) {
    uint32_t* qc = wf.qCount + parity * Q_COUNT;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        // roll the previous iteration's queue sizes into the totals and clear that bank for the next iteration
        uint32_t* prev = wf.qCount + (parity ^ 1u) * Q_COUNT;
        unsigned long long ext = prev[Q_TRACED];
        for (int q = 0; q < GSB_BSDF_TYPE_COUNT; q++) ext += prev[q];
        wf.stats[ST_EXT] += ext;
        wf.stats[ST_SHADOW] += prev[Q_SHADOW];
        wf.stats[ST_ITERS] += 1;
        for (int q = 0; q < Q_COUNT; q++) prev[q] = 0;
    }
    ExtendSource src(sc, wf, ra, qc);
    TraceStats ts = {0u, 0u, 0u};
    traceStream<false, STATS, HASINST, POOL>(sc, src, ra.refillMin, &ts, (uint32_t*)&wf.stats[ST_OVERFLOW]);
    // extension rays traced by this warp -> this iteration's counter bank
    uint32_t traced = src.traced;
    for (int o = 16; o > 0; o >>= 1) traced += __shfl_xor_sync(0xffffffffu, traced, o);
    if ((threadIdx.x & 31u) == 0u && traced) atomicAdd(&qc[Q_TRACED], traced);
    if (STATS) flushTraceStats(wf.stats, ts, src.traced);
}

// ---- shade ----------------------------------------------------------------------------------------
GSB_D void cpAsync16(void* smem, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
GSB_D void cpAsync4(void* smem, const void* gmem) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem) : "memory");
}
GSB_D void cpAsyncCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
GSB_D void cpAsyncWait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Material sorting without a global queue: every warp scans the slots 32 at a time in slot order and keeps those whose
// hit carries a BSDF of this kernel's TYPE.  A tile with at least directMin matching lanes is shaded as it is — the
// path-state streams are then read and written coalesced; sparser tiles are compacted through a 64-entry per-warp
// buffer in shared memory and shaded 32 at a time (entries wait there across in-place tiles).
// EXT: the scene uses an extension that shading has to know about (reflectance textures, environment emitter); the plain
// instantiation carries none of that code (r2m: the texture branch alone cost the diffuse kernel 11 % when compiled in).
#ifndef GSB_SHADE_DEFER_SHADOW
#define GSB_SHADE_DEFER_SHADOW 1
#endif
template <int TYPE, bool HASINST, bool EXT>
__global__ void __launch_bounds__(kBlock, TYPE == 0 ? GSB_SHADE_MIN_BLOCKS_DIFFUSE : GSB_SHADE_MIN_BLOCKS) k_shade(SceneDev sc, WavefrontState wf, uint32_t parity, uint32_t slots, uint32_t directMin) {
    uint32_t* qc = wf.qCount + parity * Q_COUNT;
    __shared__ uint32_t s_buf[kBlock / 32][64];
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t* buf = s_buf[threadIdx.x >> 5];
    const uint32_t warpsTotal = gridDim.x * (kBlock / 32), warpId = blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5);
    const uint32_t tiles = (slots + 31u) / 32u;
    uint32_t pendingCount = 0;
    uint32_t shaded = 0;   // warp-uniform: vertices this warp shaded (gsb_counters::shaded_vertices)
    // Software pipeline over batches of 32 vertices: the per-slot state of batch k (ray origin / direction, hit record,
    // weight, instance word: 68 bytes per vertex) is fetched with cp.async into one of two shared-memory stages while batch
    // k - 1 is shaded out of the other one, so the DRAM latency of the state streams overlaps the arithmetic and the
    // dependent scene-record loads of the previous batch instead of stalling every warp at the top of its batch (r2h ncu:
    // long-scoreboard 8.4 cycles per issue, 3.8 TB/s of DRAM with ~1/3 of the resident threads waiting on those loads).
    __shared__ float4 s_state[kBlock / 32][2][4][32];     // [warp][stage][origin, direction, hit, weight][lane]
    __shared__ uint32_t s_inst[kBlock / 32][2][32];
    float4 (*st)[4][32] = s_state[threadIdx.x >> 5];
    uint32_t (*sti)[32] = s_inst[threadIdx.x >> 5];
    // Deferred shadow-queue writes (below) for the diffuse kernel of world-only scenes: there it buys 4.5 % of the kernel
    // (Cornell); on the field and on the gallery's 8-blocks-per-SM kernels the 2 KB per warp it takes from L1 cost more
    // than the hidden atomic gives back (r2x: field shade 14.0 -> 14.7 ms with it on everywhere).
    constexpr bool kDefer = GSB_SHADE_DEFER_SHADOW && TYPE == 0 && !HASINST;
    __shared__ float4 s_shadow[kDefer ? kBlock / 32 : 1][kDefer ? 4 : 1][kDefer ? 32 : 1];   // [warp][origin, direction, nee, base][lane]
    float4 (*sh)[kDefer ? 32 : 1] = s_shadow[kDefer ? (threadIdx.x >> 5) : 0];
    uint32_t pendMask = 0u, pendBase = 0u, pendRank = 0u;
    bool pendNeed = false, pendHasBase = false;
    auto flushShadow = [&]() {
        if constexpr (kDefer) {
            if (pendMask) {                               // warp-uniform
                const uint32_t sp = __shfl_sync(0xffffffffu, pendBase, 0) + pendRank;
                if (pendNeed) {
                    wf.shadowO[sp] = sh[0][lane];
                    wf.shadowD[sp] = sh[1][lane];
                    wf.shadowNee[sp] = sh[2][lane];
                    if (pendHasBase) wf.shadowBase[sp] = sh[3][lane];
                }
                pendMask = 0u;
            }
        }
    };
    uint32_t tile = warpId;
    // the instance word of the tile scanned next is loaded one scan ahead (its DRAM latency hides behind the batch in between)
    uint32_t aheadWord = (tile < tiles && tile * 32u + lane < slots) ? wf.hitInst[tile * 32u + lane] : GSB_NO_HIT;
    bool scanDone = false, havePrev = false, prevActive = false;
    uint32_t prevSlot = 0, stage = 0;
    for (;;) {
        // ---- scan until the next batch of vertices is known (or the slots are used up)
        bool haveNew = false, newActive = false;
        uint32_t newSlot = 0;
        while (!scanDone) {
            const bool scanning = tile < tiles;
            const uint32_t cand = tile * 32u + lane;
            tile += warpsTotal;
            const uint32_t word = aheadWord;
            // the instance word carries the BSDF type; GSB_NO_HIT decodes to none
            const bool match = scanning && cand < slots && (word >> kInstTypeShift) == (uint32_t)TYPE;
            const uint32_t mm = __ballot_sync(0xffffffffu, match);
            // The next tile's word is requested only now, AFTER this tile's word has been used: issued before the vote (r2w
            // ncu, SASS: LDG three instructions ahead of the ISETP on `word`) the test waited on the scoreboard the new load
            // had just taken - a full DRAM round trip at the head of every batch, 18.7 % of the kernel's stall samples.
            asm volatile("" ::: "memory");
            aheadWord = (tile < tiles && tile * 32u + lane < slots) ? wf.hitInst[tile * 32u + lane] : GSB_NO_HIT;
            if ((uint32_t)__popc(mm) >= directMin) {
                // a tile that matches (almost) completely is shaded in place: coalesced state streams and no trip through
                // the compaction buffer, at the price of its few non-matching lanes idling
                newActive = match;
                newSlot = cand;
                haveNew = true;
                break;
            }
            if (match) buf[pendingCount + (uint32_t)__popc(mm & ((1u << lane) - 1u))] = cand;
            pendingCount += (uint32_t)__popc(mm);
            __syncwarp();
            if (pendingCount >= 32u || (!scanning && pendingCount > 0u)) {
                const uint32_t n = pendingCount < 32u ? pendingCount : 32u;
                newActive = lane < n;
                if (newActive) newSlot = buf[pendingCount - n + lane];
                pendingCount -= n;
                __syncwarp();
                haveNew = true;
                if (!scanning && pendingCount == 0u) scanDone = true;
                break;
            }
            if (!scanning) scanDone = true;
        }
        // ---- its state starts its trip from DRAM
        if (haveNew) {
            if (newActive) {
                cpAsync16(&st[stage][0][lane], rayOrigin(wf, newSlot));
                cpAsync16(&st[stage][1][lane], rayDirection(wf, newSlot));
                cpAsync16(&st[stage][2][lane], wf.hitTUVP + newSlot);
                cpAsync16(&st[stage][3][lane], wf.weightDW + newSlot);
                cpAsync4(&sti[stage][lane], wf.hitInst + newSlot);
            }
            cpAsyncCommit();
        }
        if (!havePrev) {
            if (!haveNew) break;
            havePrev = true; prevActive = newActive; prevSlot = newSlot; stage ^= 1u;
            continue;
        }
        // ---- shade the previous batch out of the other stage
        if (haveNew) cpAsyncWait<1>(); else cpAsyncWait<0>();
        __syncwarp();
        const uint32_t rs = stage ^ 1u;
        const bool active = prevActive;
        const uint32_t slot = prevSlot;
        bool terminate = false, needShadow = false, hasBase = false;
        float4 shO, shD, shBase, shNee;
        shaded += (uint32_t)__popc(__ballot_sync(0xffffffffu, active));
        if (active) {
            const float4 ro = st[rs][0][lane], rd = st[rs][1][lane];
            const float4 hit = st[rs][2][lane];
            const uint32_t instWord = sti[rs][lane];
            uint32_t flags = __float_as_uint(rd.w);
            // weight and directWeight of a path's first vertex are 1 (raygen.rgen:42-47): nothing was stored for them
            float4 wdw = st[rs][3][lane];
            if ((flags & 0xffffu) == 0u) wdw = make_float4(1.0f, 1.0f, 1.0f, 1.0f);
            uint32_t rng = __float_as_uint(ro.w);                                 // rayhit.rchit:668
            float3 weight = xyz(wdw);
            const float directWeight = wdw.w;
            const uint32_t prim = __float_as_uint(hit.w);
            const uint32_t inst = instWord & kInstMask;

            // :671-688 instance record, positions through object->world, normals through transformInvT
            float3 pos[3], nrm[3], emission;
            uint32_t handle, twofaced;
            if (!HASINST || (instWord & kInstBaked)) {
                // baked instance: both products were formed once when the scene was set (api.cu: k_bakeInstance, same
                // operation order), hit.w is the triangle's position in the world level: two independent fetches
                const float4 small = __ldg(sc.instSmall + inst);
                const float4* wp = sc.worldTris + 9 * (size_t)prim + 6;   // the copy in x, y, z order
                const float4* wn = sc.worldNrm + 3 * (size_t)prim;
#pragma unroll
                for (int k = 0; k < 3; k++) { pos[k] = xyz(__ldg(wp + k)); nrm[k] = xyz(__ldg(wn + k)); }
                emission = xyz(small);
                handle = __float_as_uint(small.w) & 0xffffu;
                twofaced = __float_as_uint(small.w) >> 16;
            } else {
                const float4* si = reinterpret_cast<const float4*>(sc.shadeInst + inst);
                const float4 m0 = __ldg(si), m1 = __ldg(si + 1), m2 = __ldg(si + 2);
                const float4 n0 = __ldg(si + 3), n1 = __ldg(si + 4), n2 = __ldg(si + 5);
                const uint4 p0 = __ldg(reinterpret_cast<const uint4*>(si) + 6);
                const uint4 p1 = __ldg(reinterpret_cast<const uint4*>(si) + 7);
                const float4* vpos = reinterpret_cast<const float4*>(((uint64_t)p0.y << 32) | p0.x);
                const float4* vnrm = reinterpret_cast<const float4*>(((uint64_t)p0.w << 32) | p0.z);
                handle = p1.x; twofaced = p1.y;
                emission = f3(n0.w, n1.w, n2.w);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    const float4 p = __ldg(vpos + 3 * (size_t)prim + k);
                    const float4 n = __ldg(vnrm + 3 * (size_t)prim + k);
                    pos[k] = f3(m0.x * p.x + m0.y * p.y + m0.z * p.z + m0.w, m1.x * p.x + m1.y * p.y + m1.z * p.z + m1.w,
                                m2.x * p.x + m2.y * p.y + m2.z * p.z + m2.w);
                    nrm[k] = f3(n0.x * n.x + n0.y * n.y + n0.z * n.z, n1.x * n.x + n1.y * n.y + n1.z * n.z,
                                n2.x * n.x + n2.y * n.y + n2.z * n.z);
                }
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
            // textured reflectance (extension; the diffuse and the rough-plastic records carry a hasTexture field the reference
            // never uses): the colour looked up at the hit's interpolated texture coordinate stands in for the record's constant
            // one.  A slow path taken by textured records only: the coordinates ride in the .w of the mesh's vertex arrays.
            bool textured = false;
            gsb_diffuse_bsdf texDiffuse;
            gsb_rough_plastic_bsdf texPlastic;
            if (EXT && (TYPE == GSB_BSDF_DIFFUSE || TYPE == GSB_BSDF_ROUGH_PLASTIC)) {
                const uint32_t bidx = handle & 0xffffu;
                const int tex = TYPE == GSB_BSDF_DIFFUSE ? ((const gsb_diffuse_bsdf*)sc.bsdf[GSB_BSDF_DIFFUSE])[bidx].hasTexture
                                                         : ((const gsb_rough_plastic_bsdf*)sc.bsdf[GSB_BSDF_ROUGH_PLASTIC])[bidx].hasTexture;
                if (tex > 0 && (uint32_t)tex <= sc.numTextures) {
                    uint32_t origPrim = prim;
                    if (!HASINST || (instWord & kInstBaked)) origPrim = __float_as_uint(__ldg(sc.worldTris + 9 * (size_t)prim + 6).w);
                    const ShadeInstance* sip = sc.shadeInst + inst;
                    float tu = 0.0f, tv = 0.0f;
#pragma unroll
                    for (int k = 0; k < 3; k++) {
                        const float bk = k == 0 ? bary.x : (k == 1 ? bary.y : bary.z);
                        tu += bk * __ldg(sip->vpos + 3 * (size_t)origPrim + k).w;
                        tv += bk * __ldg(sip->vnrm + 3 * (size_t)origPrim + k).w;
                    }
                    const float3 kd = textureEval(sc.textures[tex - 1], tu, tv);
                    textured = true;
                    if (TYPE == GSB_BSDF_DIFFUSE) {
                        texDiffuse = ((const gsb_diffuse_bsdf*)sc.bsdf[GSB_BSDF_DIFFUSE])[bidx];
                        texDiffuse.reflectance[0] = kd.x; texDiffuse.reflectance[1] = kd.y; texDiffuse.reflectance[2] = kd.z;
                    } else {
                        texPlastic = ((const gsb_rough_plastic_bsdf*)sc.bsdf[GSB_BSDF_ROUGH_PLASTIC])[bidx];
                        texPlastic.diffuse[0] = kd.x; texPlastic.diffuse[1] = kd.y; texPlastic.diffuse[2] = kd.z;
                    }
                }
            }
            if (!(EXT && textured)) sampleBSDFT<TYPE>(sc, handle & 0xffffu, rng, wo, wi, bsdfRes);            // :716
            else if (TYPE == GSB_BSDF_DIFFUSE) diffuseSample(texDiffuse, rng, wo, wi, bsdfRes);
            else roughPlasticSample(texPlastic, rng, wo, wi, bsdfRes);
            const float NoW = fabsf(wi.z);                                           // :717
            wi = onbUntransform(onb, wi);                                            // :718

            bool haveLight;
            const LightOutput lightRes = sampleLight<EXT>(sc, rng, position, haveLight);  // :720
            const float3 L = normalize(lightRes.position - position);                // :722
            const float3 wL = onbTransform(onb, L);                                  // :723
            const float Ldist = length(lightRes.position - position);                // :724
            const float NoL = fabsf(dot(SN, L));                                     // :725
            const float lightPdf = lightRes.pdf;                                     // :726
            BSDFOutput lightBsdfRes;
            lightBsdfRes.bsdf = f3(0.0f); lightBsdfRes.pdf = 0.0f; lightBsdfRes.isDelta = false;
            if (!(EXT && textured)) evalBSDFT<TYPE>(sc, handle & 0xffffu, wo, wL, lightBsdfRes);             // :729
            else if (TYPE == GSB_BSDF_DIFFUSE) diffuseEval(texDiffuse, wo, wL, lightBsdfRes);
            else roughPlasticEval(texPlastic, wo, wL, lightBsdfRes);
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
                // surface emission rides along only when there is some (bit 31 of the slot word says so)
                hasBase = emitted.x != 0.0f || emitted.y != 0.0f || emitted.z != 0.0f;
                shD = make_float4(L.x, L.y, L.z, __uint_as_float(slot | (hasBase ? kShadowHasBase : 0u)));
                shBase = make_float4(emitted.x, emitted.y, emitted.z, 0.0f);
                shNee = make_float4(nee.x, nee.y, nee.z, dwVisible);
            } else if (emitted.x != 0.0f || emitted.y != 0.0f || emitted.z != 0.0f) {
                // the running radiance of the sample is touched only by vertices that add to it
                const float cutoff = sc.params.cutoff;                               // raygen.rgen:60-63
                if (emitted.x < cutoff && emitted.y < cutoff && emitted.z < cutoff) {
                    float4 res = wf.result[slot];
                    res.x += emitted.x; res.y += emitted.y; res.z += emitted.z;
                    wf.result[slot] = res;
                }
            }
            if (!terminate) {
                *rayOrigin(wf, slot) = make_float4(newO.x, newO.y, newO.z, __uint_as_float(rng));   // rayhit.rchit:759
                *rayDirection(wf, slot) = make_float4(newD.x, newD.y, newD.z, __uint_as_float(flags));
                wf.weightDW[slot] = make_float4(weight.x, weight.y, weight.z, 1.0f);  // directWeight 1 unless the shadow stage finds the light
                if (wf.sortKey) wf.sortKey[slot] = sortKey(sc, newO, newD);
            }
        }
        if (terminate) wf.hitInst[slot] = GSB_NO_HIT;   // the path has ended: k_regen adds the sample to the image next iteration
        if constexpr (kDefer) {
            // The batch's place in the shadow queue is asked for here and used one batch later: the entries wait in shared
            // memory meanwhile.  Used at once, the atomic's round trip to L2 stalled every warp at the end of every batch
            // (r2w ncu: 10.5 % of the kernel's stall samples on the shuffle that broadcasts its result).
            flushShadow();
            const uint32_t mask = __ballot_sync(0xffffffffu, needShadow);
            pendMask = mask;
            pendNeed = needShadow;
            pendHasBase = needShadow && hasBase;
            pendRank = (uint32_t)__popc(mask & ((1u << lane) - 1u));
            if (mask) {
                if (lane == 0u) pendBase = atomicAdd(&qc[Q_SHADOW], (uint32_t)__popc(mask));
                if (needShadow) {
                    sh[0][lane] = shO; sh[1][lane] = shD; sh[2][lane] = shNee;
                    if (hasBase) sh[3][lane] = shBase;
                }
            }
        } else {
            const uint32_t sp = queueSlot(qc, Q_SHADOW, needShadow);
            if (needShadow) {
                wf.shadowO[sp] = shO;
                wf.shadowD[sp] = shD;
                if (hasBase) wf.shadowBase[sp] = shBase;
                wf.shadowNee[sp] = shNee;
            }
        }
        if (!haveNew) break;
        prevActive = newActive; prevSlot = newSlot; stage ^= 1u;
        __syncwarp();   // the stage just read is the next one to be filled
    }
    flushShadow();
    if (lane == 0u && shaded) atomicAdd(&wf.stats[ST_VERTICES], (unsigned long long)shaded);
}

// ---- shadow ---------------------------------------------------------------------------------------
// The ray stream of k_shadow: work index = entry of the shadow queue written by k_shade.
struct ShadowSource {
    const SceneDev& sc;
    const WavefrontState& wf;
    uint32_t* qc;
    uint32_t count;
    WorkRange range;
    uint32_t traced = 0;

    GSB_D ShadowSource(const SceneDev& s, const WavefrontState& w, uint32_t* q) : sc(s), wf(w), qc(q), count(q[Q_SHADOW]) {}
    GSB_D bool drained() const { return range.exhausted; }

    GSB_D bool fetch(bool want, RayJob& job) {
        const uint32_t i = range.take(&qc[Q_NEXT_SHADOW], count, want);
        const uint32_t ahead = range.peek(count);
        if (ahead != kNoTag) {
            prefetchL1(wf.shadowO + ahead);
            prefetchL1(wf.shadowD + ahead);
        }
        if (i == kNoTag) return false;
        const float4 o = wf.shadowO[i], d = wf.shadowD[i];
        job.ox = o.x; job.oy = o.y; job.oz = o.z;
        job.dx = d.x; job.dy = d.y; job.dz = d.z;
        job.tmin = 0.01f; job.tmax = o.w;                                            // rayhit.rchit:737-748
        job.tag = i;
        job.aux = __float_as_uint(d.w);   // the path slot; bit 31 (kShadowHasBase) stays in until commit
        traced++;
        return true;
    }

    GSB_D void commit(bool have, const HitRecord& hit, uint32_t i, uint32_t slot) {
        if (!have) return;
        const bool visible = hit.inst == GSB_NO_HIT;
        const bool hasBase = (slot & kShadowHasBase) != 0u;
        slot &= ~kShadowHasBase;
        float3 emitted = f3(0.0f);
        if (visible) {
            const float4 nee = wf.shadowNee[i];
            emitted = emitted + xyz(nee);                                            // rayhit.rchit:752
            wf.weightDW[slot].w = nee.w;                                             // :786 (neeDone)
        }
        if (hasBase) emitted = emitted + xyz(wf.shadowBase[i]);                      // :763-768
        const float cutoff = sc.params.cutoff;                                       // raygen.rgen:60-63
        if ((emitted.x != 0.0f || emitted.y != 0.0f || emitted.z != 0.0f) &&
            emitted.x < cutoff && emitted.y < cutoff && emitted.z < cutoff) {
            float4 res = wf.result[slot];
            res.x += emitted.x; res.y += emitted.y; res.z += emitted.z;
            wf.result[slot] = res;
        }
    }
};

template <bool STATS, bool HASINST, bool POOL>
__global__ void __launch_bounds__(kBlock, GSB_TRACE_MIN_BLOCKS) k_shadow(SceneDev sc, WavefrontState wf, uint32_t parity, uint32_t refillMin) {
    ShadowSource src(sc, wf, wf.qCount + parity * Q_COUNT);
    TraceStats ts = {0u, 0u, 0u};
    traceStream<true, STATS, HASINST, POOL>(sc, src, refillMin, &ts, (uint32_t*)&wf.stats[ST_OVERFLOW]);
    if (STATS) flushTraceStats(wf.stats, ts, src.traced);
}

// one thread, end of the body of the device-side render loop: keep looping while slots are live (and, as a safety net
// against a hung device, while the iteration bound of the call is not reached)
__global__ void k_loopCondition(cudaGraphConditionalHandle handle, WavefrontState wf, unsigned long long maxIters) {
    const bool more = wf.stats[ST_LIVE] != 0ull && wf.stats[ST_LOOPITERS] + 2ull <= maxIters;
    wf.stats[ST_LOOPITERS] += 2ull;
    cudaGraphSetConditional(handle, more ? 1u : 0u);
}

// one thread, once per render call: roll the last iteration's queue sizes into the totals
__global__ void k_foldCounters(WavefrontState wf, uint32_t parity, unsigned long long paths) {
    uint32_t* bank = wf.qCount + parity * Q_COUNT;
    unsigned long long ext = bank[Q_TRACED];
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

__global__ void k_iota(uint32_t* out, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}

void freeSort(WavefrontState& wf) {
    void* ptrs[] = {wf.sortKey, wf.sortKeyOut, wf.order, wf.iota, wf.sortTemp};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    wf.sortKey = wf.sortKeyOut = wf.order = wf.iota = nullptr;
    wf.sortTemp = nullptr;
    wf.sortTempBytes = 0;
}

void allocWavefront(gsb_context* ctx, uint32_t slots, bool sorting) {
    WavefrontState& wf = ctx->wf;
    if (!sorting && wf.sortKey) freeSort(wf);
    if (sorting && (!wf.sortKey || wf.capacity < slots)) {
        freeSort(wf);
        const size_t m = std::max(slots, wf.capacity);
        ensure(wf.sortKey, m); ensure(wf.sortKeyOut, m); ensure(wf.order, m); ensure(wf.iota, m);
        k_iota<<<(unsigned)((m + 255) / 256), 256, 0, ctx->stream>>>(wf.iota, (uint32_t)m);
        size_t bytes = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, bytes, wf.sortKey, wf.sortKeyOut, wf.iota, wf.order, (int)m, 0, 32, ctx->stream);
        GSB_CUDA(cudaMalloc(&wf.sortTemp, std::max<size_t>(bytes, 16)));
        wf.sortTempBytes = bytes;
    }
    if (wf.capacity >= slots && wf.qCount) return;
    const size_t n = slots;
    ensure(wf.ray, 2 * n); ensure(wf.weightDW, n); ensure(wf.result, n);
    ensure(wf.hitTUVP, n); ensure(wf.hitInst, n);

    ensure(wf.shadowO, n); ensure(wf.shadowD, n); ensure(wf.shadowBase, n); ensure(wf.shadowNee, n);
    if (!wf.qCount) ensure(wf.qCount, (size_t)2 * Q_COUNT);
    if (!wf.stats) {
        ensure(wf.stats, (size_t)ST_COUNT);
        GSB_CUDA(cudaMemsetAsync(wf.stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
    }
    wf.capacity = slots;
}

template <int TYPE>
void launchShade(gsb_context* ctx, const SceneDev& sc, int grid, uint32_t parity, uint32_t slots) {
    const bool ext = ctx->numTextures > 0 || ctx->env.type != 0;
    if (ctx->numTraceInstances > 0) {
        if (ext) k_shade<TYPE, true, true><<<grid, kBlock, 0, ctx->stream>>>(sc, ctx->wf, parity, slots, ctx->shadeDirectMin);
        else k_shade<TYPE, true, false><<<grid, kBlock, 0, ctx->stream>>>(sc, ctx->wf, parity, slots, ctx->shadeDirectMin);
    } else {
        if (ext) k_shade<TYPE, false, true><<<grid, kBlock, 0, ctx->stream>>>(sc, ctx->wf, parity, slots, ctx->shadeDirectMin);
        else k_shade<TYPE, false, false><<<grid, kBlock, 0, ctx->stream>>>(sc, ctx->wf, parity, slots, ctx->shadeDirectMin);
    }
}

}  // namespace

void freeWavefront(gsb_context* ctx) {
    WavefrontState& wf = ctx->wf;
    void* ptrs[] = {wf.ray, wf.weightDW, wf.result, wf.hitTUVP, wf.hitInst,
                    wf.shadowO, wf.shadowD, wf.shadowBase, wf.shadowNee, wf.qCount, wf.stats,
                    wf.sortKey, wf.sortKeyOut, wf.order, wf.iota, wf.sortTemp};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    wf = WavefrontState();
}

void finishPending(gsb_context* ctx) {
    if (ctx->pending.empty()) return;
    cudaStream_t s = ctx->stream;
    unsigned long long iters = 0;
    GSB_CUDA(cudaMemcpyAsync(&iters, &ctx->wf.stats[ST_ITERS], sizeof(iters), cudaMemcpyDeviceToHost, s));
    GSB_CUDA(cudaStreamSynchronize(s));
    GSB_CUDA(cudaGetLastError());
    const PendingRender& last = ctx->pending.back();
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, last.start, last.stop);
    ctx->lastRenderMs[0] = ms;
    for (int k = 1; k < 5; k++) ctx->lastRenderMs[k] = 0.0;
    const unsigned long long ran = iters - ctx->seenIters;   // wavefront iterations of all pending calls together
    ctx->launches += ran * last.launchesPerIter + ran / 2;    // stage kernels + one loop condition per two iterations
    ctx->lastRenderIters = ctx->pending.size() == 1 ? ran : ran / ctx->pending.size();
    ctx->seenIters = iters;
    for (PendingRender& pr : ctx->pending) {
        cudaGraphExecDestroy(pr.exec);
        cudaGraphDestroy(pr.graph);
        cudaEventDestroy(pr.start);
        cudaEventDestroy(pr.stop);
    }
    ctx->pending.clear();
}

void renderImpl(gsb_context* ctx, uint32_t width, uint32_t height, uint32_t first, uint32_t count, uint32_t stride) {
    cudaStream_t s = ctx->stream;
    const uint32_t npix = width * height;
    if (ctx->profiling || !ctx->useGraphLoop || ctx->pending.size() >= 64) finishPending(ctx);
    if (ctx->width != width || ctx->height != height || !ctx->dAccum) {
        ensure(ctx->dAccum, (size_t)npix);
        GSB_CUDA(cudaMemsetAsync(ctx->dAccum, 0, (size_t)npix * sizeof(float4), s));
        ctx->width = width;
        ctx->height = height;
    }
    if (count == 0) return;
    // paths in flight: enough to fill the machine several times over, small enough that the path state stays near L2
    const unsigned long long totalWork = (unsigned long long)npix * count;
    uint32_t slots = (uint32_t)std::min<unsigned long long>(totalWork, ctx->targetSlots);
    if (slots < totalWork) slots = std::max(32u, slots & ~31u);
    const SceneDev sc = ctx->sceneDev();
    const bool sorting = sc.sortBits > 0u;
    allocWavefront(ctx, slots, sorting);
    WavefrontState& wf = ctx->wf;
    RenderArgs ra;
    ra.width = width; ra.height = height; ra.npix = npix;
    ra.first = first; ra.count = count; ra.stride = stride;
    ra.slots = slots;
    ra.totalWork = totalWork;
    ra.accum = ctx->dAccum;
    ra.invNpix = 1.0 / (double)npix;
    ra.refillMin = ctx->refillMin;
    ra.camZ = (std::fmax((float)width, (float)height) / 2.0f) / std::tan(ctx->fov / 2.0f);
    ra.order = sorting ? wf.order : nullptr;
    GSB_CUDA(cudaMemsetAsync(wf.qCount, 0, 2 * Q_COUNT * sizeof(uint32_t), s));
    const unsigned long long liveInit[2] = {(unsigned long long)slots, 0ull};
    GSB_CUDA(cudaMemcpyAsync(wf.stats, liveInit, sizeof(liveInit), cudaMemcpyHostToDevice, s));
    const unsigned long long nextWork = slots;   // the first `slots` work items start out in the slots (k_generate)
    GSB_CUDA(cudaMemcpyAsync(&wf.stats[ST_NEXTWORK], &nextWork, sizeof(nextWork), cudaMemcpyHostToDevice, s));

    cudaEvent_t evStart = nullptr, evStop = nullptr;
    GSB_CUDA(cudaEventCreate(&evStart));
    GSB_CUDA(cudaEventCreate(&evStop));
    GSB_CUDA(cudaEventRecord(evStart, s));

    const int fullGrid = (int)((slots + kBlock - 1) / kBlock);
    int dev = ctx->device, sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int queueGrid = std::min(fullGrid, sms * 16);
    // persistent traversal kernels: exactly as many blocks as the machine holds at once
    auto residentGrid = [&](const void* kernel) {
        int perSm = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, kernel, kBlock, 0);
        return std::max(1, std::min(fullGrid, sms * std::max(perSm, 1)));
    };
    // the traversal kernels exist per (instrumented, scene has instances that are not baked)
    using ExtendKernel = void (*)(SceneDev, WavefrontState, RenderArgs, uint32_t);
    using ShadowKernel = void (*)(SceneDev, WavefrontState, uint32_t, uint32_t);
    static const ExtendKernel extendKernels[8] = {k_extend<false, false, false>, k_extend<false, true, false>, k_extend<true, false, false>, k_extend<true, true, false>,
                                                  k_extend<false, false, true>,  k_extend<false, true, true>,  k_extend<true, false, true>,  k_extend<true, true, true>};
    static const ShadowKernel shadowKernels[8] = {k_shadow<false, false, false>, k_shadow<false, true, false>, k_shadow<true, false, false>, k_shadow<true, true, false>,
                                                  k_shadow<false, false, true>,  k_shadow<false, true, true>,  k_shadow<true, false, true>,  k_shadow<true, true, true>};
    const int hasInst = ctx->numTraceInstances > 0 ? 1 : 0;
    const ExtendKernel extendKernel = extendKernels[(ctx->poolExtend ? 4 : 0) | ((ctx->instrumented & 1) ? 2 : 0) | hasInst];
    const ShadowKernel shadowKernel = shadowKernels[(ctx->poolShadow ? 4 : 0) | ((ctx->instrumented & 2) ? 2 : 0) | hasInst];
    const int extendGrid = residentGrid((const void*)extendKernel);
    const int shadowGrid = residentGrid((const void*)shadowKernel);
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
    const int regenGrid = (int)((slots + kRegenSpan - 1) / kRegenSpan);
    // key bits that matter: 3 per grid level + the octant
    const int sortEndBit = (int)std::min(32u, 3u * sc.sortBits + (sc.sortMode == 2u ? 6u : 3u));
    auto iteration = [&]() {
        mark();
        k_regen<<<regenGrid, kRegenBlock, 0, s>>>(sc, wf, ra);
        if (sorting) {
            size_t bytes = wf.sortTempBytes;
            cub::DeviceRadixSort::SortPairs(wf.sortTemp, bytes, wf.sortKey, wf.sortKeyOut, wf.iota, wf.order, (int)slots, 0, sortEndBit, s);
        }
        mark();
        extendKernel<<<extendGrid, kBlock, 0, s>>>(sc, wf, ra, parity);
        mark();
        if (ctx->typeUsed[0]) launchShade<0>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[1]) launchShade<1>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[2]) launchShade<2>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[3]) launchShade<3>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[4]) launchShade<4>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[5]) launchShade<5>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[6]) launchShade<6>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[7]) launchShade<7>(ctx, sc, queueGrid, parity, slots);
        if (ctx->typeUsed[8]) launchShade<8>(ctx, sc, queueGrid, parity, slots);
        mark();
        shadowKernel<<<shadowGrid, kBlock, 0, s>>>(sc, wf, parity, ctx->shadowControl);
        mark();
        parity ^= 1u;
        ctx->launches += 3 + (uint64_t)typesUsed;
        ctx->lastRenderIters++;
    };

    // An upper bound on the iterations of this call: every iteration advances every live slot's path by one bounce.
    const unsigned long long pathsPerSlot = (totalWork + slots - 1) / slots;
    const unsigned long long maxIters = (pathsPerSlot + 1ull) * (unsigned long long)(ctx->params.max_depth + 4) + 16ull;

    bool looped = false;
    if (!prof && ctx->useGraphLoop) {
        // Device-side loop: ONE graph launch per render call.  A WHILE conditional node whose body is two wavefront
        // iterations (one per counter bank) plus the loop condition; the host is not involved until the frame is done,
        // so there are no launch gaps and no sensitivity to host scheduling.
        cudaGraph_t graph = nullptr;
        cudaGraphExec_t exec = nullptr;
        do {
            if (cudaGraphCreate(&graph, 0) != cudaSuccess) break;
            cudaGraphConditionalHandle handle;
            if (cudaGraphConditionalHandleCreate(&handle, graph, 1, cudaGraphCondAssignDefault) != cudaSuccess) break;
            cudaGraphNodeParams np = {};
            np.type = cudaGraphNodeTypeConditional;
            np.conditional.handle = handle;
            np.conditional.type = cudaGraphCondTypeWhile;
            np.conditional.size = 1;
            cudaGraphNode_t node;
            if (cudaGraphAddNode(&node, graph, nullptr, 0, &np) != cudaSuccess) break;
            cudaGraph_t body = np.conditional.phGraph_out[0];
            if (cudaStreamBeginCaptureToGraph(s, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed) != cudaSuccess) break;
            iteration();
            iteration();
            k_loopCondition<<<1, 1, 0, s>>>(handle, wf, maxIters);
            cudaGraph_t captured = nullptr;
            if (cudaStreamEndCapture(s, &captured) != cudaSuccess) break;
            if (cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) break;
            GSB_CUDA(cudaMemsetAsync(&wf.stats[ST_LOOPITERS], 0, sizeof(unsigned long long), s));
            if (cudaGraphLaunch(exec, s) != cudaSuccess) break;
            looped = true;
        } while (false);
        if (!looped) {
            cudaGetLastError();  // clear; fall back to the host-driven loop below
            ctx->useGraphLoop = false;
            parity = 0;
        }
        if (looped) {
            // asynchronous from here on: the graph, its executable and the timing events are released by
            // finishPending() once the stream has drained (gsb_synchronize, read-back, counters, ...)
            k_foldCounters<<<1, 1, 0, s>>>(wf, 1u, totalWork);   // an even number of iterations ran: the last bank is 1
            GSB_CUDA(cudaEventRecord(evStop, s));
            PendingRender pr;
            pr.exec = exec; pr.graph = graph; pr.start = evStart; pr.stop = evStop;
            pr.launchesPerIter = 3u + (uint32_t)typesUsed;
            ctx->pending.push_back(pr);
            ctx->launches -= 2ull * (3 + (uint64_t)typesUsed);   // iteration() ran twice on the host during capture
            ctx->launches += 1;
            return;
        }
        if (exec) cudaGraphExecDestroy(exec);
        if (graph) cudaGraphDestroy(graph);
    }

    // Host loop (profiling mode, or no conditional graph nodes): batches of iterations; the live-slot count of
    // batch b is read back while batch b+1 is already queued, so the device rarely waits for the host.
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
        if ((unsigned long long)b * batch > maxIters + 2 * batch) break;
    }
    k_foldCounters<<<1, 1, 0, s>>>(wf, parity ^ 1u, totalWork);
    ctx->launches += 1;
    GSB_CUDA(cudaEventRecord(evStop, s));
    GSB_CUDA(cudaEventSynchronize(evStop));
    GSB_CUDA(cudaGetLastError());
    float ms = 0.0f;
    GSB_CUDA(cudaEventElapsedTime(&ms, evStart, evStop));
    ctx->seenIters += ctx->lastRenderIters;
    ctx->lastRenderMs[0] = ms;
    for (int k = 1; k < 5; k++) ctx->lastRenderMs[k] = 0.0;
    if (prof) {
        for (size_t i = 0; i + 4 < pev.size(); i += 5) {
            // marks of one iteration: regenerate (+ sort) | extend | shade | shadow  ->  finish, extend, shade, shadow
            static const int where[4] = {4, 1, 2, 3};
            for (int k = 0; k < 4; k++) {
                float t = 0.0f;
                cudaEventElapsedTime(&t, pev[i + k], pev[i + k + 1]);
                ctx->lastRenderMs[where[k]] += t;
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

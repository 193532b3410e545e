// api.cu — the extern "C" layer of include/gpuspectral_b200.h: context, scene upload,
// acceleration-structure builds, ray batches, framebuffer read-back, diagnostic entry points.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "bsdf.cuh"
#include "context.h"
#include "traverse.cuh"

namespace gsb {

static thread_local std::string g_lastError;
void setLastError(const std::string& msg) { g_lastError = msg; }

namespace {

template <class F>
gsb_status guarded(F&& f) {
    try {
        return f();
    } catch (const CudaFailure& e) {
        char buf[512];
        snprintf(buf, sizeof(buf), "CUDA error %d (%s) at %s:%d: %s", (int)e.err, cudaGetErrorString(e.err), e.file, e.line, e.what);
        setLastError(buf);
        return GSB_ERR_CUDA;
    } catch (const std::bad_alloc&) {
        setLastError("out of host memory");
        return GSB_ERR_OOM;
    } catch (const std::exception& e) {
        setLastError(e.what());
        return GSB_ERR_INVALID;
    }
}

gsb_status fail(gsb_status st, const std::string& msg) {
    setLastError(msg);
    return st;
}

const uint32_t kBsdfStride[GSB_BSDF_TYPE_COUNT] = {
    sizeof(gsb_diffuse_bsdf), sizeof(gsb_smooth_dielectric_bsdf), sizeof(gsb_smooth_conductor_bsdf), sizeof(gsb_smooth_plastic_bsdf),
    sizeof(gsb_rough_conductor_bsdf), sizeof(gsb_smooth_floor_bsdf), sizeof(gsb_rough_floor_bsdf), sizeof(gsb_rough_plastic_bsdf),
    sizeof(gsb_rough_dielectric_bsdf)};

template <class T>
T* uploadArray(const T* host, size_t count, cudaStream_t s) {
    void* p = nullptr;
    GSB_CUDA(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
    if (count) GSB_CUDA(cudaMemcpyAsync(p, host, count * sizeof(T), cudaMemcpyHostToDevice, s));
    return (T*)p;
}

struct Bind {
    int prev = -1;
    explicit Bind(int dev) : dev_(dev) {
        cudaGetDevice(&prev);
        if (prev != dev) cudaSetDevice(dev);
    }
    ~Bind() {
        if (prev >= 0 && prev != dev_) cudaSetDevice(prev);   // the caller's current device is not ours to change
    }
    int dev_ = 0;
};

// ---- kernels of this translation unit ----------------------------------------------------------------
// The ray stream of gsb_trace_batch: work index = position in the caller's ray array.
template <bool ANYHIT>
struct BatchSource {
    const gsb_ray* rays;
    gsb_hit* hits;
    uint32_t n;
    uint32_t* counter;
    WorkRange range;
    uint32_t traced = 0;

    GSB_D bool drained() const { return range.exhausted; }
    GSB_D bool fetch(bool want, RayJob& job) {
        const uint32_t i = range.take(counter, n, want);
        if (i == kNoTag) return false;
        const float4 a = __ldg(reinterpret_cast<const float4*>(rays) + 2 * (size_t)i);
        const float4 b = __ldg(reinterpret_cast<const float4*>(rays) + 2 * (size_t)i + 1);
        job.ox = a.x; job.oy = a.y; job.oz = a.z; job.tmin = a.w;
        job.dx = b.x; job.dy = b.y; job.dz = b.z; job.tmax = b.w;
        job.tag = i;
        job.aux = 0u;
        traced++;
        return true;
    }
    GSB_D void commit(bool have, const HitRecord& h, uint32_t i, uint32_t) {
        if (!have) return;
        gsb_hit o;
        if (ANYHIT) {
            o.t = 0.0f; o.u = 0.0f; o.v = 0.0f;
            o.inst = h.inst == GSB_NO_HIT ? GSB_NO_HIT : 0u;
            o.prim = o.inst;
        } else {
            o.t = h.t; o.u = h.u; o.v = h.v; o.prim = h.prim;
            o.inst = h.inst == GSB_NO_HIT ? GSB_NO_HIT : (h.inst & kInstMask);   // gl_InstanceID out of the instance word
        }
        hits[i] = o;
    }
};

template <bool ANYHIT, bool STATS, bool HASINST, bool POOL>
__global__ void __launch_bounds__(kTraceBlock) k_traceBatch(SceneDev sc, const gsb_ray* __restrict__ rays, uint32_t n,
                                                            gsb_hit* __restrict__ hits, unsigned long long* stats, uint32_t refillMin) {
    BatchSource<ANYHIT> src;
    src.rays = rays; src.hits = hits; src.n = n;
    src.counter = reinterpret_cast<uint32_t*>(&stats[ST_BATCHNEXT]);
    TraceStats ts = {0u, 0u, 0u};
    traceStream<ANYHIT, STATS, HASINST, POOL>(sc, src, refillMin, &ts, (uint32_t*)&stats[ST_OVERFLOW]);
    if (STATS) {
        uint32_t nn = ts.nodes, nt = ts.tris, ni = ts.insts, nr = src.traced;
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
}

// Bake rule, device side: vertex k of triangle t of an instance in world space, x' = ((m0*x + m1*y) + m2*z) + m3 per
// row, every operation rounded on its own (no FMA) — the order of gl_ObjectToWorldEXT * vec4(p, 1) (rayhit.rchit:679-681)
// and of the checker — and its normal through transformInvT, (n0*x + n1*y) + n2*z per row (rayhit.rchit:686-688).
// .w of the positions carries what the world level reports: v0.w = gl_PrimitiveID, v1.w = the instance word.
__global__ void k_bakeInstance(const float4* __restrict__ vpos, const float4* __restrict__ vnrm, uint32_t ntris, float4 m0, float4 m1,
                               float4 m2, float4 n0, float4 n1, float4 n2, uint32_t instWord, float4* __restrict__ outPos,
                               float4* __restrict__ outNrm) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntris) return;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const float4 p = __ldg(vpos + 3 * (size_t)t + k);
        const float x = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m0.x, p.x), __fmul_rn(m0.y, p.y)), __fmul_rn(m0.z, p.z)), m0.w);
        const float y = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m1.x, p.x), __fmul_rn(m1.y, p.y)), __fmul_rn(m1.z, p.z)), m1.w);
        const float z = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m2.x, p.x), __fmul_rn(m2.y, p.y)), __fmul_rn(m2.z, p.z)), m2.w);
        outPos[3 * (size_t)t + k] = make_float4(x, y, z, __uint_as_float(k == 0 ? t : (k == 1 ? instWord : 0u)));
        const float4 n = __ldg(vnrm + 3 * (size_t)t + k);
        const float nx = __fadd_rn(__fadd_rn(__fmul_rn(n0.x, n.x), __fmul_rn(n0.y, n.y)), __fmul_rn(n0.z, n.z));
        const float ny = __fadd_rn(__fadd_rn(__fmul_rn(n1.x, n.x), __fmul_rn(n1.y, n.y)), __fmul_rn(n1.z, n.z));
        const float nz = __fadd_rn(__fadd_rn(__fmul_rn(n2.x, n.x), __fmul_rn(n2.y, n.y)), __fmul_rn(n2.z, n.z));
        outNrm[3 * (size_t)t + k] = make_float4(nx, ny, nz, 0.0f);
    }
}

// world-level normals from bake order into leaf order (position k of the world level holds baked triangle leafPrim[k])
__global__ void k_gatherNormals(const uint32_t* __restrict__ leafPrim, uint32_t n, const float4* __restrict__ in, float4* __restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t p = leafPrim[k];
#pragma unroll
    for (int j = 0; j < 3; j++) out[3 * (size_t)k + j] = __ldg(in + 3 * (size_t)p + j);
}

__global__ void k_resolve(const float4* __restrict__ accum, uint32_t npix, float4* __restrict__ hdr, uchar4* __restrict__ ldr) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 a = accum[p];
    const float inv = a.w > 0.0f ? 1.0f / a.w : 0.0f;
    const float3 c = f3(a.x * inv, a.y * inv, a.z * inv);
    if (hdr) hdr[p] = make_float4(c.x, c.y, c.z, 1.0f);
    if (ldr) {
        // linear -> sRGB 8-bit UNORM, the conversion the B8G8R8A8_SRGB swapchain applies on write
        auto enc = [](float v) {
            float e;
            if (!(v > 0.0f)) e = 0.0f;
            else if (v >= 1.0f) e = 1.0f;
            else e = v <= 0.0031308f ? 12.92f * v : 1.055f * powf(v, 1.0f / 2.4f) - 0.055f;
            return (unsigned char)floorf(e * 255.0f + 0.5f);
        };
        ldr[p] = make_uchar4(enc(c.x), enc(c.y), enc(c.z), 255);
    }
}

__global__ void k_testRng(const uint32_t* idx, const uint32_t* ts, uint32_t n, uint32_t nDraws, uint32_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t* o = out + (size_t)i * (1 + nDraws);
    uint32_t rng = pcgHash(tea(idx[i], ts[i]));
    o[0] = rng;
    for (uint32_t k = 0; k < nDraws; k++) o[1 + k] = randPcg(rng);
}

// the primary ray exactly as startPath (wavefront.cu) forms it: camZ evaluated once on the host, same expression order,
// same compile flags (the Makefile gives this translation unit the shading kernels' division / square-root precision)
__global__ void k_testCamera(SceneDev sc, uint32_t W, uint32_t H, float camZ, const uint32_t* px, const uint32_t* py, uint32_t n, gsb_ray* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float sx = (float)W, sy = (float)H;
    const float fx = (float)px[i] - sx / 2.0f;
    const float fy = (float)py[i] - sy / 2.0f;
    const float z = camZ;
    const float3 dir = normalize(f3(-fx, fy, z));
    const float* V = sc.view;
    float3 d = f3(V[0] * dir.x + V[4] * dir.y + V[8] * dir.z, V[1] * dir.x + V[5] * dir.y + V[9] * dir.z,
                  V[2] * dir.x + V[6] * dir.y + V[10] * dir.z);
    d.y *= -1.0f;
    gsb_ray r;
    r.origin[0] = sc.eye[0]; r.origin[1] = sc.eye[1]; r.origin[2] = sc.eye[2];
    r.dir[0] = d.x; r.dir[1] = d.y; r.dir[2] = d.z;
    r.tmin = 0.0f;
    r.tmax = 1e10f;
    out[i] = r;
}

__global__ void k_testBsdfSample(SceneDev sc, uint32_t handle, const float* wo, const uint32_t* seed, uint32_t n, float* wi, float* f,
                                 float* pdf, uint32_t* isDelta, uint32_t* seedOut) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t rng = seed[i];
    float3 w = f3(0.0f);
    BSDFOutput res;
    res.bsdf = f3(0.0f); res.pdf = 0.0f; res.isDelta = false;
    sampleBSDF(sc, handle, rng, f3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), w, res);
    wi[3 * i] = w.x; wi[3 * i + 1] = w.y; wi[3 * i + 2] = w.z;
    f[3 * i] = res.bsdf.x; f[3 * i + 1] = res.bsdf.y; f[3 * i + 2] = res.bsdf.z;
    pdf[i] = res.pdf;
    isDelta[i] = res.isDelta ? 1u : 0u;
    seedOut[i] = rng;
}

__global__ void k_testBsdfEval(SceneDev sc, uint32_t handle, const float* wo, const float* wi, uint32_t n, float* f, float* pdf, uint32_t* isDelta) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    BSDFOutput res;
    res.bsdf = f3(0.0f); res.pdf = 0.0f; res.isDelta = false;
    evalBSDF(sc, handle, f3(wo[3 * i], wo[3 * i + 1], wo[3 * i + 2]), f3(wi[3 * i], wi[3 * i + 1], wi[3 * i + 2]), res);
    f[3 * i] = res.bsdf.x; f[3 * i + 1] = res.bsdf.y; f[3 * i + 2] = res.bsdf.z;
    pdf[i] = res.pdf;
    isDelta[i] = res.isDelta ? 1u : 0u;
}

__global__ void k_testSampleLight(SceneDev sc, const float* pos, const uint32_t* seed, uint32_t n, float* lpos, float* emission, float* pdf, uint32_t* seedOut) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t rng = seed[i];
    bool have;
    const LightOutput r = sampleLight(sc, rng, f3(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]), have);
    lpos[3 * i] = r.position.x; lpos[3 * i + 1] = r.position.y; lpos[3 * i + 2] = r.position.z;
    emission[3 * i] = r.emission.x; emission[3 * i + 1] = r.emission.y; emission[3 * i + 2] = r.emission.z;
    pdf[i] = r.pdf;
    seedOut[i] = rng;
}

// device scratch that frees itself
// Temporary device buffers of one API call.  With a stream they come from the stream-ordered pool (no device-wide
// synchronisation in cudaFree, and the pool keeps the memory: see the release threshold set in gsb_create).
struct Scratch {
    std::vector<void*> ptrs;
    cudaStream_t stream = nullptr;
    bool ordered = false;
    Scratch() {}
    explicit Scratch(cudaStream_t s) : stream(s), ordered(true) {}
    template <class T>
    T* alloc(size_t count) {
        void* p = nullptr;
        if (ordered) GSB_CUDA(cudaMallocAsync(&p, std::max<size_t>(count, 1) * sizeof(T), stream));
        else GSB_CUDA(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
        ptrs.push_back(p);
        return (T*)p;
    }
    template <class T>
    T* upload(const T* host, size_t count, cudaStream_t s) {
        T* p = alloc<T>(count);
        if (count) GSB_CUDA(cudaMemcpyAsync(p, host, count * sizeof(T), cudaMemcpyHostToDevice, s));
        return p;
    }
    ~Scratch() {
        for (void* p : ptrs) {
            if (ordered) cudaFreeAsync(p, stream);
            else cudaFree(p);
        }
    }
};

void ensureStats(gsb_context* ctx) {
    if (ctx->wf.stats) return;
    void* p = nullptr;
    GSB_CUDA(cudaMalloc(&p, ST_COUNT * sizeof(unsigned long long)));
    ctx->wf.stats = (unsigned long long*)p;
    GSB_CUDA(cudaMemsetAsync(ctx->wf.stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
}

bool sceneReady(gsb_context* ctx) { return ctx->haveInstances && ctx->tlas.buffer; }

void checkOverflow(gsb_context* ctx) {
    unsigned long long ov = 0, live = 0;
    GSB_CUDA(cudaMemcpyAsync(&ov, &ctx->wf.stats[ST_OVERFLOW], sizeof(ov), cudaMemcpyDeviceToHost, ctx->stream));
    GSB_CUDA(cudaMemcpyAsync(&live, &ctx->wf.stats[ST_LIVE], sizeof(live), cudaMemcpyDeviceToHost, ctx->stream));
    GSB_CUDA(cudaStreamSynchronize(ctx->stream));
    if (ov) {
        GSB_CUDA(cudaMemsetAsync(&ctx->wf.stats[ST_OVERFLOW], 0, sizeof(ov), ctx->stream));
        throw std::runtime_error("traversal stack overflow: BVH deeper than the traversal stack");
    }
    if (live) {
        // the iteration bound of the render loop (a safety net against a hung device) stopped a frame whose paths were not all
        // finished: samples are missing from the image, which must not pass for a completed render
        GSB_CUDA(cudaMemsetAsync(&ctx->wf.stats[ST_LIVE], 0, sizeof(live), ctx->stream));
        throw std::runtime_error("render stopped at its iteration bound with " + std::to_string(live) + " path slots still live: the frame is incomplete");
    }
}

// Waits for asynchronous render calls and reports a traversal-stack overflow any of them hit.
void drain(gsb_context* ctx) {
    if (ctx->pending.empty()) return;
    finishPending(ctx);
    checkOverflow(ctx);
}

}  // namespace
}  // namespace gsb

using namespace gsb;

// ---- gsb_trace_batch_static: the ray-batch entry for callers that may not allocate ---------------------
// Scratch of that entry, held by every GPU that loads this library until the process exits: a control block (the
// kernel's work counter and traversal-stack overflow flag, the ST_* words k_traceBatch touches), then the device
// copies of the caller's rays and of the hits.  In HOST code the name below is only a stand-in for the symbol, not a
// device address: a kernel is given the pointer cudaGetSymbolAddress returns for the device that is current.
constexpr size_t kStaticControlBytes = 256;
static_assert(ST_COUNT * sizeof(unsigned long long) <= kStaticControlBytes, "control block too small");
constexpr size_t kStaticRaysOffset = kStaticControlBytes;
constexpr size_t kStaticHitsOffset = kStaticRaysOffset + (size_t)GSB_STATIC_BATCH_MAX_RAYS * sizeof(gsb_ray);
constexpr size_t kStaticScratchBytes = kStaticHitsOffset + (size_t)GSB_STATIC_BATCH_MAX_RAYS * sizeof(gsb_hit);
static_assert(kStaticHitsOffset % 16 == 0, "hit array must keep 16-byte alignment");
static __device__ __align__(256) unsigned char g_staticBatchScratch[kStaticScratchBytes];

// Device address of the scratch on `device`, which must be the current device: resolved once per device.
static unsigned char* staticBatchScratch(int device) {
    constexpr int kMaxDevices = 64;
    static std::mutex lock;
    static unsigned char* address[kMaxDevices] = {};
    if (device < 0 || device >= kMaxDevices) throw std::runtime_error("gsb_trace_batch_static: device ordinal out of range");
    std::lock_guard<std::mutex> hold(lock);
    if (!address[device]) {
        void* p = nullptr;
        GSB_CUDA(cudaGetSymbolAddress(&p, g_staticBatchScratch));
        if (!p) throw std::runtime_error("gsb_trace_batch_static: cudaGetSymbolAddress returned no address");
        address[device] = static_cast<unsigned char*>(p);
    }
    return address[device];
}

// transformInvT = glm::inverse(glm::transpose(M)) (PathTracer.cpp:62) in float32: glm's 4x4 inverse is the cofactor
// expansion of external/glm/glm/detail/func_matrix.inl:294-352; restated operation for operation so that the normal
// matrix carries the bits the reference's host code computes.  T: column-major object->world; out: rows of the 3x3 part.
static void normalMatrixGlm(const float* T, float out[3][3]) {
    float m[4][4];   // m[c][r] = transpose(M)[c][r] = M[r][c] = T[r*4 + c]
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) m[c][r] = T[r * 4 + c];
    const float c00 = m[2][2] * m[3][3] - m[3][2] * m[2][3], c02 = m[1][2] * m[3][3] - m[3][2] * m[1][3], c03 = m[1][2] * m[2][3] - m[2][2] * m[1][3];
    const float c04 = m[2][1] * m[3][3] - m[3][1] * m[2][3], c06 = m[1][1] * m[3][3] - m[3][1] * m[1][3], c07 = m[1][1] * m[2][3] - m[2][1] * m[1][3];
    const float c08 = m[2][1] * m[3][2] - m[3][1] * m[2][2], c10 = m[1][1] * m[3][2] - m[3][1] * m[1][2], c11 = m[1][1] * m[2][2] - m[2][1] * m[1][2];
    const float c12 = m[2][0] * m[3][3] - m[3][0] * m[2][3], c14 = m[1][0] * m[3][3] - m[3][0] * m[1][3], c15 = m[1][0] * m[2][3] - m[2][0] * m[1][3];
    const float c16 = m[2][0] * m[3][2] - m[3][0] * m[2][2], c18 = m[1][0] * m[3][2] - m[3][0] * m[1][2], c19 = m[1][0] * m[2][2] - m[2][0] * m[1][2];
    const float c20 = m[2][0] * m[3][1] - m[3][0] * m[2][1], c22 = m[1][0] * m[3][1] - m[3][0] * m[1][1], c23 = m[1][0] * m[2][1] - m[2][0] * m[1][1];
    const float f0[4] = {c00, c00, c02, c03}, f1[4] = {c04, c04, c06, c07}, f2[4] = {c08, c08, c10, c11};
    const float f3[4] = {c12, c12, c14, c15}, f4[4] = {c16, c16, c18, c19}, f5[4] = {c20, c20, c22, c23};
    const float v0[4] = {m[1][0], m[0][0], m[0][0], m[0][0]}, v1[4] = {m[1][1], m[0][1], m[0][1], m[0][1]};
    const float v2[4] = {m[1][2], m[0][2], m[0][2], m[0][2]}, v3[4] = {m[1][3], m[0][3], m[0][3], m[0][3]};
    const float sa[4] = {1.0f, -1.0f, 1.0f, -1.0f}, sb[4] = {-1.0f, 1.0f, -1.0f, 1.0f};
    volatile float inv4[4][4];   // volatile: keeps the host compiler from contracting a*b - c*d into fused operations
    for (int k = 0; k < 4; k++) {
        inv4[0][k] = ((v1[k] * f0[k] - v2[k] * f1[k]) + v3[k] * f2[k]) * sa[k];
        inv4[1][k] = ((v0[k] * f0[k] - v2[k] * f3[k]) + v3[k] * f4[k]) * sb[k];
        inv4[2][k] = ((v0[k] * f1[k] - v1[k] * f3[k]) + v3[k] * f5[k]) * sa[k];
        inv4[3][k] = ((v0[k] * f2[k] - v1[k] * f4[k]) + v2[k] * f5[k]) * sb[k];
    }
    const float d0 = m[0][0] * inv4[0][0], d1 = m[0][1] * inv4[1][0], d2 = m[0][2] * inv4[2][0], d3 = m[0][3] * inv4[3][0];
    const float oneOverDet = 1.0f / ((d0 + d1) + (d2 + d3));
    for (int r = 0; r < 3; r++)
        for (int k = 0; k < 3; k++) out[r][k] = inv4[k][r] * oneOverDet;
}

// Traversal variant per scene (traverse.cuh): rays that end after a handful of passes (a world level of a few nodes and
// nothing else: the Cornell box) are refilled straight from the source once 24 lanes are idle; longer-lived rays go
// through the per-warp ray pool, where a lane is refilled as soon as it is idle.  GSB_POOL = bit 0 extend, bit 1 shadow
// overrides; GSB_POOL_REFILL = idle lanes that trigger a refill out of the pool.
static void choosePolicy(gsb_context* ctx) {
    const bool longRays = ctx->numTraceInstances == 0 && ctx->world.numNodes > 16;   // two-level scenes: measured better without the pool (r2g / r2h)
    ctx->poolExtend = ctx->poolShadow = longRays;
    if (const char* e = getenv("GSB_POOL")) { const int v = atoi(e); ctx->poolExtend = (v & 1) != 0; ctx->poolShadow = (v & 2) != 0; }
    uint32_t poolRefill = 1u;
    if (const char* e = getenv("GSB_POOL_REFILL")) poolRefill = (uint32_t)std::min(32, std::max(1, atoi(e)));
    // idle lanes that trigger a refill straight from the source (measured, r2e: taking a ray is two loads since k_regen
    // starts the paths, so refilling early pays: Cornell 24 -> 12 lanes = -10 % k_extend; two-level scenes 8 for both kernels;
    // the any-hit kernel of world-only scenes 16).  GSB_REFILL / GSB_SHADOW_REFILL override.
    const bool twoLevel = ctx->numTraceInstances > 0;
    const uint32_t direct = ctx->directRefill ? ctx->directRefill : (twoLevel ? 8u : 12u);
    const uint32_t directShadow = ctx->shadowRefill ? ctx->shadowRefill : (twoLevel ? 8u : 16u);   // r3j: 24 -> 16, Cornell k_shadow 62.5 -> 59.1 ms, gallery flat
    ctx->refillMin = (ctx->refillMin & ~0xffu) | (ctx->poolExtend ? poolRefill : direct);
    // Phase quorum of the closest-hit kernel (traverse.cuh: a node / triangle round runs when at least `vote` lanes want it;
    // a triangle round also when at least `triMin` triangles wait): 2 and 8, and a quorum of 1 for a world level of a few
    // nodes, where the lanes are in step anyway and waiting for company only idles them (r3j, Cornell k_extend per step:
    // 4 / 4: 104.1 ms, 2 / 8: 101.9, 1 / 8: 100.4; gallery 198.7 / 195.9 / 200.2 (1 / 4); field 99.0 / 98.9 / 100.3 (1 / 4)).
    // GSB_VOTE / GSB_TRIMIN override.
    const bool tinyWorld = !twoLevel && ctx->world.numNodes <= 8u;
    if (!ctx->voteFromEnv) ctx->refillMin = (ctx->refillMin & ~0xff00u) | ((tinyWorld ? 1u : 2u) << 8);
    if (!ctx->triMinFromEnv) ctx->refillMin = (ctx->refillMin & ~0xff0000u) | (8u << 16);
    ctx->shadowControl = ctx->poolShadow ? poolRefill : directShadow;
    // The any-hit kernel runs without a lane quorum for its phases (a quorum of 1 or 2 cost Cornell's k_shadow 59.0 -> 60.3 /
    // 61.0 ms) but leaves its triangle loop when fewer than 4 lanes still have triangles waiting (r3k, k_shadow per step:
    // Cornell 59.0 -> 57.9 ms, gallery 110.5 -> 103.4, field 36.2 -> 34.6; 2 .. 6 are equal).  GSB_SHADOW_VOTE / GSB_SHADOW_TRIMIN.
    uint32_t shadowVote = 0u, shadowTriMin = 4u;
    if (const char* e = getenv("GSB_SHADOW_VOTE")) shadowVote = (uint32_t)std::min(32, std::max(0, atoi(e)));
    if (const char* e = getenv("GSB_SHADOW_TRIMIN")) shadowTriMin = (uint32_t)std::min(32, std::max(0, atoi(e)));
    ctx->shadowControl |= (shadowVote << 8) | (shadowTriMin << 16);
    // Ray sorting between bounces (wavefront.cu: sortKey): pays when the BVH is much larger than the caches, i.e. when
    // incoherent rays cost memory latency and lane divergence; GSB_SORT = grid bits per axis (0 = off) overrides.
    ctx->sortBits = ctx->numTraceInstances > 0 ? 6u : 0u;
    if (const char* e = getenv("GSB_SORT")) ctx->sortBits = (uint32_t)std::min(9, std::max(0, atoi(e)));
}

SceneDev gsb_context::sceneDev() const {
    SceneDev sc;
    memset(&sc, 0, sizeof(sc));
    sc.worldNodes = world.buffer;
    sc.worldTris = reinterpret_cast<const float4*>(world.buffer + world.leafOffset16);
    sc.worldNrm = dWorldNrm;
    sc.instSmall = dInstSmall;
    sc.tlasNodes = tlas.buffer;
    sc.tlasInst = reinterpret_cast<const TraceInstance*>(tlas.buffer + tlas.leafOffset16);
    sc.shadeInst = dShade;
    sc.instBsdf = dInstBsdf;
    for (int t = 0; t < GSB_BSDF_TYPE_COUNT; t++) sc.bsdf[t] = dBsdf[t];
    sc.lights = dLights;
    sc.numLights = numLights;
    sc.textures = dTextures;
    sc.numTextures = numTextures;
    sc.env = env;
    sc.numInstances = (uint32_t)instances.size();
    memcpy(sc.eye, eye, sizeof(eye));
    memcpy(sc.view, view, sizeof(view));
    sc.fov = fov;
    sc.params = params;
    sc.sortBits = sortBits;
    sc.sortMode = 0u;
    if (const char* e = getenv("GSB_SORT_MODE")) sc.sortMode = (uint32_t)std::min(2, std::max(0, atoi(e)));
    for (int a = 0; a < 3; a++) {
        // bounds of everything a ray can hit: the world level and the top level
        float lo = 1e30f, hi = -1e30f;
        if (world.numPrims) { lo = std::min(lo, world.lo[a]); hi = std::max(hi, world.hi[a]); }
        if (numTraceInstances) { lo = std::min(lo, tlas.lo[a]); hi = std::max(hi, tlas.hi[a]); }
        if (!(hi > lo)) { lo = 0.0f; hi = 1.0f; }
        sc.sortLo[a] = lo;
        sc.sortScale[a] = (float)(1u << sortBits) / (hi - lo);
    }
    return sc;
}

// Sampling tables of a latitude-longitude environment (shared rule with the checker): per row the running sum of the texel
// luminances (Rec. 709 weights, after the per-channel scale), normalised; over the rows the running sum of
// row total x sin(theta at the row centre), normalised.  Accumulated in double, stored as float.
static void buildEnvironmentTables(const gsb_environment& e, std::vector<float>& rowCdf, std::vector<float>& colCdf) {
    const uint32_t W = e.width, H = e.height;
    rowCdf.assign((size_t)H + 1, 0.0f);
    colCdf.assign((size_t)H * (W + 1), 0.0f);
    std::vector<double> rowW(H);
    double total = 0.0;
    for (uint32_t y = 0; y < H; y++) {
        std::vector<double> c((size_t)W + 1, 0.0);
        for (uint32_t x = 0; x < W; x++) {
            const float* p = e.texels + 3 * ((size_t)y * W + x);
            const double lum = 0.2126 * (double)(p[0] * e.scale[0]) + 0.7152 * (double)(p[1] * e.scale[1]) + 0.0722 * (double)(p[2] * e.scale[2]);
            c[x + 1] = c[x] + (lum > 0.0 ? lum : 0.0);
        }
        for (uint32_t x = 0; x <= W; x++) colCdf[(size_t)y * (W + 1) + x] = (float)(c[W] > 0.0 ? c[x] / c[W] : (double)x / W);
        colCdf[(size_t)y * (W + 1) + W] = 1.0f;
        rowW[y] = c[W] * std::sin(3.14159265358979323846 * (y + 0.5) / H);
        total += rowW[y];
    }
    double acc = 0.0;
    for (uint32_t y = 0; y < H; y++) {
        acc += rowW[y];
        rowCdf[y + 1] = (float)(total > 0.0 ? acc / total : (double)(y + 1) / H);
    }
    rowCdf[H] = 1.0f;
}

static void freeEnvironment(gsb_context* ctx) {
    cudaFree(ctx->dEnvTexels);
    cudaFree(ctx->dEnvRowCdf);
    cudaFree(ctx->dEnvColCdf);
    ctx->dEnvTexels = nullptr;
    ctx->dEnvRowCdf = ctx->dEnvColCdf = nullptr;
    memset(&ctx->env, 0, sizeof(ctx->env));
}

static void freeTextures(gsb_context* ctx) {
    for (float4* t : ctx->textureTexels) cudaFree(t);
    ctx->textureTexels.clear();
    cudaFree(ctx->dTextures);
    ctx->dTextures = nullptr;
    ctx->numTextures = 0;
}

extern "C" {

const char* gsb_last_error(void) { return g_lastError.c_str(); }
const char* gsb_version(void) { return "gpuspectral_b200 0.1 (sm_100a)"; }

gsb_status gsb_create(int device_ordinal, gsb_context** out_ctx) {
    if (!out_ctx) return fail(GSB_ERR_INVALID, "gsb_create: out_ctx is null");
    *out_ctx = nullptr;
    return guarded([&]() -> gsb_status {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0)
            return fail(GSB_ERR_CUDA, std::string("gsb_create: no CUDA device available (") + cudaGetErrorString(e) + ")");
        if (device_ordinal < 0 || device_ordinal >= count) return fail(GSB_ERR_INVALID, "gsb_create: device ordinal out of range");
        GSB_CUDA(cudaSetDevice(device_ordinal));
        gsb_context* ctx = new gsb_context();
        ctx->device = device_ordinal;
        GSB_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        {
            // keep freed stream-ordered allocations (BVH build temporaries, read-back staging) in the pool instead of
            // handing them back to the driver at every synchronisation
            cudaMemPool_t pool = nullptr;
            if (cudaDeviceGetDefaultMemPool(&pool, device_ordinal) == cudaSuccess) {
                unsigned long long keep = ~0ull;
                cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
            }
            cudaGetLastError();
        }
        ensureStats(ctx);
        if (const char* e = getenv("GSB_GRAPH")) ctx->useGraphLoop = atoi(e) != 0;
        if (const char* e = getenv("GSB_REFILL")) {
            const long v = atol(e);
            if (v >= 1 && v <= 32) { ctx->refillMin = (ctx->refillMin & ~0xffu) | (uint32_t)v; ctx->directRefill = (uint32_t)v; }
        }
        if (const char* e = getenv("GSB_SHADOW_REFILL")) {
            const long v = atol(e);
            if (v >= 1 && v <= 32) ctx->shadowRefill = (uint32_t)v;
        }
        if (const char* e = getenv("GSB_VOTE")) {
            const long v = atol(e);
            if (v >= 0 && v <= 32) { ctx->refillMin = (ctx->refillMin & ~0xff00u) | ((uint32_t)v << 8); ctx->voteFromEnv = true; }
        }
        if (const char* e = getenv("GSB_TRIMIN")) {
            const long v = atol(e);
            if (v >= 0 && v <= 32) { ctx->refillMin = (ctx->refillMin & ~0xff0000u) | ((uint32_t)v << 16); ctx->triMinFromEnv = true; }
        }
        if (const char* e = getenv("GSB_SHADE_DIRECT")) {
            const long v = atol(e);
            if (v >= 1 && v <= 33) ctx->shadeDirectMin = (uint32_t)v;
        }
        if (const char* e = getenv("GSB_SLOTS")) {
            const long v = atol(e);
            if (v >= 1024 && v <= (1l << 26)) ctx->targetSlots = (uint32_t)v;
        }
        *out_ctx = ctx;
        return GSB_OK;
    });
}

gsb_status gsb_destroy(gsb_context* ctx) {
    if (!ctx) return GSB_OK;
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        cudaStreamSynchronize(ctx->stream);
        for (MeshRec& m : ctx->meshes) {
            if (!m.alive) continue;
            cudaFree(m.vpos);
            cudaFree(m.vnrm);
            freeBvh(m.blas);
        }
        freeBvh(ctx->tlas);
        freeBvh(ctx->world);
        cudaFree(ctx->dWorldVerts);
        cudaFree(ctx->dWorldNrm);
        cudaFree(ctx->dNrmSoup);
        freeTextures(ctx);
        freeEnvironment(ctx);
        cudaFree(ctx->dInstSmall);
        cudaFree(ctx->dShade);
        cudaFree(ctx->dInstBsdf);
        for (void* p : ctx->dBsdf) cudaFree(p);
        cudaFree(ctx->dLights);
        cudaFree(ctx->dAccum);
        freeWavefront(ctx);
        if (ctx->ownStream) cudaStreamDestroy(ctx->stream);
        delete ctx;
        return GSB_OK;
    });
}

gsb_status gsb_mesh_create(gsb_context* ctx, const float* positions3, const float* normals3, uint32_t nverts, uint32_t* out_mesh) {
    return gsb_mesh_create_uv(ctx, positions3, normals3, nullptr, nverts, out_mesh);
}

gsb_status gsb_scene_set_environment(gsb_context* ctx, const gsb_environment* env) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_scene_set_environment: null context");
    if (env && env->type > GSB_ENVIRONMENT_LATLONG) return fail(GSB_ERR_INVALID, "gsb_scene_set_environment: unknown environment type");
    if (env && env->type == GSB_ENVIRONMENT_LATLONG && (!env->texels || env->width == 0 || env->height == 0))
        return fail(GSB_ERR_INVALID, "gsb_scene_set_environment: the latitude-longitude image has no texels");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        freeEnvironment(ctx);
        if (!env || env->type == GSB_ENVIRONMENT_NONE) return GSB_OK;
        EnvDev& d = ctx->env;
        d.type = env->type; d.width = env->width; d.height = env->height;
        for (int k = 0; k < 3; k++) d.scale[k] = env->scale[k];
        for (int k = 0; k < 9; k++) d.rot[k] = env->to_world[k];
        d.distance = 1.0e5f;
        if (env->type == GSB_ENVIRONMENT_LATLONG) {
            const size_t n = (size_t)env->width * env->height;
            std::vector<float4> px(n);
            for (size_t k = 0; k < n; k++) px[k] = make_float4(env->texels[3 * k], env->texels[3 * k + 1], env->texels[3 * k + 2], 0.0f);
            std::vector<float> rowCdf, colCdf;
            buildEnvironmentTables(*env, rowCdf, colCdf);
            ctx->dEnvTexels = uploadArray(px.data(), n, ctx->stream);
            ctx->dEnvRowCdf = uploadArray(rowCdf.data(), rowCdf.size(), ctx->stream);
            ctx->dEnvColCdf = uploadArray(colCdf.data(), colCdf.size(), ctx->stream);
            GSB_CUDA(cudaStreamSynchronize(ctx->stream));
            d.texels = ctx->dEnvTexels; d.rowCdf = ctx->dEnvRowCdf; d.colCdf = ctx->dEnvColCdf;
        }
        return GSB_OK;
    });
}

gsb_status gsb_scene_set_textures(gsb_context* ctx, const gsb_texture* textures, uint32_t count) {
    if (!ctx || (count && !textures)) return fail(GSB_ERR_INVALID, "gsb_scene_set_textures: null argument");
    for (uint32_t i = 0; i < count; i++) {
        if (textures[i].type > GSB_TEXTURE_BITMAP) return fail(GSB_ERR_INVALID, "gsb_scene_set_textures: unknown texture type");
        if (textures[i].type == GSB_TEXTURE_BITMAP && (!textures[i].texels || textures[i].width == 0 || textures[i].height == 0))
            return fail(GSB_ERR_INVALID, "gsb_scene_set_textures: bitmap " + std::to_string(i) + " has no texels");
    }
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        freeTextures(ctx);
        std::vector<TextureDev> recs(count);
        for (uint32_t i = 0; i < count; i++) {
            const gsb_texture& t = textures[i];
            TextureDev& r = recs[i];
            memset(&r, 0, sizeof(r));
            r.type = t.type; r.width = t.width; r.height = t.height;
            r.uscale = t.uscale; r.vscale = t.vscale; r.uoffset = t.uoffset; r.voffset = t.voffset;
            r.color0 = make_float4(t.color0[0], t.color0[1], t.color0[2], 0.0f);
            r.color1 = make_float4(t.color1[0], t.color1[1], t.color1[2], 0.0f);
            if (t.type == GSB_TEXTURE_BITMAP) {
                const size_t n = (size_t)t.width * t.height;
                std::vector<float4> px(n);
                for (size_t k = 0; k < n; k++) px[k] = make_float4(t.texels[3 * k], t.texels[3 * k + 1], t.texels[3 * k + 2], 0.0f);
                float4* d = uploadArray(px.data(), (uint32_t)n, ctx->stream);
                GSB_CUDA(cudaStreamSynchronize(ctx->stream));
                ctx->textureTexels.push_back(d);
                r.texels = d;
            }
        }
        ctx->dTextures = uploadArray(recs.data(), count, ctx->stream);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->numTextures = count;
        return GSB_OK;
    });
}

gsb_status gsb_mesh_create_uv(gsb_context* ctx, const float* positions3, const float* normals3, const float* uv2, uint32_t nverts,
                              uint32_t* out_mesh) {
    if (!ctx || !out_mesh) return fail(GSB_ERR_INVALID, "gsb_mesh_create: null argument");
    if (nverts % 3 != 0) return fail(GSB_ERR_INVALID, "gsb_mesh_create: vertex count is not a multiple of 3");
    if (nverts && (!positions3 || !normals3)) return fail(GSB_ERR_INVALID, "gsb_mesh_create: null vertex arrays");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        MeshRec m;
        m.alive = true;
        m.ntris = nverts / 3;
        std::vector<float4> p4(nverts), n4(nverts);
        for (uint32_t i = 0; i < nverts; i++) {
            // the texture coordinate rides in the two free .w components: u with the position, v with the normal
            p4[i] = make_float4(positions3[3 * (size_t)i], positions3[3 * (size_t)i + 1], positions3[3 * (size_t)i + 2], uv2 ? uv2[2 * (size_t)i] : 0.0f);
            n4[i] = make_float4(normals3[3 * (size_t)i], normals3[3 * (size_t)i + 1], normals3[3 * (size_t)i + 2], uv2 ? uv2[2 * (size_t)i + 1] : 0.0f);
        }
        m.vpos = uploadArray(p4.data(), nverts, ctx->stream);
        m.vnrm = uploadArray(n4.data(), nverts, ctx->stream);
        // The bottom level is built when the instance list first needs it (gsb_scene_set_instances): the meshes of a scene
        // are then built side by side in one batch, and a mesh that is only ever baked into the world level needs none.
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));   // p4 / n4 go out of scope
        uint32_t id = (uint32_t)ctx->meshes.size();
        for (uint32_t k = 0; k < ctx->meshes.size(); k++)
            if (!ctx->meshes[k].alive) { id = k; break; }
        if (id == ctx->meshes.size()) ctx->meshes.push_back(m);
        else ctx->meshes[id] = m;
        *out_mesh = id;
        return GSB_OK;
    });
}

gsb_status gsb_mesh_destroy(gsb_context* ctx, uint32_t mesh) {
    if (!ctx || mesh >= ctx->meshes.size() || !ctx->meshes[mesh].alive) return fail(GSB_ERR_INVALID, "gsb_mesh_destroy: bad mesh handle");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        for (const gsb_instance& in : ctx->instances)
            if (in.mesh == mesh) {
                ctx->haveInstances = false;  // the scene referenced it: force a new gsb_scene_set_instances
                break;
            }
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        MeshRec& m = ctx->meshes[mesh];
        cudaFree(m.vpos);
        cudaFree(m.vnrm);
        freeBvh(m.blas);
        m = MeshRec();
        return GSB_OK;
    });
}

gsb_status gsb_scene_set_bsdfs(gsb_context* ctx, uint32_t type, const void* params, uint32_t count) {
    if (!ctx || type >= GSB_BSDF_TYPE_COUNT) return fail(GSB_ERR_INVALID, "gsb_scene_set_bsdfs: bad type");
    if (count && !params) return fail(GSB_ERR_INVALID, "gsb_scene_set_bsdfs: null params");
    if (count > 65536) return fail(GSB_ERR_INVALID, "gsb_scene_set_bsdfs: more than 65536 records of one type (16-bit handle index)");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->dBsdf[type]);
        ctx->dBsdf[type] = uploadArray((const uint8_t*)params, (size_t)count * kBsdfStride[type], ctx->stream);
        ctx->bsdfCount[type] = count;
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_scene_set_lights(gsb_context* ctx, const gsb_triangle_light* lights, uint32_t count) {
    if (!ctx || (count && !lights)) return fail(GSB_ERR_INVALID, "gsb_scene_set_lights: null argument");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        cudaFree(ctx->dLights);
        ctx->dLights = uploadArray(lights, count, ctx->stream);
        ctx->numLights = count;
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

// Everything the kernels keep per instance, derived from its gsb_instance (PathTracer.cpp:58-75): shading record,
// the 16-byte record of a baked hit, and for a TLAS instance its trace record and padded world bounds.
struct InstanceRecords {
    ShadeInstance sh;
    float4 small;
    TraceInstance tr;
    float wlo[3], whi[3];
    uint32_t instWord;
};

static void instanceRecords(const gsb_context* ctx, uint32_t i, bool baked, InstanceRecords& out) {
    const gsb_instance& in = ctx->instances[i];
    const MeshRec& mesh = ctx->meshes[in.mesh];
    const float* T = in.transform;  // column-major: T[c*4 + r]
    double M[3][4];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 4; c++) M[r][c] = T[c * 4 + r];
    // inverse of the linear part by cofactors, float64, rounded once to float32
    const double c00 = M[1][1] * M[2][2] - M[1][2] * M[2][1], c01 = M[1][2] * M[2][0] - M[1][0] * M[2][2],
                 c02 = M[1][0] * M[2][1] - M[1][1] * M[2][0];
    const double det = M[0][0] * c00 + M[0][1] * c01 + M[0][2] * c02;
    const double id = 1.0 / det;
    double inv[3][3];
    inv[0][0] = c00 * id;
    inv[1][0] = c01 * id;
    inv[2][0] = c02 * id;
    inv[0][1] = (M[0][2] * M[2][1] - M[0][1] * M[2][2]) * id;
    inv[1][1] = (M[0][0] * M[2][2] - M[0][2] * M[2][0]) * id;
    inv[2][1] = (M[0][1] * M[2][0] - M[0][0] * M[2][1]) * id;
    inv[0][2] = (M[0][1] * M[1][2] - M[0][2] * M[1][1]) * id;
    inv[1][2] = (M[0][2] * M[1][0] - M[0][0] * M[1][2]) * id;
    inv[2][2] = (M[0][0] * M[1][1] - M[0][1] * M[1][0]) * id;
    // transformInvT = glm::inverse(glm::transpose(M)) in float32 as the reference's host computes it (PathTracer.cpp:62)
    float nrm[3][3];
    normalMatrixGlm(T, nrm);
    ShadeInstance& sh = out.sh;
    for (int r = 0; r < 3; r++) {
        sh.o2w[r] = make_float4((float)M[r][0], (float)M[r][1], (float)M[r][2], (float)M[r][3]);
        sh.nrm[r] = make_float4(nrm[r][0], nrm[r][1], nrm[r][2], in.emission[r]);
    }
    sh.vpos = mesh.vpos;
    sh.vnrm = mesh.vnrm;
    sh.bsdf = in.bsdf;
    sh.twofaced = in.twofaced;
    sh.pad[0] = sh.pad[1] = 0;
    const uint32_t smallBits = (in.bsdf & 0xffffu) | (std::min(in.twofaced, 0xffffu) << 16);
    out.small = make_float4(in.emission[0], in.emission[1], in.emission[2], 0.0f);
    memcpy(&out.small.w, &smallBits, sizeof(float));
    out.instWord = i | ((in.bsdf >> 16) << kInstTypeShift) | (baked ? kInstBaked : 0u);
    if (baked) return;
    TraceInstance& tr = out.tr;
    for (int r = 0; r < 3; r++) {
        const double tw = -(inv[r][0] * M[0][3] + inv[r][1] * M[1][3] + inv[r][2] * M[2][3]);
        tr.w2o[r] = make_float4((float)inv[r][0], (float)inv[r][1], (float)inv[r][2], (float)tw);
    }
    tr.blas = mesh.blas.buffer;
    // meshes of <= 4 triangles are traced without their (single) node: the TLAS box test already covered them
    tr.triOffset16 = mesh.blas.leafOffset16 | ((mesh.ntris >= 1 && mesh.ntris <= 4) ? (mesh.ntris << kDirectShift) : 0u);
    tr.instId = out.instWord;
    // padded world bounds of the padded object box
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int corner = 0; corner < 8; corner++) {
        const double p[3] = {(corner & 1) ? mesh.blas.hi[0] : mesh.blas.lo[0], (corner & 2) ? mesh.blas.hi[1] : mesh.blas.lo[1],
                             (corner & 4) ? mesh.blas.hi[2] : mesh.blas.lo[2]};
        for (int r = 0; r < 3; r++) {
            const double w = M[r][0] * p[0] + M[r][1] * p[1] + M[r][2] * p[2] + M[r][3];
            lo[r] = std::fmin(lo[r], w);
            hi[r] = std::fmax(hi[r], w);
        }
    }
    double scale = 0.0;
    for (int a = 0; a < 3; a++) scale = std::fmax(scale, std::fmax(std::fabs(lo[a]), std::fabs(hi[a])));
    const double pad = scale * (double)kBoxPadRel;
    for (int a = 0; a < 3; a++) {
        out.wlo[a] = std::nextafter((float)(lo[a] - pad), -INFINITY);
        out.whi[a] = std::nextafter((float)(hi[a] + pad), INFINITY);
    }
}

static void bakeInstance(gsb_context* ctx, uint32_t i, const InstanceRecords& r) {
    const MeshRec& mesh = ctx->meshes[ctx->instances[i].mesh];
    k_bakeInstance<<<(mesh.ntris + 127) / 128, 128, 0, ctx->stream>>>(mesh.vpos, mesh.vnrm, mesh.ntris, r.sh.o2w[0], r.sh.o2w[1], r.sh.o2w[2],
                                                                      r.sh.nrm[0], r.sh.nrm[1], r.sh.nrm[2], r.instWord,
                                                                      ctx->dWorldVerts + 3 * ctx->bakeOffset[i], ctx->dNrmSoup + 3 * ctx->bakeOffset[i]);
}

// Builds everything that depends on the instance list (ctx->instances): per-instance records, the world level over the
// baked instances, the top level over the others.  What VulkanRays::buildTLAS + PathTracer::prepareScene do per frame.
static void rebuildScene(gsb_context* ctx) {
    const uint32_t count = (uint32_t)ctx->instances.size();
    std::vector<ShadeInstance> shade(count);
    std::vector<uint32_t> handles(count);
    std::vector<float4> small(count);
    // Bake rule (shared with the checker: kBakeMaxTris / kBakeBudget): in gl_InstanceID order, an instance
    // whose mesh has at most kBakeMaxTris triangles goes to the world level while the baked total fits the budget.
    ctx->bakeOffset.assign(count, ~0ull);   // first world-level triangle of a baked instance
    ctx->traceIndex.assign(count, -1);
    ctx->hostTrace.clear();
    ctx->hostWlo.clear();
    ctx->hostWhi.clear();
    uint64_t baked = 0;
    for (uint32_t i = 0; i < count; i++) {
        const uint32_t nt = ctx->meshes[ctx->instances[i].mesh].ntris;
        if (nt > 0 && nt <= kBakeMaxTris && baked + nt <= kBakeBudget) { ctx->bakeOffset[i] = baked; baked += nt; }
    }
    {   // bottom levels still missing for the meshes that are traced through one: all of them in one batch
        std::vector<BlasBuildRequest> todo;
        std::vector<bool> queued(ctx->meshes.size(), false);
        for (uint32_t i = 0; i < count; i++) {
            const uint32_t mi = ctx->instances[i].mesh;
            MeshRec& m = ctx->meshes[mi];
            if (ctx->bakeOffset[i] != ~0ull || m.blas.buffer || queued[mi]) continue;
            queued[mi] = true;
            todo.push_back(BlasBuildRequest{m.vpos, m.ntris, &m.blas, false});
        }
        if (!todo.empty()) buildBlasBatch(todo.data(), todo.size(), ctx->stream);
    }
    cudaFree(ctx->dWorldVerts);
    cudaFree(ctx->dWorldNrm);
    cudaFree(ctx->dNrmSoup);
    ctx->dWorldVerts = ctx->dWorldNrm = ctx->dNrmSoup = nullptr;
    if (baked) {
        void* q = nullptr;
        GSB_CUDA(cudaMalloc(&q, (size_t)baked * 3 * sizeof(float4)));
        ctx->dWorldVerts = (float4*)q;
        GSB_CUDA(cudaMalloc(&q, (size_t)baked * 3 * sizeof(float4)));
        ctx->dWorldNrm = (float4*)q;
        GSB_CUDA(cudaMalloc(&q, (size_t)baked * 3 * sizeof(float4)));
        ctx->dNrmSoup = (float4*)q;   // world normals in bake order, gathered into leaf order after the build
    }
    for (bool& u : ctx->typeUsed) u = false;
    for (uint32_t i = 0; i < count; i++) {
        InstanceRecords r;
        const bool isBaked = ctx->bakeOffset[i] != ~0ull;
        instanceRecords(ctx, i, isBaked, r);
        shade[i] = r.sh;
        small[i] = r.small;
        handles[i] = ctx->instances[i].bsdf;
        ctx->typeUsed[ctx->instances[i].bsdf >> 16] = true;
        if (isBaked) { bakeInstance(ctx, i, r); continue; }
        ctx->traceIndex[i] = (int32_t)ctx->hostTrace.size();
        ctx->hostTrace.push_back(r.tr);
        for (int a = 0; a < 3; a++) { ctx->hostWlo.push_back(r.wlo[a]); ctx->hostWhi.push_back(r.whi[a]); }
    }
    GSB_CUDA(cudaGetLastError());
    cudaFree(ctx->dShade);
    cudaFree(ctx->dInstBsdf);
    ctx->dShade = uploadArray(shade.data(), count, ctx->stream);
    ctx->dInstBsdf = uploadArray(handles.data(), count, ctx->stream);
    cudaFree(ctx->dInstSmall);
    ctx->dInstSmall = uploadArray(small.data(), count, ctx->stream);
    freeBvh(ctx->world);
    // the treelet setting that won the world level's comparison is kept while the level keeps its size
    if (ctx->worldSettingTris != baked) { ctx->worldSetting = -1; ctx->worldSettingTris = baked; }
    buildBlas(ctx->dWorldVerts, (uint32_t)baked, ctx->stream, ctx->world, true, &ctx->worldSetting);
    if (baked) k_gatherNormals<<<((uint32_t)baked + 127) / 128, 128, 0, ctx->stream>>>(ctx->world.leafPrim, (uint32_t)baked, ctx->dNrmSoup, ctx->dWorldNrm);
    GSB_CUDA(cudaStreamSynchronize(ctx->stream));
    freeBvh(ctx->tlas);
    ctx->numTraceInstances = (uint32_t)ctx->hostTrace.size();
    ctx->numBakedTriangles = (uint32_t)baked;
    buildTlas(ctx->hostWlo.data(), ctx->hostWhi.data(), ctx->hostTrace.data(), ctx->numTraceInstances, ctx->stream, ctx->tlas);
    GSB_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->haveInstances = true;
    choosePolicy(ctx);
    double ms = ctx->tlas.buildMs + ctx->world.buildMs;
    std::vector<bool> seen(ctx->meshes.size(), false);
    for (uint32_t i = 0; i < count; i++) {
        const uint32_t m = ctx->instances[i].mesh;
        if (ctx->bakeOffset[i] == ~0ull && !seen[m]) { seen[m] = true; ms += ctx->meshes[m].blas.buildMs; }
    }
    ctx->lastBuildMs = ms;
    ctx->lastRefitMs = 0.0;
    ctx->refits = 0;
}

gsb_status gsb_scene_set_instances(gsb_context* ctx, const gsb_instance* instances, uint32_t count) {
    if (!ctx || (count && !instances)) return fail(GSB_ERR_INVALID, "gsb_scene_set_instances: null argument");
    if (count > kMaxInstances) return fail(GSB_ERR_INVALID, "gsb_scene_set_instances: more than 2^27 - 1 instances");
    for (uint32_t i = 0; i < count; i++) {
        if (instances[i].mesh >= ctx->meshes.size() || !ctx->meshes[instances[i].mesh].alive)
            return fail(GSB_ERR_INVALID, "gsb_scene_set_instances: instance " + std::to_string(i) + " references an unknown mesh");
        if ((instances[i].bsdf >> 16) >= GSB_BSDF_TYPE_COUNT)
            return fail(GSB_ERR_INVALID, "gsb_scene_set_instances: instance " + std::to_string(i) + " has an unknown BSDF type");
    }
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        ctx->instances.assign(instances, instances + count);
        rebuildScene(ctx);
        return GSB_OK;
    });
}

// Moved instances (animated scenes; the reference rebuilds its TLAS from {blas, mat4} every frame, PathTracer.cpp:10-19):
// the transforms of instances [first, first + count) are replaced, their records recomputed, baked ones re-baked into the
// world level, and both levels REFIT over their existing topology (bvh.h: refitTriangles / refitTlas).  When the summed
// node area of a refit level exceeds kRefitRebuildRatio x its value at build time the scene is rebuilt instead.
gsb_status gsb_scene_update_transforms(gsb_context* ctx, uint32_t first, uint32_t count, const float* transforms) {
    if (!ctx || (!transforms && count)) return fail(GSB_ERR_INVALID, "gsb_scene_update_transforms: null argument");
    if (!ctx->haveInstances) return fail(GSB_ERR_STATE, "gsb_scene_update_transforms: no scene (call gsb_scene_set_instances first)");
    if ((uint64_t)first + count > ctx->instances.size()) return fail(GSB_ERR_INVALID, "gsb_scene_update_transforms: instance range out of bounds");
    if (count == 0) return GSB_OK;
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        constexpr double kRefitRebuildRatio = 2.0;
        const char* force = getenv("GSB_REFIT");      // "0": always rebuild (A/B measurements)
        if (force && atoi(force) == 0) {
            for (uint32_t k = 0; k < count; k++) memcpy(ctx->instances[first + k].transform, transforms + 16 * (size_t)k, 16 * sizeof(float));
            rebuildScene(ctx);
            return GSB_OK;
        }
        // the yardstick of both levels, measured once on the trees as built
        auto worldScale = [&]() {
            float sc = 0.0f;
            for (int a = 0; a < 3; a++) sc = std::fmax(sc, std::fmax(std::fabs(ctx->world.lo[a]), std::fabs(ctx->world.hi[a])));
            return sc;
        };
        if (ctx->world.numPrims && ctx->world.areaAtBuild == 0.0) ctx->world.areaAtBuild = refitTriangles(ctx->world, worldScale(), ctx->stream, true);
        if (ctx->numTraceInstances && ctx->tlas.areaAtBuild == 0.0)
            ctx->tlas.areaAtBuild = refitTlas(ctx->tlas, ctx->hostWlo.data(), ctx->hostWhi.data(), ctx->hostTrace.data(), ctx->numTraceInstances, ctx->stream, true);
        bool touchedWorld = false, touchedTlas = false;
        for (uint32_t k = 0; k < count; k++) {
            const uint32_t i = first + k;
            memcpy(ctx->instances[i].transform, transforms + 16 * (size_t)k, 16 * sizeof(float));
            InstanceRecords r;
            const bool isBaked = ctx->bakeOffset[i] != ~0ull;
            instanceRecords(ctx, i, isBaked, r);
            GSB_CUDA(cudaMemcpyAsync(ctx->dShade + i, &r.sh, sizeof(ShadeInstance), cudaMemcpyHostToDevice, ctx->stream));
            if (isBaked) {
                bakeInstance(ctx, i, r);
                touchedWorld = true;
            } else {
                const int32_t t = ctx->traceIndex[i];
                ctx->hostTrace[t] = r.tr;
                for (int a = 0; a < 3; a++) { ctx->hostWlo[3 * t + a] = r.wlo[a]; ctx->hostWhi[3 * t + a] = r.whi[a]; }
                touchedTlas = true;
            }
        }
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));   // `r` lives on the stack: its copies must have left
        double ms = 0.0;
        bool degraded = false;
        if (touchedWorld) {
            // the baked triangles moved: leaf records and normals again from the soup (same leaf order), boxes bottom-up.
            // The padding follows the largest coordinate of the level, like at build time.
            const uint32_t n = ctx->world.numPrims;
            repackTriangles(ctx->world, ctx->dWorldVerts, true, ctx->stream);
            k_gatherNormals<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->world.leafPrim, n, ctx->dNrmSoup, ctx->dWorldNrm);
            const float sc = soupScale(ctx->dWorldVerts, n * 3, ctx->stream);
            const double area = refitTriangles(ctx->world, sc, ctx->stream);
            ms += ctx->world.refitMs;
            degraded = degraded || area > kRefitRebuildRatio * ctx->world.areaAtBuild;
        }
        if (touchedTlas) {
            const double area = refitTlas(ctx->tlas, ctx->hostWlo.data(), ctx->hostWhi.data(), ctx->hostTrace.data(), ctx->numTraceInstances, ctx->stream);
            ms += ctx->tlas.refitMs;
            degraded = degraded || area > kRefitRebuildRatio * ctx->tlas.areaAtBuild;
        }
        if (degraded) {
            rebuildScene(ctx);
            return GSB_OK;
        }
        ctx->lastRefitMs = ms;
        ctx->refits++;
        return GSB_OK;
    });
}

gsb_status gsb_set_camera(gsb_context* ctx, const float eye[3], const float to_world[16], float fov_radians) {
    if (!ctx || !eye || !to_world) return fail(GSB_ERR_INVALID, "gsb_set_camera: null argument");
    memcpy(ctx->eye, eye, sizeof(float) * 3);
    memcpy(ctx->view, to_world, sizeof(float) * 16);
    ctx->fov = fov_radians;
    return GSB_OK;
}

gsb_status gsb_default_params(gsb_params* out) {
    if (!out) return fail(GSB_ERR_INVALID, "gsb_default_params: null argument");
    out->max_depth = 50;
    out->rr_start = 10;
    out->cutoff = 20.0f;
    out->nee = 1;
    out->pixel_filter = 0;
    return GSB_OK;
}

gsb_status gsb_set_params(gsb_context* ctx, const gsb_params* params) {
    if (!ctx || !params) return fail(GSB_ERR_INVALID, "gsb_set_params: null argument");
    if (params->max_depth < 0 || params->max_depth > 65000) return fail(GSB_ERR_INVALID, "gsb_set_params: max_depth out of range");
    if (params->pixel_filter < 0 || params->pixel_filter > 2) return fail(GSB_ERR_INVALID, "gsb_set_params: pixel_filter must be 0 (none), 1 (box) or 2 (tent)");
    ctx->params = *params;
    return GSB_OK;
}

gsb_status gsb_get_bvh_stats(gsb_context* ctx, gsb_bvh_stats* out) {
    if (!ctx || !out) return fail(GSB_ERR_INVALID, "gsb_get_bvh_stats: null argument");
    memset(out, 0, sizeof(*out));
    std::vector<bool> seen(ctx->meshes.size(), false), blasUsed(ctx->meshes.size(), false);
    uint64_t baked = 0;
    for (const gsb_instance& in : ctx->instances) {
        const MeshRec& m = ctx->meshes[in.mesh];
        out->instanced_triangles += m.ntris;
        if (!seen[in.mesh]) {
            seen[in.mesh] = true;
            out->num_meshes++;
            out->unique_triangles += m.ntris;
        }
        // the bake rule of gsb_scene_set_instances: a baked instance does not go through its mesh's own BLAS
        if (m.ntris > 0 && m.ntris <= kBakeMaxTris && baked + m.ntris <= kBakeBudget) {
            baked += m.ntris;
            out->baked_instances++;
        } else if (!blasUsed[in.mesh]) {
            blasUsed[in.mesh] = true;
            out->blas_nodes += m.blas.numNodes;
            out->bvh_bytes += m.blas.bytes;
            out->sah_cost_before += m.blas.sahBefore;
            out->sah_cost_after += m.blas.sahAfter;
        }
    }
    out->num_instances = ctx->instances.size();
    out->tlas_nodes = ctx->tlas.numNodes;
    out->bvh_bytes += ctx->tlas.bytes + ctx->world.bytes;
    out->baked_triangles = ctx->numBakedTriangles;
    out->world_nodes = ctx->world.numNodes;
    out->sah_cost_before += ctx->world.sahBefore;
    out->sah_cost_after += ctx->world.sahAfter;
    out->build_ms = ctx->lastBuildMs;
    out->refits = ctx->refits;
    out->refit_ms = ctx->lastRefitMs;
    return GSB_OK;
}

gsb_status gsb_render(gsb_context* ctx, uint32_t width, uint32_t height, uint32_t first_sample, uint32_t num_samples, uint32_t sample_stride) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_render: null context");
    if (width == 0 || height == 0 || (uint64_t)width * height > (1ull << 28)) return fail(GSB_ERR_INVALID, "gsb_render: bad frame size");
    if (sample_stride == 0) return fail(GSB_ERR_INVALID, "gsb_render: sample_stride must be >= 1");
    if (!sceneReady(ctx)) return fail(GSB_ERR_STATE, "gsb_render: no scene (call gsb_scene_set_instances first)");
    for (const gsb_instance& in : ctx->instances)
        if ((in.bsdf & 0xffffu) >= ctx->bsdfCount[in.bsdf >> 16])
            return fail(GSB_ERR_STATE, "gsb_render: an instance references a BSDF record that was not uploaded");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        renderImpl(ctx, width, height, first_sample, num_samples, sample_stride);
        if (ctx->pending.empty()) checkOverflow(ctx);   // host-driven loop: the call has completed
        return GSB_OK;
    });
}

gsb_status gsb_clear(gsb_context* ctx) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_clear: null context");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        if (ctx->dAccum) GSB_CUDA(cudaMemsetAsync(ctx->dAccum, 0, (size_t)ctx->width * ctx->height * sizeof(float4), ctx->stream));
        GSB_CUDA(cudaMemsetAsync(ctx->wf.stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
        ctx->seenIters = 0;
        ctx->launches = 0;
        return GSB_OK;
    });
}

gsb_status gsb_synchronize(gsb_context* ctx) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_synchronize: null context");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

static gsb_status readImage(gsb_context* ctx, float* hdr, uint8_t* ldr) {
    if (!ctx || (!hdr && !ldr)) return fail(GSB_ERR_INVALID, "gsb_read: null argument");
    if (!ctx->dAccum) return fail(GSB_ERR_STATE, "gsb_read: nothing rendered yet");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        const uint32_t npix = ctx->width * ctx->height;
        Scratch sc(ctx->stream);
        float4* dHdr = hdr ? sc.alloc<float4>(npix) : nullptr;
        uchar4* dLdr = ldr ? sc.alloc<uchar4>(npix) : nullptr;
        k_resolve<<<(npix + 255) / 256, 256, 0, ctx->stream>>>(ctx->dAccum, npix, dHdr, dLdr);
        if (hdr) GSB_CUDA(cudaMemcpyAsync(hdr, dHdr, (size_t)npix * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
        if (ldr) GSB_CUDA(cudaMemcpyAsync(ldr, dLdr, (size_t)npix * sizeof(uchar4), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_read_hdr(gsb_context* ctx, float* out_rgba) { return readImage(ctx, out_rgba, nullptr); }
gsb_status gsb_read_ldr(gsb_context* ctx, uint8_t* out_rgba8) { return readImage(ctx, nullptr, out_rgba8); }

gsb_status gsb_read_sum(gsb_context* ctx, float* out_rgba_sum) {
    if (!ctx || !out_rgba_sum) return fail(GSB_ERR_INVALID, "gsb_read_sum: null argument");
    if (!ctx->dAccum) return fail(GSB_ERR_STATE, "gsb_read_sum: nothing rendered yet");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaMemcpyAsync(out_rgba_sum, ctx->dAccum, (size_t)ctx->width * ctx->height * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_write_sum(gsb_context* ctx, uint32_t width, uint32_t height, const float* rgba_sum) {
    if (!ctx || !rgba_sum) return fail(GSB_ERR_INVALID, "gsb_write_sum: null argument");
    if (width == 0 || height == 0 || (uint64_t)width * height > (1ull << 28)) return fail(GSB_ERR_INVALID, "gsb_write_sum: bad frame size");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        renderImpl(ctx, width, height, 0, 0, 1);   // count 0: (re)allocates the frame, renders nothing
        GSB_CUDA(cudaMemcpyAsync(ctx->dAccum, rgba_sum, (size_t)width * height * sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_device_sum_ptr(gsb_context* ctx, void** out_ptr, size_t* out_bytes) {
    if (!ctx || !out_ptr || !out_bytes) return fail(GSB_ERR_INVALID, "gsb_device_sum_ptr: null argument");
    if (!ctx->dAccum) return fail(GSB_ERR_STATE, "gsb_device_sum_ptr: nothing rendered yet");
    *out_ptr = ctx->dAccum;
    *out_bytes = (size_t)ctx->width * ctx->height * sizeof(float4);
    return GSB_OK;
}

gsb_status gsb_get_counters(gsb_context* ctx, gsb_counters* out) {
    if (!ctx || !out) return fail(GSB_ERR_INVALID, "gsb_get_counters: null argument");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        unsigned long long h[ST_COUNT];
        GSB_CUDA(cudaMemcpyAsync(h, ctx->wf.stats, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        out->paths = h[ST_PATHS];
        out->extension_rays = h[ST_EXT];
        out->shadow_rays = h[ST_SHADOW];
        out->iterations = h[ST_ITERS];
        out->nodes_visited = h[ST_NODES];
        out->tris_tested = h[ST_TRIS];
        out->inst_entered = h[ST_INSTS];
        out->rays_counted = h[ST_RAYS];
        out->shaded_vertices = h[ST_VERTICES];
        return GSB_OK;
    });
}

gsb_status gsb_reset_counters(gsb_context* ctx) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_reset_counters: null context");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaMemsetAsync(ctx->wf.stats, 0, ST_COUNT * sizeof(unsigned long long), ctx->stream));
        ctx->seenIters = 0;
        ctx->launches = 0;
        return GSB_OK;
    });
}

gsb_status gsb_set_profiling(gsb_context* ctx, int enabled) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_set_profiling: null context");
    ctx->profiling = enabled != 0;
    return GSB_OK;
}

gsb_status gsb_set_instrumented(gsb_context* ctx, int mode) {
    if (!ctx || mode < 0 || mode > 3) return fail(GSB_ERR_INVALID, "gsb_set_instrumented: bad argument");
    ctx->instrumented = mode;
    return GSB_OK;
}

gsb_status gsb_set_stream(gsb_context* ctx, void* cuda_stream) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_set_stream: null context");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        if (ctx->ownStream) cudaStreamDestroy(ctx->stream);
        if (cuda_stream) {
            ctx->stream = (cudaStream_t)cuda_stream;
            ctx->ownStream = false;
        } else {
            GSB_CUDA(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
            ctx->ownStream = true;
        }
        return GSB_OK;
    });
}

gsb_status gsb_get_launch_info(gsb_context* ctx, uint64_t* out_launches, uint64_t* out_last_render_iterations) {
    if (!ctx) return fail(GSB_ERR_INVALID, "gsb_get_launch_info: null context");
    const gsb_status st = guarded([&]() -> gsb_status { Bind bind(ctx->device); drain(ctx); return GSB_OK; });
    if (st != GSB_OK) return st;
    if (out_launches) *out_launches = ctx->launches;
    if (out_last_render_iterations) *out_last_render_iterations = ctx->lastRenderIters;
    return GSB_OK;
}

gsb_status gsb_last_render_ms(gsb_context* ctx, double out_ms[5]) {
    if (!ctx || !out_ms) return fail(GSB_ERR_INVALID, "gsb_last_render_ms: null argument");
    const gsb_status st = guarded([&]() -> gsb_status { Bind bind(ctx->device); drain(ctx); return GSB_OK; });
    if (st != GSB_OK) return st;
    for (int k = 0; k < 5; k++) out_ms[k] = ctx->lastRenderMs[k];
    return GSB_OK;
}

static gsb_status traceBatch(gsb_context* ctx, const gsb_ray* rays, uint32_t count, gsb_hit* out_hits, int any_hit, bool stats,
                             int repeats, double* out_ms) {
    if (!ctx || (count && !rays)) return fail(GSB_ERR_INVALID, "gsb_trace_batch: null argument");
    if (!sceneReady(ctx)) return fail(GSB_ERR_STATE, "gsb_trace_batch: no scene (call gsb_scene_set_instances first)");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        drain(ctx);
        Scratch scr;
        gsb_ray* dRays = scr.upload(rays, count, ctx->stream);
        gsb_hit* dHits = scr.alloc<gsb_hit>(count);
        const SceneDev sc = ctx->sceneDev();
        int sms = 148, perSm = 4;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        const bool hasInst = ctx->numTraceInstances > 0;
        using BatchKernel = void (*)(SceneDev, const gsb_ray*, uint32_t, gsb_hit*, unsigned long long*, uint32_t);
        static const BatchKernel kernels[16] = {
            k_traceBatch<false, false, false, false>, k_traceBatch<false, false, true, false>, k_traceBatch<false, true, false, false>, k_traceBatch<false, true, true, false>,
            k_traceBatch<true, false, false, false>,  k_traceBatch<true, false, true, false>,  k_traceBatch<true, true, false, false>,  k_traceBatch<true, true, true, false>,
            k_traceBatch<false, false, false, true>,  k_traceBatch<false, false, true, true>,  k_traceBatch<false, true, false, true>,  k_traceBatch<false, true, true, true>,
            k_traceBatch<true, false, false, true>,   k_traceBatch<true, false, true, true>,   k_traceBatch<true, true, false, true>,   k_traceBatch<true, true, true, true>};
        // the batch entry runs the same traversal variant the render kernels of this scene use (pool or direct refill)
        const bool pool = any_hit ? ctx->poolShadow : ctx->poolExtend;
        const BatchKernel kernel = kernels[(pool ? 8 : 0) | (any_hit ? 4 : 0) | (stats ? 2 : 0) | (hasInst ? 1 : 0)];
        const uint32_t control = any_hit ? ctx->shadowControl : ctx->refillMin;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, (const void*)kernel, kTraceBlock, 0);
        const int grid = std::max(1, std::min((int)((count + kTraceBlock - 1) / kTraceBlock), sms * std::max(perSm, 1)));
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (out_ms) {
            GSB_CUDA(cudaEventCreate(&e0));
            GSB_CUDA(cudaEventCreate(&e1));
        }
        for (int r = 0; r < repeats + (out_ms ? 1 : 0); r++) {
            if (out_ms && r == 1) GSB_CUDA(cudaEventRecord(e0, ctx->stream));  // launch 0 is a warm-up
            if (count == 0) break;
            GSB_CUDA(cudaMemsetAsync(&ctx->wf.stats[ST_BATCHNEXT], 0, sizeof(unsigned long long), ctx->stream));  // the stream's work counter
            kernel<<<grid, kTraceBlock, 0, ctx->stream>>>(sc, dRays, count, dHits, ctx->wf.stats, control);
        }
        GSB_CUDA(cudaGetLastError());
        if (out_ms) {
            GSB_CUDA(cudaEventRecord(e1, ctx->stream));
            GSB_CUDA(cudaEventSynchronize(e1));
            float ms = 0.0f;
            GSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            *out_ms = repeats > 0 ? (double)ms / repeats : 0.0;
            cudaEventDestroy(e0);
            cudaEventDestroy(e1);
        }
        if (out_hits && count) GSB_CUDA(cudaMemcpyAsync(out_hits, dHits, (size_t)count * sizeof(gsb_hit), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        checkOverflow(ctx);
        return GSB_OK;
    });
}

gsb_status gsb_trace_batch(gsb_context* ctx, const gsb_ray* rays, uint32_t count, gsb_hit* out_hits, int any_hit) {
    if (count && !out_hits) return fail(GSB_ERR_INVALID, "gsb_trace_batch: null output");
    return traceBatch(ctx, rays, count, out_hits, any_hit, false, 1, nullptr);
}

gsb_status gsb_trace_batch_stats(gsb_context* ctx, const gsb_ray* rays, uint32_t count, gsb_hit* out_hits, int any_hit) {
    if (count && !out_hits) return fail(GSB_ERR_INVALID, "gsb_trace_batch_stats: null output");
    return traceBatch(ctx, rays, count, out_hits, any_hit, true, 1, nullptr);
}

gsb_status gsb_trace_batch_timed(gsb_context* ctx, const gsb_ray* rays, uint32_t count, int any_hit, int repeats, double* out_ms_per_launch) {
    if (!out_ms_per_launch || repeats < 1) return fail(GSB_ERR_INVALID, "gsb_trace_batch_timed: bad argument");
    return traceBatch(ctx, rays, count, nullptr, any_hit, false, repeats, out_ms_per_launch);
}

// No counterpart in the reference.  gsb_trace_batch with its scratch
// (device rays, device hits, work counter) in g_staticBatchScratch instead of cudaMalloc'ed memory: no cudaMalloc /
// cudaFree / cudaMallocAsync, no event or stream creation, no allocating library call.  All calls on a device share
// the array, so calls that overlap on the GPU (two contexts on different streams, two threads) race; the control block
// is cleared at the start of every call, rays and hits are fully overwritten for the `count` entries used.
gsb_status gsb_trace_batch_static(gsb_context* ctx, const gsb_ray* rays, uint32_t count, gsb_hit* out_hits, int any_hit) {
    if (!ctx || (count && (!rays || !out_hits))) return fail(GSB_ERR_INVALID, "gsb_trace_batch_static: null argument");
    if (count > GSB_STATIC_BATCH_MAX_RAYS)
        return fail(GSB_ERR_INVALID, "gsb_trace_batch_static: " + std::to_string(count) + " rays exceed the static scratch of " +
                                         std::to_string((unsigned)GSB_STATIC_BATCH_MAX_RAYS) + " rays (GSB_STATIC_BATCH_MAX_RAYS); split the batch or use gsb_trace_batch");
    if (!sceneReady(ctx)) return fail(GSB_ERR_STATE, "gsb_trace_batch_static: no scene (call gsb_scene_set_instances first)");
    if (!ctx->pending.empty())
        return fail(GSB_ERR_STATE, "gsb_trace_batch_static: an asynchronous gsb_render is pending (call gsb_synchronize first)");
    if (count == 0) return GSB_OK;
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        unsigned char* base = staticBatchScratch(ctx->device);
        unsigned long long* control = reinterpret_cast<unsigned long long*>(base);
        gsb_ray* dRays = reinterpret_cast<gsb_ray*>(base + kStaticRaysOffset);
        gsb_hit* dHits = reinterpret_cast<gsb_hit*>(base + kStaticHitsOffset);
        cudaStream_t s = ctx->stream;
        GSB_CUDA(cudaMemsetAsync(control, 0, kStaticControlBytes, s));   // work counter and overflow flag start at zero
        GSB_CUDA(cudaMemcpyAsync(dRays, rays, (size_t)count * sizeof(gsb_ray), cudaMemcpyHostToDevice, s));
        const SceneDev sc = ctx->sceneDev();
        const bool hasInst = ctx->numTraceInstances > 0;
        const bool pool = any_hit ? ctx->poolShadow : ctx->poolExtend;   // the traversal variant the render kernels of this scene use
        using BatchKernel = void (*)(SceneDev, const gsb_ray*, uint32_t, gsb_hit*, unsigned long long*, uint32_t);
        static const BatchKernel kernels[8] = {
            k_traceBatch<false, false, false, false>, k_traceBatch<false, false, true, false>, k_traceBatch<true, false, false, false>, k_traceBatch<true, false, true, false>,
            k_traceBatch<false, false, false, true>,  k_traceBatch<false, false, true, true>,  k_traceBatch<true, false, false, true>,  k_traceBatch<true, false, true, true>};
        const BatchKernel kernel = kernels[(pool ? 4 : 0) | (any_hit ? 2 : 0) | (hasInst ? 1 : 0)];
        int sms = 148, perSm = 4;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, (const void*)kernel, kTraceBlock, 0);
        const int grid = std::max(1, std::min((int)((count + kTraceBlock - 1) / kTraceBlock), sms * std::max(perSm, 1)));
        kernel<<<grid, kTraceBlock, 0, s>>>(sc, dRays, count, dHits, control, any_hit ? ctx->shadowControl : ctx->refillMin);
        GSB_CUDA(cudaGetLastError());
        unsigned long long overflow = 0;
        GSB_CUDA(cudaMemcpyAsync(out_hits, dHits, (size_t)count * sizeof(gsb_hit), cudaMemcpyDeviceToHost, s));
        GSB_CUDA(cudaMemcpyAsync(&overflow, control + ST_OVERFLOW, sizeof(overflow), cudaMemcpyDeviceToHost, s));
        GSB_CUDA(cudaStreamSynchronize(s));
        if (overflow) throw std::runtime_error("traversal stack overflow: BVH deeper than the traversal stack");
        return GSB_OK;
    });
}

// ---- diagnostics ------------------------------------------------------------------------------------
gsb_status gsb_test_rng(gsb_context* ctx, const uint32_t* idx, const uint32_t* timestamp, uint32_t n, uint32_t n_draws, uint32_t* out) {
    if (!ctx || !idx || !timestamp || !out) return fail(GSB_ERR_INVALID, "gsb_test_rng: null argument");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        Scratch scr;
        uint32_t* dIdx = scr.upload(idx, n, ctx->stream);
        uint32_t* dTs = scr.upload(timestamp, n, ctx->stream);
        uint32_t* dOut = scr.alloc<uint32_t>((size_t)n * (1 + n_draws));
        k_testRng<<<(n + 127) / 128, 128, 0, ctx->stream>>>(dIdx, dTs, n, n_draws, dOut);
        GSB_CUDA(cudaMemcpyAsync(out, dOut, (size_t)n * (1 + n_draws) * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_test_camera_rays(gsb_context* ctx, uint32_t width, uint32_t height, const uint32_t* px, const uint32_t* py, uint32_t n, gsb_ray* out) {
    if (!ctx || !px || !py || !out) return fail(GSB_ERR_INVALID, "gsb_test_camera_rays: null argument");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        Scratch scr;
        uint32_t* dx = scr.upload(px, n, ctx->stream);
        uint32_t* dy = scr.upload(py, n, ctx->stream);
        gsb_ray* dOut = scr.alloc<gsb_ray>(n);
        const float camZ = (std::fmax((float)width, (float)height) / 2.0f) / std::tan(ctx->fov / 2.0f);   // as renderImpl
        k_testCamera<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->sceneDev(), width, height, camZ, dx, dy, n, dOut);
        GSB_CUDA(cudaMemcpyAsync(out, dOut, (size_t)n * sizeof(gsb_ray), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

static bool handleOk(gsb_context* ctx, uint32_t handle) {
    return (handle >> 16) < GSB_BSDF_TYPE_COUNT && (handle & 0xffffu) < ctx->bsdfCount[handle >> 16];
}

gsb_status gsb_test_bsdf_sample(gsb_context* ctx, uint32_t handle, const float* wo, const uint32_t* seed, uint32_t n, float* wi, float* f,
                                float* pdf, uint32_t* is_delta, uint32_t* seed_out) {
    if (!ctx || !wo || !seed || !wi || !f || !pdf || !is_delta || !seed_out) return fail(GSB_ERR_INVALID, "gsb_test_bsdf_sample: null argument");
    if (!handleOk(ctx, handle)) return fail(GSB_ERR_INVALID, "gsb_test_bsdf_sample: BSDF handle out of range");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        Scratch scr;
        float* dWo = scr.upload(wo, 3 * (size_t)n, ctx->stream);
        uint32_t* dSeed = scr.upload(seed, n, ctx->stream);
        float* dWi = scr.alloc<float>(3 * (size_t)n);
        float* dF = scr.alloc<float>(3 * (size_t)n);
        float* dPdf = scr.alloc<float>(n);
        uint32_t* dDelta = scr.alloc<uint32_t>(n);
        uint32_t* dSeedOut = scr.alloc<uint32_t>(n);
        k_testBsdfSample<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->sceneDev(), handle, dWo, dSeed, n, dWi, dF, dPdf, dDelta, dSeedOut);
        GSB_CUDA(cudaMemcpyAsync(wi, dWi, 3 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(f, dF, 3 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(pdf, dPdf, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(is_delta, dDelta, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(seed_out, dSeedOut, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_test_bsdf_eval(gsb_context* ctx, uint32_t handle, const float* wo, const float* wi, uint32_t n, float* f, float* pdf, uint32_t* is_delta) {
    if (!ctx || !wo || !wi || !f || !pdf || !is_delta) return fail(GSB_ERR_INVALID, "gsb_test_bsdf_eval: null argument");
    if (!handleOk(ctx, handle)) return fail(GSB_ERR_INVALID, "gsb_test_bsdf_eval: BSDF handle out of range");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        Scratch scr;
        float* dWo = scr.upload(wo, 3 * (size_t)n, ctx->stream);
        float* dWi = scr.upload(wi, 3 * (size_t)n, ctx->stream);
        float* dF = scr.alloc<float>(3 * (size_t)n);
        float* dPdf = scr.alloc<float>(n);
        uint32_t* dDelta = scr.alloc<uint32_t>(n);
        k_testBsdfEval<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->sceneDev(), handle, dWo, dWi, n, dF, dPdf, dDelta);
        GSB_CUDA(cudaMemcpyAsync(f, dF, 3 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(pdf, dPdf, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(is_delta, dDelta, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

gsb_status gsb_test_sample_light(gsb_context* ctx, const float* pos, const uint32_t* seed, uint32_t n, float* lpos, float* emission, float* pdf,
                                 uint32_t* seed_out) {
    if (!ctx || !pos || !seed || !lpos || !emission || !pdf || !seed_out) return fail(GSB_ERR_INVALID, "gsb_test_sample_light: null argument");
    return guarded([&]() -> gsb_status {
        Bind bind(ctx->device);
        Scratch scr;
        float* dPos = scr.upload(pos, 3 * (size_t)n, ctx->stream);
        uint32_t* dSeed = scr.upload(seed, n, ctx->stream);
        float* dL = scr.alloc<float>(3 * (size_t)n);
        float* dE = scr.alloc<float>(3 * (size_t)n);
        float* dPdf = scr.alloc<float>(n);
        uint32_t* dSeedOut = scr.alloc<uint32_t>(n);
        k_testSampleLight<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->sceneDev(), dPos, dSeed, n, dL, dE, dPdf, dSeedOut);
        GSB_CUDA(cudaMemcpyAsync(lpos, dL, 3 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(emission, dE, 3 * (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(pdf, dPdf, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaMemcpyAsync(seed_out, dSeedOut, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
        GSB_CUDA(cudaStreamSynchronize(ctx->stream));
        return GSB_OK;
    });
}

}  // extern "C"

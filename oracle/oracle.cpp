// oracle.cpp — CPU restatement of the reference path tracer.  TEST INFRASTRUCTURE ONLY.
//
// Follows, function by function, the GLSL of the reference (paths relative to
// /root/reference/src/GPUSpectral/assets/shaders/):
//   pt_common.glsl   RNG (tea/pcgHash/randPcg/randUniform), Onb, payload and scene records
//   raygen.rgen      primary ray, bounce loop, cutoff, Russian roulette, running mean
//   rayhit.rchit     hit reconstruction, 8 BSDF models, triangle-light NEE, MIS, emission
//   miss.rmiss / shadowmiss.rmiss
// and the host glue of renderer/PathTracer.cpp:58-93 for the derived instance matrices.
//
// What has no source in the reference is the ray/triangle intersection itself (the Vulkan
// driver / RT hardware behind traceRayEXT): PARITY UNPINNED for that part.  The oracle
// DEFINES it as: per instance, the world ray is taken to object space with the inverse
// instance matrix (direction not renormalised, so t is shared with world space) — or, for
// instances of small meshes, the triangles are taken to world space once instead (the bake
// rule next to kBakeMaxTris) and the ray is used as it is — every
// triangle is tested with the watertight test of Woop, Benthin & Wald (JCGT 2013) in
// float32 with individually rounded operations (no FMA contraction), candidates need
// tmin < t < tmax (Vulkan's triangle interval), and the closest hit is the minimum of
// (t, instance, primitive) in lexicographic order.  Mode 2 of or_trace_batch is an
// independent float64 Moeller-Trumbore brute force used to bound the error of that rule.
//
// All shading arithmetic is float32 like the shaders.  RNG draws are sequenced explicitly
// left to right (GLSL evaluates constructor/call arguments in order; C++ does not).
//
// PINNED (everything except the intersection rule above): oracle/_ref/libgsref.so is the reference's own shader text
// compiled as C++ against the reference's glm (oracle/Makefile, oracle/ref/); tests/test_oracle_vs_reference_cpu.py
// requires this file to equal it BIT FOR BIT — RNG, every BSDF sample / eval through the reference's own switch, light
// sampling, camera rays, per-bounce payloads and whole frames (running mean included) — live on >= 1e5 inputs per
// function, and against frozen outputs of that library (tests/golden/ref_vectors.npz) where the reference is absent.
// Vector helpers therefore use glm's formulas (normalize = v * (1 / sqrt(dot(v, v))), min / max as GLSL defines them,
// the cofactor 4x4 inverse for transformInvT).
//
// Build: see oracle/Makefile (-O2 -ffp-contract=off: nothing here may be fused).

#include "oracle.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ---------------------------------------------------------------------------------------
// GLSL-flavoured vector helpers
// ---------------------------------------------------------------------------------------
struct vec3 {
    float x, y, z;
    vec3() : x(0), y(0), z(0) {}
    explicit vec3(float s) : x(s), y(s), z(s) {}
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
inline vec3 operator+(vec3 a, vec3 b) { return vec3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline vec3 operator-(vec3 a, vec3 b) { return vec3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline vec3 operator-(vec3 a) { return vec3(-a.x, -a.y, -a.z); }
inline vec3 operator*(vec3 a, vec3 b) { return vec3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline vec3 operator*(vec3 a, float s) { return vec3(a.x * s, a.y * s, a.z * s); }
inline vec3 operator*(float s, vec3 a) { return vec3(s * a.x, s * a.y, s * a.z); }
inline vec3 operator/(vec3 a, float s) { return vec3(a.x / s, a.y / s, a.z / s); }
inline vec3 operator/(vec3 a, vec3 b) { return vec3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline vec3 operator+(vec3 a, float s) { return vec3(a.x + s, a.y + s, a.z + s); }
inline vec3 operator-(float s, vec3 a) { return vec3(s - a.x, s - a.y, s - a.z); }
inline vec3 operator-(vec3 a, float s) { return vec3(a.x - s, a.y - s, a.z - s); }
inline float dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) {
    return vec3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y);
}
inline float length(vec3 a) { return std::sqrt(dot(a, a)); }
// glm::normalize (external/glm/glm/detail/func_geometric.inl:82-90): v * inversesqrt(dot(v, v)), inversesqrt = 1 / sqrt
// (func_exponential.inl:134-139) — the formula the reference's host math library uses, one division and three products
inline vec3 normalize(vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
inline vec3 vsqrt(vec3 a) { return vec3(std::sqrt(a.x), std::sqrt(a.y), std::sqrt(a.z)); }
// GLSL faceforward(N, I, Nref): dot(Nref, I) < 0 ? N : -N
inline vec3 faceforward(vec3 N, vec3 I, vec3 Nref) { return dot(Nref, I) < 0.0f ? N : -N; }
// GLSL 4.60 section 8.3: max(x, y) "returns y if x < y, otherwise x", min(x, y) "returns y if y < x, otherwise x",
// clamp = min(max(x, lo), hi) — also glm's definitions (func_common.inl:17-30,505-509).  They differ from fmax / fmin
// only when an operand is NaN (the first operand then wins), which the shader text reaches at grazing directions.
inline float gmax(float x, float y) { return (x < y) ? y : x; }
inline float gmin(float x, float y) { return (y < x) ? y : x; }
inline float clampf(float v, float lo, float hi) { return gmin(gmax(v, lo), hi); }

// GLSL "#define M_PI 3.14159265358979323846" is a 32-bit float literal in the shaders.
constexpr float kPi = 3.14159265358979323846f;

// ---------------------------------------------------------------------------------------
// RNG — pt_common.glsl:85-120
// ---------------------------------------------------------------------------------------
struct Rng {
    uint32_t state;
    uint64_t draws = 0;
    explicit Rng(uint32_t s) : state(s) {}
    // randPcg, pt_common.glsl:87-93
    uint32_t randPcg() {
        uint32_t st = state;
        state = state * 747796405u + 2891336453u;
        uint32_t word = ((st >> ((st >> 28u) + 4u)) ^ st) * 277803737u;
        ++draws;
        return (word >> 22u) ^ word;
    }
    // randUniform, pt_common.glsl:102-104.  float(0xffffffffu) rounds to 2^32, so the scale is
    // exactly 2^-32 and the result lies in [0, 1] inclusive.
    float randUniform() { return (float)randPcg() * (1.0f / (float)0xffffffffu); }
};

// pcgHash, pt_common.glsl:95-100
inline uint32_t pcgHash(uint32_t v) {
    uint32_t state = v * 747796405u + 2891336453u;
    uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
    return (word >> 22u) ^ word;
}

// tea, pt_common.glsl:106-120
inline uint32_t tea(uint32_t val0, uint32_t val1) {
    uint32_t v0 = val0, v1 = val1, s0 = 0;
    for (uint32_t n = 0; n < 4; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}

// ---------------------------------------------------------------------------------------
// Onb — pt_common.glsl:122-151
// ---------------------------------------------------------------------------------------
struct Onb { vec3 tangent, binormal, normal; };
inline Onb onbCreate(vec3 n) {
    Onb onb;
    onb.normal = normalize(n);
    if (std::fabs(onb.normal.x) > std::fabs(onb.normal.z))
        onb.binormal = vec3(-onb.normal.y, onb.normal.x, 0.0f);
    else
        onb.binormal = vec3(0.0f, -onb.normal.z, onb.normal.y);
    onb.binormal = normalize(onb.binormal);
    onb.tangent = cross(onb.binormal, onb.normal);
    return onb;
}
inline vec3 onbUntransform(const Onb& o, vec3 v) { return o.tangent * v.x + o.binormal * v.y + o.normal * v.z; }
inline vec3 onbTransform(const Onb& o, vec3 v) { return vec3(dot(v, o.tangent), dot(v, o.binormal), dot(v, o.normal)); }

// ---------------------------------------------------------------------------------------
// Scene records
// ---------------------------------------------------------------------------------------
struct BvhNode {           // oracle's own CPU BVH (binary, object space)
    float lo[3], hi[3];
    int32_t left, right;   // children, or -1
    uint32_t first, count; // triangle range in `order` when leaf
};

struct Mesh {
    std::vector<float> pos;   // 3 floats / vertex
    std::vector<float> nrm;
    std::vector<float> uv;    // 2 floats / vertex; empty = all zero (texture extension)
    uint32_t ntris = 0;
    std::vector<BvhNode> nodes;
    std::vector<uint32_t> order;
    float lo[3], hi[3];       // padded object bounds
};

struct Instance {
    gsb_instance src;
    float o2w[12];            // rows of the 3x4 object->world matrix (gl_ObjectToWorldEXT)
    float w2o[12];            // rows of the 3x4 world->object matrix
    float nrmM[9];            // upper-left 3x3 of inverse(transpose(M)), row-major (PathTracer.cpp:62)
    double wlo[3], whi[3];    // padded world bounds
    std::shared_ptr<Mesh> world;  // baked instance: its triangles in world space (+ their BVH); null = traced in object space
};

// Bake rule (shared with the product, gpuspectral_b200/csrc/api.cu): walking the instances in gl_InstanceID order, an
// instance whose mesh has at most kBakeMaxTris triangles is baked — its vertices are taken to world space once,
// x' = ((m0*x + m1*y) + m2*z) + m3 per row in float32 (the operation order of gl_ObjectToWorldEXT * vec4(p, 1) in
// rayhit.rchit:679-681), and rays meet them untransformed — as long as the baked total stays within kBakeBudget.
constexpr uint32_t kBakeMaxTris = 16384;
constexpr uint64_t kBakeBudget = 1ull << 22;

}  // namespace

struct or_scene {
    std::vector<Mesh> meshes;
    std::vector<Instance> instances;
    std::vector<gsb_diffuse_bsdf> diffuse;
    std::vector<gsb_smooth_dielectric_bsdf> smoothDielectric;
    std::vector<gsb_smooth_conductor_bsdf> smoothConductor;
    std::vector<gsb_smooth_plastic_bsdf> smoothPlastic;
    std::vector<gsb_rough_conductor_bsdf> roughConductor;
    std::vector<gsb_smooth_floor_bsdf> smoothFloor;
    std::vector<gsb_rough_floor_bsdf> roughFloor;
    std::vector<gsb_rough_plastic_bsdf> roughPlastic;
    std::vector<gsb_rough_dielectric_bsdf> roughDielectric;   // extension, see roughDielectricSample
    std::vector<gsb_triangle_light> lights;
    // reflectance textures (extension; gsb_texture of the public header with the texels copied)
    struct Texture { gsb_texture desc; std::vector<float> texels; };
    std::vector<Texture> textures;
    // environment emitter (extension; gsb_environment of the public header with the texels copied and the sampling tables)
    struct Environment {
        gsb_environment desc = {};
        std::vector<float> texels, rowCdf, colCdf;
        float distance = 1.0e5f;
    } env;
    float eye[3] = {0, 0, 0};
    float view[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    float fov = 1.0f;
    gsb_params params = {50, 10, 20.0f, 1, 0};
};

namespace {

// ---------------------------------------------------------------------------------------
// Intersection (defined by the oracle; see header comment)
// ---------------------------------------------------------------------------------------
struct Hit {
    float t, u, v;
    uint32_t inst, prim;
};

struct ObjRay {           // object-space ray + Woop precomputation
    float o[3], d[3];
    int kx, ky, kz;
    float Sx, Sy, Sz;
};

// M = nullptr: the ray is used as it is (baked instances, whose triangles are stored in world space)
inline void makeObjRay(const float* M /*3x4 rows*/, const float* o, const float* d, ObjRay& r) {
    for (int i = 0; i < 3; i++) {
        if (!M) { r.o[i] = o[i]; r.d[i] = d[i]; continue; }
        const float* m = M + 4 * i;
        r.o[i] = ((m[0] * o[0] + m[1] * o[1]) + m[2] * o[2]) + m[3];
        r.d[i] = (m[0] * d[0] + m[1] * d[1]) + m[2] * d[2];
    }
    float ax = std::fabs(r.d[0]), ay = std::fabs(r.d[1]), az = std::fabs(r.d[2]);
    int kz = (ax > ay) ? ((ax > az) ? 0 : 2) : ((ay > az) ? 1 : 2);
    int kx = kz + 1; if (kx == 3) kx = 0;
    int ky = kx + 1; if (ky == 3) ky = 0;
    if (r.d[kz] < 0.0f) std::swap(kx, ky);
    r.kx = kx; r.ky = ky; r.kz = kz;
    r.Sx = r.d[kx] / r.d[kz];
    r.Sy = r.d[ky] / r.d[kz];
    r.Sz = 1.0f / r.d[kz];
}

// Watertight ray/triangle test, float32, every operation individually rounded.
// Returns true and (t,u,v) when the ray's line crosses the triangle (no interval test here).
inline bool woopTest(const ObjRay& r, const float* v0, const float* v1, const float* v2, float& t, float& bu, float& bv) {
    const float A[3] = {v0[0] - r.o[0], v0[1] - r.o[1], v0[2] - r.o[2]};
    const float B[3] = {v1[0] - r.o[0], v1[1] - r.o[1], v1[2] - r.o[2]};
    const float C[3] = {v2[0] - r.o[0], v2[1] - r.o[1], v2[2] - r.o[2]};
    const float Ax = A[r.kx] - r.Sx * A[r.kz], Ay = A[r.ky] - r.Sy * A[r.kz];
    const float Bx = B[r.kx] - r.Sx * B[r.kz], By = B[r.ky] - r.Sy * B[r.kz];
    const float Cx = C[r.kx] - r.Sx * C[r.kz], Cy = C[r.ky] - r.Sy * C[r.kz];
    float U = Cx * By - Cy * Bx;
    float V = Ax * Cy - Ay * Cx;
    float W = Bx * Ay - By * Ax;
    if (U == 0.0f || V == 0.0f || W == 0.0f) {
        double CxBy = (double)Cx * (double)By, CyBx = (double)Cy * (double)Bx;
        U = (float)(CxBy - CyBx);
        double AxCy = (double)Ax * (double)Cy, AyCx = (double)Ay * (double)Cx;
        V = (float)(AxCy - AyCx);
        double BxAy = (double)Bx * (double)Ay, ByAx = (double)By * (double)Ax;
        W = (float)(BxAy - ByAx);
    }
    if ((U < 0.0f || V < 0.0f || W < 0.0f) && (U > 0.0f || V > 0.0f || W > 0.0f)) return false;
    const float det = (U + V) + W;
    if (det == 0.0f) return false;
    const float Az = r.Sz * A[r.kz], Bz = r.Sz * B[r.kz], Cz = r.Sz * C[r.kz];
    const float T = (U * Az + V * Bz) + W * Cz;
    t = T / det;
    bu = V / det;
    bv = W / det;
    return true;
}

// closest-hit candidate acceptance: tmin < t < tmax, and (t, inst, prim) lexicographic minimum
inline void offer(Hit& best, float tmin, float tmax, float t, float u, float v, uint32_t inst, uint32_t prim) {
    if (!(t > tmin && t < tmax)) return;
    if (t < best.t || (t == best.t && (inst < best.inst || (inst == best.inst && prim < best.prim)))) {
        best.t = t; best.u = u; best.v = v; best.inst = inst; best.prim = prim;
    }
}

constexpr float kPadRel = 1.0f / 32768.0f;  // box padding relative to the coordinate scale

void buildMeshBvh(Mesh& m) {
    const uint32_t n = m.ntris;
    m.order.resize(n);
    for (uint32_t i = 0; i < n; i++) m.order[i] = i;
    std::vector<float> cen(3 * (size_t)n);
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = 0; i < n; i++) {
        const float* p = &m.pos[9 * (size_t)i];
        for (int a = 0; a < 3; a++) {
            float mn = std::fmin(p[a], std::fmin(p[3 + a], p[6 + a]));
            float mx = std::fmax(p[a], std::fmax(p[3 + a], p[6 + a]));
            cen[3 * (size_t)i + a] = 0.5f * (mn + mx);
            lo[a] = std::fmin(lo[a], mn);
            hi[a] = std::fmax(hi[a], mx);
        }
    }
    float scale = 0.0f;
    for (int a = 0; a < 3; a++) scale = std::fmax(scale, std::fmax(std::fabs(lo[a]), std::fabs(hi[a])));
    const float pad = scale * kPadRel;
    for (int a = 0; a < 3; a++) { m.lo[a] = lo[a] - pad; m.hi[a] = hi[a] + pad; }
    m.nodes.clear();
    if (n == 0) return;
    struct Work { int32_t node; uint32_t first, count; };
    std::vector<Work> stack;
    m.nodes.push_back(BvhNode{});
    stack.push_back({0, 0, n});
    while (!stack.empty()) {
        Work w = stack.back();
        stack.pop_back();
        float blo[3] = {INFINITY, INFINITY, INFINITY}, bhi[3] = {-INFINITY, -INFINITY, -INFINITY};
        float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = w.first; i < w.first + w.count; i++) {
            const float* p = &m.pos[9 * (size_t)m.order[i]];
            for (int a = 0; a < 3; a++) {
                blo[a] = std::fmin(blo[a], std::fmin(p[a], std::fmin(p[3 + a], p[6 + a])));
                bhi[a] = std::fmax(bhi[a], std::fmax(p[a], std::fmax(p[3 + a], p[6 + a])));
                clo[a] = std::fmin(clo[a], cen[3 * (size_t)m.order[i] + a]);
                chi[a] = std::fmax(chi[a], cen[3 * (size_t)m.order[i] + a]);
            }
        }
        BvhNode nd;
        for (int a = 0; a < 3; a++) { nd.lo[a] = blo[a] - pad; nd.hi[a] = bhi[a] + pad; }
        nd.left = nd.right = -1;
        nd.first = w.first;
        nd.count = w.count;
        if (w.count > 4) {
            int axis = 0;
            if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
            if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
            uint32_t mid = w.count / 2;
            std::nth_element(m.order.begin() + w.first, m.order.begin() + w.first + mid,
                             m.order.begin() + w.first + w.count, [&](uint32_t a, uint32_t b) {
                                 float ca = cen[3 * (size_t)a + axis], cb = cen[3 * (size_t)b + axis];
                                 return ca < cb || (ca == cb && a < b);
                             });
            nd.left = (int32_t)m.nodes.size();
            nd.right = nd.left + 1;
            m.nodes.push_back(BvhNode{});
            m.nodes.push_back(BvhNode{});
            stack.push_back({nd.left, w.first, mid});
            stack.push_back({nd.right, w.first + mid, w.count - mid});
        }
        m.nodes[w.node] = nd;
    }
}

// slab test in float64 on the padded float boxes: conservative for the float32 triangle test
inline bool slab(const double* lo, const double* hi, const double* o, const double* d, double tmin, double tmax) {
    double t0 = tmin, t1 = tmax;
    for (int a = 0; a < 3; a++) {
        if (d[a] == 0.0) {
            if (o[a] < lo[a] || o[a] > hi[a]) return false;
        } else {
            double ta = (lo[a] - o[a]) / d[a], tb = (hi[a] - o[a]) / d[a];
            if (ta > tb) std::swap(ta, tb);
            if (ta > t0) t0 = ta;
            if (tb < t1) t1 = tb;
            if (t0 > t1) return false;
        }
    }
    return true;
}

void traceInstanceBrute(const Mesh& m, const ObjRay& r, uint32_t inst, float tmin, float tmax, Hit& best, bool anyHit, bool& occluded) {
    for (uint32_t p = 0; p < m.ntris; p++) {
        const float* v = &m.pos[9 * (size_t)p];
        float t, u, w;
        if (!woopTest(r, v, v + 3, v + 6, t, u, w)) continue;
        if (anyHit) {
            if (t > tmin && t < tmax) { occluded = true; return; }
        } else {
            offer(best, tmin, tmax, t, u, w, inst, p);
        }
    }
}

void traceInstanceBvh(const Mesh& m, const ObjRay& r, uint32_t inst, float tmin, float tmax, Hit& best, bool anyHit, bool& occluded) {
    if (m.nodes.empty()) return;
    const double o[3] = {r.o[0], r.o[1], r.o[2]}, d[3] = {r.d[0], r.d[1], r.d[2]};
    int32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const BvhNode& nd = m.nodes[stack[--sp]];
        const double lo[3] = {nd.lo[0], nd.lo[1], nd.lo[2]}, hi[3] = {nd.hi[0], nd.hi[1], nd.hi[2]};
        const double far = anyHit ? (double)tmax : (double)best.t;  // best.t starts at tmax; ties at best.t still accepted
        if (!slab(lo, hi, o, d, (double)tmin, far)) continue;
        if (nd.left < 0) {
            for (uint32_t i = nd.first; i < nd.first + nd.count; i++) {
                const uint32_t p = m.order[i];
                const float* v = &m.pos[9 * (size_t)p];
                float t, u, w;
                if (!woopTest(r, v, v + 3, v + 6, t, u, w)) continue;
                if (anyHit) {
                    if (t > tmin && t < tmax) { occluded = true; return; }
                } else {
                    offer(best, tmin, tmax, t, u, w, inst, p);
                }
            }
        } else {
            stack[sp++] = nd.left;
            stack[sp++] = nd.right;
        }
    }
}

// float32 trace (mode 0 brute force, mode 1 BVH)
Hit traceF32(const or_scene& s, const float* o, const float* d, float tmin, float tmax, bool anyHit, bool useBvh) {
    Hit best;
    best.t = tmax; best.u = best.v = 0.0f; best.inst = best.prim = GSB_NO_HIT;
    bool occluded = false;
    const double od[3] = {o[0], o[1], o[2]}, dd[3] = {d[0], d[1], d[2]};
    for (uint32_t i = 0; i < (uint32_t)s.instances.size(); i++) {
        const Instance& in = s.instances[i];
        const Mesh& m = in.world ? *in.world : s.meshes[in.src.mesh];
        if (useBvh) {
            const double far = anyHit ? (double)tmax : (double)best.t;
            if (!slab(in.wlo, in.whi, od, dd, (double)tmin, far)) continue;
        }
        ObjRay r;
        makeObjRay(in.world ? nullptr : in.w2o, o, d, r);
        if (useBvh) traceInstanceBvh(m, r, i, tmin, tmax, best, anyHit, occluded);
        else traceInstanceBrute(m, r, i, tmin, tmax, best, anyHit, occluded);
        if (occluded) { best.inst = 0; best.prim = 0; best.t = 0.0f; return best; }
    }
    return best;
}

// float64 Moeller-Trumbore brute force in world space (mode 2)
Hit traceF64(const or_scene& s, const float* o, const float* d, float tmin, float tmax, bool anyHit) {
    Hit best;
    best.t = tmax; best.u = best.v = 0.0f; best.inst = best.prim = GSB_NO_HIT;
    double bt = tmax;
    const double O[3] = {o[0], o[1], o[2]}, D[3] = {d[0], d[1], d[2]};
    for (uint32_t i = 0; i < (uint32_t)s.instances.size(); i++) {
        const Instance& in = s.instances[i];
        const Mesh& m = s.meshes[in.src.mesh];
        for (uint32_t p = 0; p < m.ntris; p++) {
            double P[3][3];
            for (int k = 0; k < 3; k++) {
                const float* v = &m.pos[9 * (size_t)p + 3 * k];
                for (int r = 0; r < 3; r++) {
                    const float* row = in.o2w + 4 * r;
                    P[k][r] = (double)row[0] * v[0] + (double)row[1] * v[1] + (double)row[2] * v[2] + (double)row[3];
                }
            }
            double e1[3], e2[3], pv[3], tv[3], qv[3];
            for (int a = 0; a < 3; a++) { e1[a] = P[1][a] - P[0][a]; e2[a] = P[2][a] - P[0][a]; }
            pv[0] = D[1] * e2[2] - D[2] * e2[1]; pv[1] = D[2] * e2[0] - D[0] * e2[2]; pv[2] = D[0] * e2[1] - D[1] * e2[0];
            double det = e1[0] * pv[0] + e1[1] * pv[1] + e1[2] * pv[2];
            if (det == 0.0) continue;
            double inv = 1.0 / det;
            for (int a = 0; a < 3; a++) tv[a] = O[a] - P[0][a];
            double u = (tv[0] * pv[0] + tv[1] * pv[1] + tv[2] * pv[2]) * inv;
            if (u < 0.0 || u > 1.0) continue;
            qv[0] = tv[1] * e1[2] - tv[2] * e1[1]; qv[1] = tv[2] * e1[0] - tv[0] * e1[2]; qv[2] = tv[0] * e1[1] - tv[1] * e1[0];
            double v = (D[0] * qv[0] + D[1] * qv[1] + D[2] * qv[2]) * inv;
            if (v < 0.0 || u + v > 1.0) continue;
            double t = (e2[0] * qv[0] + e2[1] * qv[1] + e2[2] * qv[2]) * inv;
            if (!(t > (double)tmin && t < (double)tmax)) continue;
            if (anyHit) { best.inst = 0; best.prim = 0; best.t = 0.0f; return best; }
            if (t < bt) {
                bt = t;
                best.t = (float)t; best.u = (float)u; best.v = (float)v; best.inst = i; best.prim = p;
            }
        }
    }
    return best;
}

// ---------------------------------------------------------------------------------------
// Sampling helpers — rayhit.rchit:89-210
// ---------------------------------------------------------------------------------------
struct vec2 { float x, y; };

// sampleConcentric, rayhit.rchit:89-105
inline vec2 sampleConcentric(Rng& rng) {
    float s0 = rng.randUniform();
    float s1 = rng.randUniform();
    vec2 u = {2.0f * s0 - 1.0f, 2.0f * s1 - 1.0f};
    if (u.x == 0.0f && u.y == 0.0f) return vec2{0.0f, 0.0f};
    float r, th;
    if (std::fabs(u.x) > std::fabs(u.y)) {
        r = u.x;
        th = (kPi / 4.0f) * (u.y / u.x);
    } else {
        r = u.y;
        th = kPi / 2.0f - (kPi / 4.0f) * (u.x / u.y);
    }
    return vec2{r * std::cos(th), r * std::sin(th)};
}

// randCosineHemisphere, rayhit.rchit:107-111
inline vec3 randCosineHemisphere(Rng& rng) {
    vec2 u = sampleConcentric(rng);
    float z = std::sqrt(gmax(0.0f, 1.0f - u.x * u.x - u.y * u.y));
    return vec3(u.x, u.y, z);
}

// cosineHemispherePdf, rayhit.rchit:113-115
inline float cosineHemispherePdf(vec3 w) { return gmax(std::fabs(w.z) / kPi, 0.000001f); }

// sampleHalf, rayhit.rchit:155-166
inline vec3 sampleHalf(Rng& rng, float alpha) {
    float ux = rng.randUniform();
    float uy = rng.randUniform();
    float phi = 2.0f * kPi * ux;
    float logSample = std::log(1.0f - uy);
    if (std::isinf(logSample)) logSample = 0.0f;
    float tan2 = -alpha * alpha * logSample;
    float cost = 1.0f / std::sqrt(1.0f + tan2);
    float sint = std::sqrt(gmax(0.0f, 1.0f - cost * cost));
    return vec3(std::cos(phi) * sint, std::sin(phi) * sint, cost);
}

// beckmannD, rayhit.rchit:177-183
inline float beckmannD(vec3 wh, float alpha) {
    float cos2 = wh.z * wh.z;
    float tan2 = (wh.x * wh.x + wh.y * wh.y) / cos2;
    float a = std::exp(-tan2 / (alpha * alpha));
    float b = kPi * alpha * alpha * cos2 * cos2;
    return a / b;
}

// ggxD, rayhit.rchit:185-192
inline float ggxD(vec3 wh, float alpha) {
    float cos2 = wh.z * wh.z;
    float tan2 = (wh.x * wh.x + wh.y * wh.y) / cos2;
    if (std::isinf(tan2)) return 0.0f;
    float b = (1.0f + tan2 / (alpha * alpha));
    float a = kPi * alpha * alpha * cos2 * cos2 * b * b;
    return 1.0f / a;
}

// ggxLambda, rayhit.rchit:194-200
inline float ggxLambda(vec3 wh, float alpha) {
    float cos2 = wh.z * wh.z;
    float tan2 = (wh.x * wh.x + wh.y * wh.y) / cos2;
    if (std::isinf(tan2)) return 0.0f;
    float a = -1.0f + std::sqrt(1.0f + alpha * alpha * tan2);
    return 0.5f * a;
}

// ggxMask, rayhit.rchit:202-204
inline float ggxMask(vec3 wo, vec3 wi, float alpha) { return 1.0f / (1.0f + ggxLambda(wo, alpha) + ggxLambda(wi, alpha)); }

// powerHeuristic, rayhit.rchit:206-210
inline float powerHeuristic(int nf, float fPdf, int ng, float gPdf) {
    float f = (float)nf * fPdf;
    float g = (float)ng * gPdf;
    return (f * f) / (f * f + g * g);
}

struct BSDFOutput { vec3 bsdf; float pdf; bool isDelta; };

// fresnel(no, cosTho, nt, cosTht), rayhit.rchit:218-226
inline float fresnel4(float no, float cosTho, float nt, float cosTht) {
    float a = nt * cosTho - no * cosTht;
    float ad = nt * cosTho + no * cosTht;
    float b = no * cosTho - nt * cosTht;
    float bd = no * cosTho + nt * cosTht;
    float A = a * a / (ad * ad);
    float B = b * b / (bd * bd);
    return 0.5f * (A + B);
}

// fresnel(vec3 wo, no, nt), rayhit.rchit:228-237
inline float fresnelW(vec3 wo, float no, float nt) {
    float sinTho = std::sqrt(gmax(wo.x * wo.x + wo.y * wo.y, 0.0f));
    float sqrtTerm = 1.0f - ((no * no) / (nt * nt)) * (sinTho * sinTho);
    if (sqrtTerm <= 0.0f) return 1.0f;
    float cosTht = std::sqrt(sqrtTerm);
    float cosTho = std::fabs(wo.z);
    return fresnel4(no, cosTho, nt, cosTht);
}

// fresnel(float cosTho, no, nt), rayhit.rchit:239-247
inline float fresnelC(float cosTho, float no, float nt) {
    float sinTho = std::sqrt(gmax(1.0f - cosTho * cosTho, 0.0f));
    float sqrtTerm = 1.0f - ((no * no) / (nt * nt)) * (sinTho * sinTho);
    if (sqrtTerm <= 0.0f) return 1.0f;
    float cosTht = std::sqrt(sqrtTerm);
    return fresnel4(no, cosTho, nt, cosTht);
}

// FresnelDieletricConductor, rayhit.rchit:269-288
inline vec3 FresnelDieletricConductor(vec3 Eta, vec3 Etak, float CosTheta) {
    float CosTheta2 = CosTheta * CosTheta;
    float SinTheta2 = 1.0f - CosTheta2;
    vec3 Eta2 = Eta * Eta;
    vec3 Etak2 = Etak * Etak;
    vec3 t0 = Eta2 - Etak2 - SinTheta2;
    vec3 a2plusb2 = vsqrt(t0 * t0 + 4.0f * Eta2 * Etak2);
    vec3 t1 = a2plusb2 + CosTheta2;
    vec3 a = vsqrt(0.5f * (a2plusb2 + t0));
    vec3 t2 = 2.0f * a * CosTheta;
    vec3 Rs = (t1 - t2) / (t1 + t2);
    vec3 t3 = CosTheta2 * a2plusb2 + SinTheta2 * SinTheta2;
    vec3 t4 = t2 * SinTheta2;
    vec3 Rp = Rs * (t3 - t4) / (t3 + t4);
    return 0.5f * (Rp + Rs);
}

// refractRay, rayhit.rchit:290-299
inline bool refractRay(vec3 wo, vec3 n, float no, float nt, vec3& wt) {
    float sinTho = std::sqrt(gmax(wo.x * wo.x + wo.y * wo.y, 0.0f));
    float sqrtTerm = 1.0f - ((no * no) / (nt * nt)) * (sinTho * sinTho);
    if (sqrtTerm <= 0.0f) return false;
    float cosTht = std::sqrt(sqrtTerm);
    wt = (no / nt) * (-wo) + ((no / nt) * dot(wo, n) - cosTht) * n;
    return true;
}

// coupledDiffuseTerm, rayhit.rchit:301-308
inline float coupledDiffuseTerm(float R0, float cosTho, float cosThi) {
    float k = 21.0f / (20.0f * kPi * (1.0f - R0));
    float a = 1.0f - cosTho;
    float b = 1.0f - cosThi;
    float a5 = a * a * a * a * a;
    float b5 = b * b * b * b * b;
    return k * (1.0f - a5) * (1.0f - b5);
}

// fresnelBlendDiffuseTerm, rayhit.rchit:310-317
inline float fresnelBlendDiffuseTerm(float R0, float cosTho, float cosThi) {
    float k = 28.0f / (23.0f * kPi);
    float a = 1.0f - 0.5f * cosTho;
    float b = 1.0f - 0.5f * cosThi;
    float a5 = a * a * a * a * a;
    float b5 = b * b * b * b * b;
    return k * (1.0f - R0) * (1.0f - a5) * (1.0f - b5);
}

// internalScatterEscapeFraction, rayhit.rchit:320-324
inline float internalScatterEscapeFraction(float R0, float no, float nt) {
    float Re = (kPi * 20.0f * R0 + 1.0f) / 21.0f;
    float eta = no / nt;
    return 1.0f - eta * eta * (1.0f - Re);
}

// schlickFresnel, rayhit.rchit:326-330
inline float schlickFresnel(float R0, float cosTho) {
    float a = 1.0f - cosTho;
    float a5 = a * a * a * a * a;
    return R0 + a5 * (1.0f - R0);
}

inline vec3 v3(const float* p) { return vec3(p[0], p[1], p[2]); }
inline vec3 mirror(vec3 wo) { return vec3(-wo.x, -wo.y, wo.z); }

// ---------------------------------------------------------------------------------------
// BSDF models — rayhit.rchit:341-617
// ---------------------------------------------------------------------------------------
// Diffuse, :341-358
void diffuseSample(const gsb_diffuse_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    (void)wo;
    vec3 kD = v3(b.reflectance);
    wi = randCosineHemisphere(rng);
    res.bsdf = kD / kPi;
    res.pdf = cosineHemispherePdf(wi);
    res.isDelta = false;
}
void diffuseEval(const gsb_diffuse_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    (void)wo;
    res.bsdf = v3(b.reflectance) / kPi;
    res.pdf = cosineHemispherePdf(wi);
    res.isDelta = false;
}

// SmoothDielectric, :362-404
void smoothDielectricSample(const gsb_smooth_dielectric_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    bool entering = wo.z > 0.0f;
    float no = entering ? b.iorOut : b.iorIn;
    float nt = entering ? b.iorIn : b.iorOut;
    float cosTho = wo.z;
    vec3 wt;
    vec3 up(0.0f, 0.0f, 1.0f);
    if (!refractRay(wo, faceforward(up, -wo, up), no, nt, wt)) {
        wi = mirror(wo);
        res.bsdf = 1.0f * vec3(1.0f / std::fabs(cosTho));
        res.isDelta = true;
        res.pdf = 1.0f;
        return;
    }
    float Fr = fresnel4(no, std::fabs(cosTho), nt, std::fabs(wt.z));
    float u = rng.randUniform();
    if (u < Fr) {
        wi = mirror(wo);
        res.bsdf = Fr * vec3(1.0f / std::fabs(cosTho));
        res.isDelta = true;
        res.pdf = Fr;
    } else {
        wi = wt;
        res.bsdf = vec3(((no * no) / (nt * nt)) * (1.0f - Fr) / std::fabs(wt.z));
        res.isDelta = true;
        res.pdf = 1.0f - Fr;
    }
}
void deltaEval(BSDFOutput& res) {  // smoothDielectricBSDFEval :400-404, smoothConductorBSDFEval :420-426
    res.bsdf = vec3(0.0f);
    res.pdf = 1.0f;
    res.isDelta = true;
}

// SmoothConductor, :406-418
void smoothConductorSample(const gsb_smooth_conductor_bsdf& b, vec3 wo, vec3& wi, BSDFOutput& res) {
    float no = b.iorOut;
    float nt = b.iorIn;
    float Fr = nt == 0.0f ? 1.0f : fresnelW(wo, no, nt);
    wi = mirror(wo);
    res.bsdf = Fr * vec3(1.0f / std::fabs(wo.z));
    res.isDelta = true;
    res.pdf = 1.0f;
}

// SmoothFloor, :428-458
void smoothFloorSample(const gsb_smooth_floor_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    float Fr = schlickFresnel(b.R0, std::fabs(wo.z));
    float u = rng.randUniform();
    if (u < Fr) {
        wi = mirror(wo);
        res.bsdf = v3(b.diffuse) * coupledDiffuseTerm(b.R0, std::fabs(wo.z), std::fabs(wi.z)) + Fr * vec3(1.0f / std::fabs(wo.z));
        res.pdf = Fr;
        res.isDelta = true;
    } else {
        wi = randCosineHemisphere(rng);
        res.bsdf = v3(b.diffuse) * coupledDiffuseTerm(b.R0, std::fabs(wo.z), std::fabs(wi.z));
        res.pdf = (1.0f - Fr) * cosineHemispherePdf(wi);
        res.isDelta = false;
    }
}
void smoothFloorEval(const gsb_smooth_floor_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    float Fr = schlickFresnel(b.R0, std::fabs(wo.z));
    res.bsdf = v3(b.diffuse) * coupledDiffuseTerm(b.R0, std::fabs(wo.z), std::fabs(wi.z));
    res.pdf = (1.0f - Fr) * cosineHemispherePdf(wi);
    res.isDelta = false;
}

// SmoothPlastic, :461-506
void smoothPlasticSample(const gsb_smooth_plastic_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    float u = rng.randUniform();
    float no = b.iorOut;
    float nt = b.iorIn;
    float Fri = fresnelC(std::fabs(wo.z), no, nt);
    if (u < Fri) {
        wi = mirror(wo);
        res.bsdf = Fri * vec3(1.0f / std::fabs(wo.z));
        res.pdf = Fri;
        res.isDelta = true;
    } else {
        wi = randCosineHemisphere(rng);
        float Fro = fresnelC(std::fabs(wi.z), no, nt);
        float Ri = internalScatterEscapeFraction(b.R0, no, nt);
        float eta = no / nt;
        vec3 diffuse = v3(b.diffuse);
        vec3 d = diffuse * eta * eta * (1.0f - Fri) * (1.0f - Fro) / (kPi * (1.0f - diffuse * Ri));
        res.bsdf = d;
        res.pdf = (1.0f - Fri) * cosineHemispherePdf(wi);
        res.isDelta = false;
    }
}
void smoothPlasticEval(const gsb_smooth_plastic_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    float no = b.iorOut;
    float nt = b.iorIn;
    float Fri = fresnelC(std::fabs(wo.z), no, nt);
    float Fro = fresnelC(std::fabs(wi.z), no, nt);
    float Ri = internalScatterEscapeFraction(b.R0, no, nt);
    float eta = no / nt;
    vec3 diffuse = v3(b.diffuse);
    vec3 d = diffuse * (1.0f - Fri) * (1.0f - Fro) * eta * eta / (kPi * (1.0f - diffuse * Ri));
    res.bsdf = d;
    res.pdf = (1.0f - Fri) * cosineHemispherePdf(wi);
    res.isDelta = false;
}

// RoughConductor, :508-530
void roughConductorSample(const gsb_rough_conductor_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    vec3 Fr = FresnelDieletricConductor(v3(b.eta), v3(b.k), std::fabs(wo.z));
    vec3 wh = sampleHalf(rng, b.alpha);
    if (wh.z <= 0.0f) wh = wh * -1.0f;
    wi = normalize(-wo + 2.0f * dot(wh, wo) * wh);
    res.bsdf = v3(b.reflectance) * Fr * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * std::fabs(wi.z) * std::fabs(wo.z));
    res.pdf = beckmannD(wh, b.alpha) * std::fabs(wh.z) / (4.0f * std::fabs(dot(wo, wh)));
    res.isDelta = false;
}
void roughConductorEval(const gsb_rough_conductor_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    vec3 Fr = FresnelDieletricConductor(v3(b.eta), v3(b.k), std::fabs(wo.z));
    vec3 wh = normalize(wo + wi);
    res.bsdf = Fr * v3(b.reflectance) * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * std::fabs(wi.z) * std::fabs(wo.z));
    res.pdf = beckmannD(wh, b.alpha) * std::fabs(wh.z) / (4.0f * std::fabs(dot(wo, wh)));
    res.isDelta = false;
}

// RoughPlastic, :532-582
void roughPlasticSample(const gsb_rough_plastic_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    float u = rng.randUniform();
    float no = b.iorOut;
    float nt = b.iorIn;
    float eta = no / nt;
    if (u < 0.5f) {
        vec3 wh = sampleHalf(rng, b.alpha);
        if (wh.z <= 0.0f) wh = wh * -1.0f;
        wi = normalize(-wo + 2.0f * dot(wh, wo) * wh);
    } else {
        wi = randCosineHemisphere(rng);
    }
    vec3 wh = normalize(wi + wo);
    float Fri = fresnelC(std::fabs(dot(wh, wo)), no, nt);
    float Fro = fresnelC(std::fabs(dot(wh, wi)), no, nt);
    float Ri = internalScatterEscapeFraction(b.R0, no, nt);
    vec3 kD = v3(b.diffuse);
    vec3 specular = vec3(Fri) * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * std::fabs(wo.z) * std::fabs(wi.z));
    vec3 d = kD * (1.0f - Fri) * (1.0f - Fro) * eta * eta / (kPi * (1.0f - kD * Ri));
    res.pdf = 0.5f * beckmannD(wh, b.alpha) * std::fabs(wh.z) / (4.0f * std::fabs(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}
void roughPlasticEval(const gsb_rough_plastic_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    float no = b.iorOut;
    float nt = b.iorIn;
    vec3 wh = normalize(wi + wo);
    float Fri = fresnelC(std::fabs(dot(wh, wo)), no, nt);
    float Fro = fresnelC(std::fabs(dot(wh, wi)), no, nt);
    float Ri = internalScatterEscapeFraction(b.R0, no, nt);
    float eta = no / nt;
    vec3 kD = v3(b.diffuse);
    vec3 specular = vec3(Fri) * ggxD(wh, b.alpha) * ggxMask(wo, wi, b.alpha) / (4.0f * std::fabs(wo.z) * std::fabs(wi.z));
    vec3 d = kD * (1.0f - Fri) * (1.0f - Fro) * eta * eta / (kPi * (1.0f - kD * Ri));
    res.pdf = 0.5f * gmax(beckmannD(wh, b.alpha) * std::fabs(wh.z), 0.01f) / (4.0f * std::fabs(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}

// RoughFloor, :583-617
void roughFloorSample(const gsb_rough_floor_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    float u = rng.randUniform();
    if (u < 0.5f) {
        vec3 wh = sampleHalf(rng, b.alpha);
        if (wh.z <= 0.0f) wh = wh * -1.0f;
        wi = normalize(-wo + 2.0f * dot(wh, wo) * wh);
    } else {
        wi = randCosineHemisphere(rng);
    }
    vec3 wh = normalize(wi + wo);
    float Fr = schlickFresnel(b.R0, std::fabs(dot(wo, wh)));
    vec3 d = v3(b.diffuse) * fresnelBlendDiffuseTerm(b.R0, std::fabs(wo.z), std::fabs(wi.z));
    vec3 specular = vec3(Fr) * ggxD(wh, b.alpha) / (4.0f * std::fabs(dot(wo, wh)) * gmax(std::fabs(wo.z), std::fabs(wi.z)));
    res.pdf = 0.5f * beckmannD(wh, b.alpha) * std::fabs(wh.z) / (4.0f * std::fabs(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}
void roughFloorEval(const gsb_rough_floor_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    vec3 wh = normalize(wi + wo);
    float Fr = schlickFresnel(b.R0, std::fabs(dot(wo, wh)));
    vec3 d = v3(b.diffuse) * fresnelBlendDiffuseTerm(b.R0, std::fabs(wo.z), std::fabs(wi.z));
    float ggx = ggxD(wh, b.alpha);
    vec3 specular = vec3(Fr) * ggx / (4.0f * std::fabs(dot(wo, wh)) * gmax(std::fabs(wo.z), std::fabs(wi.z)));
    res.pdf = 0.5f * beckmannD(wh, b.alpha) * std::fabs(wh.z) / (4.0f * std::fabs(dot(wo, wh))) + 0.5f * cosineHemispherePdf(wi);
    res.bsdf = d + specular;
    res.isDelta = false;
}

// RoughDielectric — extension with no counterpart in rayhit.rchit (the reference only carries a dead include,
// engine/Loader.cpp:88-143).  Restates Walter, Marschner, Li & Torrance, "Microfacet Models for Refraction through
// Rough Surfaces" (EGSR 2007): GGX D (eq. 33), the reference shader's ggxMask as G, reflection term eq. 20,
// transmission term eq. 21 with the half-vector Jacobian of eq. 17, written for a path that walks against the light
// (wo back along the path, light arriving along wi), hence the (no/nt)^2 factor of the smooth dielectric (:362-404).
// Draw order: two uniforms for the half vector (eq. 35-36), then one for the lobe.  Parity for this type is
// unpinned by construction: this restatement is the only oracle.
vec3 sampleHalfGGX(Rng& rng, float alpha) {
    float ux = rng.randUniform();
    float uy = rng.randUniform();
    float phi = 2.0f * kPi * ux;
    float tan2 = alpha * alpha * uy / (1.0f - uy);
    float cost = 1.0f / std::sqrt(1.0f + tan2);
    float sint = std::sqrt(gmax(0.0f, 1.0f - cost * cost));
    return vec3(std::cos(phi) * sint, std::sin(phi) * sint, cost);
}
void roughDielectricInvalid(vec3 wo, vec3& wi, BSDFOutput& res) {
    wi = vec3(-wo.x, -wo.y, wo.z);
    res.bsdf = vec3(0.0f);
    res.pdf = 0.0f;
    res.isDelta = false;
}
void roughDielectricSample(const gsb_rough_dielectric_bsdf& b, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    const bool entering = wo.z > 0.0f;
    const float no = entering ? b.iorOut : b.iorIn;
    const float nt = entering ? b.iorIn : b.iorOut;
    vec3 wh = sampleHalfGGX(rng, b.alpha);
    if (!entering) wh = wh * -1.0f;
    const float u = rng.randUniform();
    const float cosI = dot(wo, wh);
    const float D = ggxD(wh, b.alpha);
    if (!(cosI > 0.0f) || !(D > 0.0f)) { roughDielectricInvalid(wo, wi, res); return; }
    const float Fr = fresnelC(cosI, no, nt);
    res.isDelta = false;
    if (u < Fr) {
        wi = normalize(-wo + 2.0f * cosI * wh);
        if (!(wi.z * wo.z > 0.0f)) { roughDielectricInvalid(wo, wi, res); return; }
        res.bsdf = vec3(Fr * D * ggxMask(wo, wi, b.alpha) / (4.0f * std::fabs(wi.z) * std::fabs(wo.z)));
        res.pdf = Fr * D * std::fabs(wh.z) / (4.0f * cosI);
    } else {
        const float eta = no / nt;
        const float cosT2 = 1.0f - eta * eta * (1.0f - cosI * cosI);
        wi = normalize(-eta * wo + (eta * cosI - std::sqrt(gmax(cosT2, 0.0f))) * wh);
        const float cosO = dot(wi, wh);
        if (!(wi.z * wo.z < 0.0f) || !(cosO < 0.0f)) { roughDielectricInvalid(wo, wi, res); return; }
        const float denom = no * cosI + nt * cosO;
        const float jac = std::fabs(cosO) / (denom * denom);
        res.bsdf = vec3((1.0f - Fr) * D * ggxMask(wo, wi, b.alpha) * cosI * no * no * jac / (std::fabs(wi.z) * std::fabs(wo.z)));
        res.pdf = (1.0f - Fr) * D * std::fabs(wh.z) * nt * nt * jac;
    }
}
void roughDielectricEval(const gsb_rough_dielectric_bsdf& b, vec3 wo, vec3 wi, BSDFOutput& res) {
    const bool entering = wo.z > 0.0f;
    const float no = entering ? b.iorOut : b.iorIn;
    const float nt = entering ? b.iorIn : b.iorOut;
    res.isDelta = false;
    res.bsdf = vec3(0.0f);
    res.pdf = 0.0f;
    const bool reflect = wi.z * wo.z > 0.0f;
    vec3 wh = reflect ? normalize(wo + wi) : normalize(no * wo + nt * wi);
    if (dot(wo, wh) < 0.0f) wh = wh * -1.0f;
    const float cosI = dot(wo, wh), cosO = dot(wi, wh);
    const float D = ggxD(wh, b.alpha);
    if (!(cosI > 0.0f) || !(D > 0.0f) || wi.z == 0.0f || wo.z == 0.0f) return;
    if (entering ? wh.z <= 0.0f : wh.z >= 0.0f) return;
    const float Fr = fresnelC(cosI, no, nt);
    if (reflect) {
        res.bsdf = vec3(Fr * D * ggxMask(wo, wi, b.alpha) / (4.0f * std::fabs(wi.z) * std::fabs(wo.z)));
        res.pdf = Fr * D * std::fabs(wh.z) / (4.0f * cosI);
    } else {
        if (!(cosO < 0.0f)) return;
        const float denom = no * cosI + nt * cosO;
        if (denom == 0.0f) return;
        const float jac = std::fabs(cosO) / (denom * denom);
        res.bsdf = vec3((1.0f - Fr) * D * ggxMask(wo, wi, b.alpha) * cosI * no * no * jac / (std::fabs(wi.z) * std::fabs(wo.z)));
        res.pdf = (1.0f - Fr) * D * std::fabs(wh.z) * nt * nt * jac;
    }
}

// isTransimissionBSDF, :620-627 (+ the rough dielectric extension)
inline bool isTransmission(uint32_t type) { return type == GSB_BSDF_SMOOTH_DIELECTRIC || type == GSB_BSDF_ROUGH_DIELECTRIC; }

// sampleBSDF dispatch, :630-641
void sampleBSDF(const or_scene& s, uint32_t handle, Rng& rng, vec3 wo, vec3& wi, BSDFOutput& res) {
    const uint32_t idx = handle & 0xffffu;
    switch (handle >> 16) {
        case GSB_BSDF_DIFFUSE: diffuseSample(s.diffuse[idx], rng, wo, wi, res); break;
        case GSB_BSDF_SMOOTH_DIELECTRIC: smoothDielectricSample(s.smoothDielectric[idx], rng, wo, wi, res); break;
        case GSB_BSDF_SMOOTH_CONDUCTOR: smoothConductorSample(s.smoothConductor[idx], wo, wi, res); break;
        case GSB_BSDF_SMOOTH_PLASTIC: smoothPlasticSample(s.smoothPlastic[idx], rng, wo, wi, res); break;
        case GSB_BSDF_ROUGH_CONDUCTOR: roughConductorSample(s.roughConductor[idx], rng, wo, wi, res); break;
        case GSB_BSDF_SMOOTH_FLOOR: smoothFloorSample(s.smoothFloor[idx], rng, wo, wi, res); break;
        case GSB_BSDF_ROUGH_FLOOR: roughFloorSample(s.roughFloor[idx], rng, wo, wi, res); break;
        case GSB_BSDF_ROUGH_PLASTIC: roughPlasticSample(s.roughPlastic[idx], rng, wo, wi, res); break;
        case GSB_BSDF_ROUGH_DIELECTRIC: roughDielectricSample(s.roughDielectric[idx], rng, wo, wi, res); break;
    }
}

// evalBSDF dispatch, :643-654
void evalBSDF(const or_scene& s, uint32_t handle, vec3 wo, vec3 wi, BSDFOutput& res) {
    const uint32_t idx = handle & 0xffffu;
    switch (handle >> 16) {
        case GSB_BSDF_DIFFUSE: diffuseEval(s.diffuse[idx], wo, wi, res); break;
        case GSB_BSDF_SMOOTH_DIELECTRIC: deltaEval(res); break;
        case GSB_BSDF_SMOOTH_CONDUCTOR: deltaEval(res); break;
        case GSB_BSDF_SMOOTH_PLASTIC: smoothPlasticEval(s.smoothPlastic[idx], wo, wi, res); break;
        case GSB_BSDF_ROUGH_CONDUCTOR: roughConductorEval(s.roughConductor[idx], wo, wi, res); break;
        case GSB_BSDF_SMOOTH_FLOOR: smoothFloorEval(s.smoothFloor[idx], wo, wi, res); break;
        case GSB_BSDF_ROUGH_FLOOR: roughFloorEval(s.roughFloor[idx], wo, wi, res); break;
        case GSB_BSDF_ROUGH_PLASTIC: roughPlasticEval(s.roughPlastic[idx], wo, wi, res); break;
        case GSB_BSDF_ROUGH_DIELECTRIC: roughDielectricEval(s.roughDielectric[idx], wo, wi, res); break;
    }
}

size_t bsdfCount(const or_scene& s, uint32_t type) {
    switch (type) {
        case GSB_BSDF_DIFFUSE: return s.diffuse.size();
        case GSB_BSDF_SMOOTH_DIELECTRIC: return s.smoothDielectric.size();
        case GSB_BSDF_SMOOTH_CONDUCTOR: return s.smoothConductor.size();
        case GSB_BSDF_SMOOTH_PLASTIC: return s.smoothPlastic.size();
        case GSB_BSDF_ROUGH_CONDUCTOR: return s.roughConductor.size();
        case GSB_BSDF_SMOOTH_FLOOR: return s.smoothFloor.size();
        case GSB_BSDF_ROUGH_FLOOR: return s.roughFloor.size();
        case GSB_BSDF_ROUGH_PLASTIC: return s.roughPlastic.size();
        case GSB_BSDF_ROUGH_DIELECTRIC: return s.roughDielectric.size();
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// Light sampling — rayhit.rchit:117-153
// ---------------------------------------------------------------------------------------
struct LightOutput { vec3 position; vec3 emission; float pdf; };

// sampleTrangleLight, :123-145
LightOutput sampleTriangleLight(const gsb_triangle_light& light, Rng& rng, vec3 pos) {
    float e1 = rng.randUniform();
    float e2 = rng.randUniform();
    float u = 1.0f - std::sqrt(e1);
    float v = e2 * std::sqrt(e1);
    float w = 1.0f - u - v;
    vec3 v0 = v3(light.positions[0]);
    vec3 v1 = v3(light.positions[1]);
    vec3 v2 = v3(light.positions[2]);
    float A = 0.5f * std::fabs(length(cross(v2 - v0, v1 - v0)));
    vec3 normal = normalize(cross(v1 - v0, v2 - v0));
    vec3 lightPos = u * v0 + v * v1 + w * v2;
    float ldist = length(lightPos - pos);
    vec3 l = normalize(lightPos - pos);
    LightOutput res;
    res.position = lightPos;
    res.emission = v3(light.radiance) * (dot(-l, normal) > 0.0f ? 1.0f : 0.0f);
    res.pdf = ldist * ldist / (std::fabs(dot(-l, normal)) * A);
    return res;
}

// ---- environment emitter (extension, no reference source: engine/Loader.cpp:338-346 holds it as commented-out code) ----
inline vec3 envRotate(const float* R, vec3 d) {
    return vec3(R[0] * d.x + R[1] * d.y + R[2] * d.z, R[3] * d.x + R[4] * d.y + R[5] * d.z, R[6] * d.x + R[7] * d.y + R[8] * d.z);
}
inline vec3 envRotateInv(const float* R, vec3 d) {
    return vec3(R[0] * d.x + R[3] * d.y + R[6] * d.z, R[1] * d.x + R[4] * d.y + R[7] * d.z, R[2] * d.x + R[5] * d.y + R[8] * d.z);
}
vec3 envRadiance(const or_scene::Environment& e, vec3 d) {
    const vec3 sc = v3(e.desc.scale);
    if (e.desc.type == GSB_ENVIRONMENT_CONSTANT) return sc;
    const vec3 l = envRotateInv(e.desc.to_world, d);
    const float th = std::acos(std::fmin(std::fmax(l.y, -1.0f), 1.0f));
    const float ph = std::atan2(l.x, -l.z);
    const float u = ph * (0.5f / kPi) + 0.5f, v = th * (1.0f / kPi);
    const int W = (int)e.desc.width, H = (int)e.desc.height;
    const int x = std::max(std::min((int)(u * (float)W), W - 1), 0), y = std::max(std::min((int)(v * (float)H), H - 1), 0);
    return v3(&e.texels[3 * ((size_t)y * W + x)]) * sc;
}
inline int envSearch(const float* cdf, int n, float x) {   // largest i in [0, n) with cdf[i] <= x
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (cdf[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}
vec3 envSample(const or_scene::Environment& e, float e1, float e2, vec3& radiance, float& pdf) {
    if (e.desc.type == GSB_ENVIRONMENT_CONSTANT) {
        const float z = 1.0f - 2.0f * e1, r = std::sqrt(std::fmax(0.0f, 1.0f - z * z)), ph = 2.0f * kPi * e2;
        radiance = v3(e.desc.scale);
        pdf = 1.0f / (4.0f * kPi);
        return vec3(r * std::cos(ph), z, r * std::sin(ph));
    }
    const int W = (int)e.desc.width, H = (int)e.desc.height;
    const float a = std::fmin(e1, 0.99999994f), b = std::fmin(e2, 0.99999994f);
    const int y = envSearch(e.rowCdf.data(), H, a);
    const float r0 = e.rowCdf[y], r1 = e.rowCdf[y + 1];
    const float* cc = &e.colCdf[(size_t)y * (W + 1)];
    const int x = envSearch(cc, W, b);
    const float c0 = cc[x], c1 = cc[x + 1];
    const float fv = r1 > r0 ? (a - r0) / (r1 - r0) : 0.5f, fu = c1 > c0 ? (b - c0) / (c1 - c0) : 0.5f;
    const float u = ((float)x + fu) / (float)W, v = ((float)y + fv) / (float)H;
    const float th = v * kPi, ph = (u - 0.5f) * (2.0f * kPi);
    const float st = std::sin(th);
    const vec3 l(st * std::sin(ph), std::cos(th), -st * std::cos(ph));
    radiance = v3(&e.texels[3 * ((size_t)y * W + x)]) * v3(e.desc.scale);
    pdf = st > 0.0f ? (r1 - r0) * (c1 - c0) * (float)W * (float)H / (2.0f * kPi * kPi * st) : 0.0f;
    return envRotate(e.desc.to_world, l);
}

// sampleLight, :147-153.  numLights == 0 is a division by zero in the shader; the oracle
// defines it as: the three draws are still consumed, pdf = 0, no emission.
LightOutput sampleLight(const or_scene& s, Rng& rng, vec3 pos, bool& haveLight) {
    // the environment emitter (extension) counts as one more light behind the triangle lights
    const uint32_t numLights = (uint32_t)s.lights.size() + (s.env.desc.type ? 1u : 0u);
    if (numLights == 0) {
        rng.randPcg(); rng.randPcg(); rng.randPcg();
        haveLight = false;
        LightOutput res;
        res.position = pos + vec3(0.0f, 1.0f, 0.0f);
        res.emission = vec3(0.0f);
        res.pdf = 0.0f;
        return res;
    }
    haveLight = true;
    uint32_t lightIdx = rng.randPcg() % numLights;
    if (lightIdx == s.lights.size()) {
        LightOutput res;
        const float e1 = rng.randUniform();
        const float e2 = rng.randUniform();
        const vec3 d = envSample(s.env, e1, e2, res.emission, res.pdf);
        res.position = pos + d * s.env.distance;
        res.pdf *= 1.0f / (float)numLights;
        return res;
    }
    LightOutput res = sampleTriangleLight(s.lights[lightIdx], rng, pos);
    res.pdf *= 1.0f / (float)numLights;
    return res;
}

// ---------------------------------------------------------------------------------------
// Path state + closest-hit shading — pt_common.glsl:4-14, rayhit.rchit:666-797
// ---------------------------------------------------------------------------------------
struct HitPayload {
    vec3 emitted, weight, origin, direction;
    float directWeight;
    uint32_t seed;
    int wasDelta, countEmitted, done;
};

inline bool isvalid(float x) { return !std::isnan(x) && !std::isinf(x); }
inline bool isvalid(vec3 v) { return isvalid(v.x) && isvalid(v.y) && isvalid(v.z); }

struct PathStats { uint64_t ext = 0, shadow = 0, draws = 0; };

// Reflectance textures (extension, no reference source: engine/Loader.cpp:122-143 holds them as commented-out code).
// checkerboard: Mitsuba's rule on the transformed coordinate; bitmap: bilinear over texel centres, repeat, v = 1 at row 0.
vec3 textureEval(const or_scene::Texture& t, float u, float v) {
    const gsb_texture& d = t.desc;
    const float uu = u * d.uscale + d.uoffset, vv = v * d.vscale + d.voffset;
    if (d.type == GSB_TEXTURE_CHECKERBOARD) {
        const int ix = (int)std::floor(2.0f * uu), iy = (int)std::floor(2.0f * vv);
        return ((ix + iy) & 1) == 0 ? v3(d.color0) : v3(d.color1);
    }
    const float w = (float)d.width, h = (float)d.height;
    const float x = (uu - std::floor(uu)) * w - 0.5f, y = (1.0f - (vv - std::floor(vv))) * h - 0.5f;
    const float fx = std::floor(x), fy = std::floor(y);
    const float ax = x - fx, ay = y - fy;
    const int W = (int)d.width, H = (int)d.height;
    int x0 = (int)fx, y0 = (int)fy;
    x0 = ((x0 % W) + W) % W; y0 = ((y0 % H) + H) % H;
    const int x1 = (x0 + 1) % W, y1 = (y0 + 1) % H;
    auto px = [&](int xx, int yy) { const float* p = &t.texels[3 * ((size_t)yy * W + xx)]; return vec3(p[0], p[1], p[2]); };
    return (px(x0, y0) * (1.0f - ax) + px(x1, y0) * ax) * (1.0f - ay) + (px(x0, y1) * (1.0f - ax) + px(x1, y1) * ax) * ay;
}

void closestHit(const or_scene& s, const Hit& hit, HitPayload& prd, PathStats& st, or_bounce* rec) {
    Rng rng(prd.seed);  // :668
    const Instance& instance = s.instances[hit.inst];
    const Mesh& mesh = s.meshes[instance.src.mesh];
    const size_t base = 9 * (size_t)hit.prim;  // :670, vertices 3p, 3p+1, 3p+2

    // :676-688 positions through gl_ObjectToWorldEXT, normals through transformInvT
    vec3 pos[3], nrm[3];
    for (int k = 0; k < 3; k++) {
        const float* p = &mesh.pos[base + 3 * k];
        const float* n = &mesh.nrm[base + 3 * k];
        const float* M = instance.o2w;
        pos[k] = vec3(M[0] * p[0] + M[1] * p[1] + M[2] * p[2] + M[3],
                      M[4] * p[0] + M[5] * p[1] + M[6] * p[2] + M[7],
                      M[8] * p[0] + M[9] * p[1] + M[10] * p[2] + M[11]);
        const float* N = instance.nrmM;
        nrm[k] = vec3(N[0] * n[0] + N[1] * n[1] + N[2] * n[2],
                      N[3] * n[0] + N[4] * n[1] + N[5] * n[2],
                      N[6] * n[0] + N[7] * n[1] + N[8] * n[2]);
    }
    const vec3 bary(1.0f - hit.u - hit.v, hit.u, hit.v);               // :690
    vec3 position = prd.origin + prd.direction * hit.t;                // :692
    vec3 SN = normalize(bary.x * nrm[0] + bary.y * nrm[1] + bary.z * nrm[2]);  // :693
    vec3 N = normalize(cross(pos[1] - pos[0], pos[2] - pos[0]));       // :694

    const vec3 rayDir = prd.direction;                                 // :696
    const vec3 emission = v3(instance.src.emission);
    if (dot(N, -rayDir) < 0.0f) {                                      // :698-707
        if (instance.src.twofaced == 1 && emission.x == 0.0f && emission.y == 0.0f && emission.z == 0.0f) {
            N = N * -1.0f;
            SN = SN * -1.0f;
        }
    }
    Onb onb = onbCreate(SN);                                           // :712
    vec3 wo = normalize(onbTransform(onb, -rayDir));                   // :713
    BSDFOutput bsdfRes;
    bsdfRes.bsdf = vec3(0.0f); bsdfRes.pdf = 0.0f; bsdfRes.isDelta = false;
    vec3 wi;
    const uint32_t handle = instance.src.bsdf;
    // textured reflectance (extension; the diffuse and the rough-plastic records carry a hasTexture field the reference never
    // uses): the colour looked up at the hit's interpolated texture coordinate stands in for the record's constant one
    const uint32_t btype = handle >> 16, bidx = handle & 0xffffu;
    bool textured = false;
    gsb_diffuse_bsdf texDiffuse = {};
    gsb_rough_plastic_bsdf texPlastic = {};
    if (btype == GSB_BSDF_DIFFUSE || btype == GSB_BSDF_ROUGH_PLASTIC) {
        const int tex = btype == GSB_BSDF_DIFFUSE ? s.diffuse[bidx].hasTexture : s.roughPlastic[bidx].hasTexture;
        if (tex > 0 && (size_t)tex <= s.textures.size()) {
            float tu = 0.0f, tv = 0.0f;
            for (int k = 0; k < 3; k++) {
                const float bk = k == 0 ? bary.x : (k == 1 ? bary.y : bary.z);
                const size_t vi = 3 * (size_t)hit.prim + k;
                tu += bk * (mesh.uv.empty() ? 0.0f : mesh.uv[2 * vi]);
                tv += bk * (mesh.uv.empty() ? 0.0f : mesh.uv[2 * vi + 1]);
            }
            const vec3 kd = textureEval(s.textures[tex - 1], tu, tv);
            textured = true;
            if (btype == GSB_BSDF_DIFFUSE) { texDiffuse = s.diffuse[bidx]; texDiffuse.reflectance[0] = kd.x; texDiffuse.reflectance[1] = kd.y; texDiffuse.reflectance[2] = kd.z; }
            else { texPlastic = s.roughPlastic[bidx]; texPlastic.diffuse[0] = kd.x; texPlastic.diffuse[1] = kd.y; texPlastic.diffuse[2] = kd.z; }
        }
    }
    if (!textured) sampleBSDF(s, handle, rng, wo, wi, bsdfRes);        // :716
    else if (btype == GSB_BSDF_DIFFUSE) diffuseSample(texDiffuse, rng, wo, wi, bsdfRes);
    else roughPlasticSample(texPlastic, rng, wo, wi, bsdfRes);
    float NoW = std::fabs(wi.z);                                       // :717
    wi = onbUntransform(onb, wi);                                      // :718

    bool haveLight = false;
    LightOutput lightRes = sampleLight(s, rng, position, haveLight);   // :720
    vec3 L = normalize(lightRes.position - position);                  // :722
    vec3 wL = onbTransform(onb, L);                                    // :723
    float Ldist = length(lightRes.position - position);                // :724
    const float NoL = std::fabs(dot(SN, L));                           // :725
    float lightPdf = lightRes.pdf;                                     // :726

    BSDFOutput lightBsdfRes;
    lightBsdfRes.bsdf = vec3(0.0f); lightBsdfRes.pdf = 0.0f; lightBsdfRes.isDelta = false;
    if (!textured) evalBSDF(s, handle, wo, wL, lightBsdfRes);          // :729
    else if (btype == GSB_BSDF_DIFFUSE) diffuseEval(texDiffuse, wo, wL, lightBsdfRes);
    else roughPlasticEval(texPlastic, wo, wL, lightBsdfRes);
    const bool transmissive = isTransmission(handle >> 16);
    bool neeDone = false;                                              // :732
    if (s.params.nee && haveLight) {                                   // :734
        if (!bsdfRes.isDelta) {                                        // :735
            if ((dot(N, -rayDir) > 0.0f && dot(N, L) > 0.0f) || transmissive) {  // :736
                // :737-748 shadow ray, any-hit, tmin 0.01, tmax Ldist - 0.01
                const float o[3] = {position.x, position.y, position.z};
                const float d[3] = {L.x, L.y, L.z};
                Hit sh = traceF32(s, o, d, 0.01f, Ldist - 0.01f, true, true);
                st.shadow++;
                bool shadowed = sh.inst != GSB_NO_HIT;
                if (rec) { rec->shadow_cast = 1; rec->shadow_visible = shadowed ? 0 : 1; }
                if (!shadowed && lightPdf != 0.0f) {                   // :750
                    float w = powerHeuristic(1, lightPdf, 1, bsdfRes.pdf);  // :751
                    prd.emitted = prd.emitted + w * NoL * lightBsdfRes.bsdf * prd.weight * lightRes.emission / lightPdf;  // :752
                    neeDone = true;
                }
            }
        }
    }

    prd.seed = rng.state;                                              // :759
    st.draws += rng.draws;
    float lightFlag = dot(N, -rayDir) > 0.0f ? 1.0f : 0.0f;            // :760
    const bool NEE = s.params.nee != 0;
    if (NEE && prd.countEmitted == 0 && prd.wasDelta == 0)             // :763-765
        prd.emitted = prd.emitted + prd.directWeight * emission * lightFlag * prd.weight;
    if (!NEE || prd.countEmitted == 1 || prd.wasDelta == 1)            // :766-768
        prd.emitted = prd.emitted + emission * lightFlag * prd.weight;
    if (dot(wi, N) <= 0.0f && !transmissive) { prd.done = 1; return; }        // :770-773
    if (dot(N, -rayDir) <= 0.0f && !transmissive) { prd.done = 1; return; }   // :776-779
    if (!isvalid(bsdfRes.pdf) || !isvalid(bsdfRes.bsdf) || bsdfRes.pdf == 0.0f) { prd.done = 1; return; }  // :781-784
    if (neeDone) prd.directWeight = powerHeuristic(1, bsdfRes.pdf, 1, lightPdf);  // :785-790
    else prd.directWeight = 1.0f;
    prd.countEmitted = 0;                                              // :792
    prd.origin = position + 0.0001f * faceforward(N, -wi, N);          // :793
    prd.direction = wi;                                                // :794
    prd.weight = prd.weight * (bsdfRes.bsdf * NoW / bsdfRes.pdf);      // :795
    prd.wasDelta = bsdfRes.isDelta ? 1 : 0;                            // :796
}

// rayDir + camera transform — raygen.rgen:20-35
// gsb_params::pixel_filter (extension, no reference source): offset from the pixel's integer coordinate for one uniform draw
inline float filterOffset(int filter, float u) {
    if (filter == 1) return u - 0.5f;
    return u < 0.5f ? std::sqrt(2.0f * u) - 1.0f : 1.0f - std::sqrt(2.0f - 2.0f * u);
}

void primaryRay(const or_scene& s, uint32_t W, uint32_t H, uint32_t x, uint32_t y, vec3& origin, vec3& direction, float offx = 0.0f,
                float offy = 0.0f) {
    const float sx = (float)W, sy = (float)H;
    const float fx = ((float)x + offx) - sx / 2.0f;
    const float fy = ((float)y + offy) - sy / 2.0f;
    const float z = (gmax(sx, sy) / 2.0f) / std::tan(s.fov / 2.0f);
    vec3 dir = normalize(vec3(-fx, fy, z));
    const float* V = s.view;  // column-major
    direction = vec3(V[0] * dir.x + V[4] * dir.y + V[8] * dir.z,
                     V[1] * dir.x + V[5] * dir.y + V[9] * dir.z,
                     V[2] * dir.x + V[6] * dir.y + V[10] * dir.z);
    direction.y *= -1.0f;
    origin = vec3(s.eye[0], s.eye[1], s.eye[2]);
}

// raygen main — raygen.rgen:29-81: one pixel-sample
vec3 pathSample(const or_scene& s, uint32_t W, uint32_t H, uint32_t x, uint32_t y, uint32_t timestamp, PathStats& st,
                or_bounce* bounces, uint32_t maxBounces, uint32_t* nBounces) {
    vec3 origin, direction;
    uint32_t rngState = pcgHash(tea(W * y + x, timestamp));           // :37
    float offx = 0.0f, offy = 0.0f;
    if (s.params.pixel_filter) {
        // extension: the first two draws of the stream place the ray inside the pixel's footprint (:38 is only a comment)
        Rng jr(rngState);
        const float u0 = jr.randUniform(), u1 = jr.randUniform();
        offx = filterOffset(s.params.pixel_filter, u0);
        offy = filterOffset(s.params.pixel_filter, u1);
        rngState = jr.state;
        st.draws += jr.draws;
    }
    primaryRay(s, W, H, x, y, origin, direction, offx, offy);
    vec3 result(0.0f);
    HitPayload prd;
    prd.weight = vec3(1.0f);
    prd.directWeight = 1.0f;
    prd.countEmitted = 1;
    prd.wasDelta = 0;
    prd.done = 0;
    prd.seed = rngState;
    prd.origin = origin;
    prd.direction = direction;
    int depth = 0;
    uint32_t nb = 0;
    while (true) {
        prd.emitted = vec3(0.0f);
        const float o[3] = {prd.origin.x, prd.origin.y, prd.origin.z};
        const float d[3] = {prd.direction.x, prd.direction.y, prd.direction.z};
        Hit hit = traceF32(s, o, d, 0.0f, 1e10f, false, true);         // :54-58
        st.ext++;
        or_bounce* rec = (bounces && nb < maxBounces) ? &bounces[nb] : nullptr;
        if (rec) { std::memset(rec, 0, sizeof(*rec)); rec->inst = hit.inst; rec->prim = hit.prim; rec->t = hit.t; }
        if (hit.inst == GSB_NO_HIT) {
            prd.done = 1;                                              // miss.rmiss:15-18
            if (s.env.desc.type) {
                // environment emitter (extension): weighted like a surface emitter a BSDF sample reaches, :760-768
                const bool plain = !s.params.nee || prd.countEmitted != 0 || prd.wasDelta != 0;
                prd.emitted = envRadiance(s.env, prd.direction) * prd.weight * (plain ? 1.0f : prd.directWeight);
            }
        }
        else closestHit(s, hit, prd, st, rec);
        if (rec) {
            rec->emitted[0] = prd.emitted.x; rec->emitted[1] = prd.emitted.y; rec->emitted[2] = prd.emitted.z;
            rec->weight[0] = prd.weight.x; rec->weight[1] = prd.weight.y; rec->weight[2] = prd.weight.z;
            rec->seed = prd.seed;
        }
        nb++;
        Rng rr(prd.seed);                                              // :59 (this copy is never written back to prd.seed)
        const float cutoff = s.params.cutoff;
        if (prd.emitted.x < cutoff && prd.emitted.y < cutoff && prd.emitted.z < cutoff) result = result + prd.emitted;  // :60-63
        if (depth > s.params.rr_start) {                               // :66-71
            float q = clampf(gmax(gmax(prd.weight.x, prd.weight.y), prd.weight.z), 0.05f, 1.0f);
            if (rr.randUniform() > q) break;
            prd.weight = prd.weight / q;
        }
        if (depth > s.params.max_depth) break;                         // :73-75
        if (prd.done == 1) break;                                      // :77-78
        ++depth;
    }
    if (nBounces) *nBounces = nb;
    return result;
}

void deriveInstance(Instance& in) {
    const float* T = in.src.transform;  // column-major: T[c*4+r]
    double M[3][4];
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 4; c++) {
            M[r][c] = T[c * 4 + r];
            in.o2w[r * 4 + c] = T[c * 4 + r];
        }
    // inverse of the affine part in float64, rounded once to float32
    const double a = M[0][0], b = M[0][1], c = M[0][2], d = M[1][0], e = M[1][1], f = M[1][2], g = M[2][0], h = M[2][1], i = M[2][2];
    const double det = a * (e * i - f * h) - b * (d * i - f * g) + c * (d * h - e * g);
    const double id = 1.0 / det;
    double inv[3][3] = {{(e * i - f * h) * id, (c * h - b * i) * id, (b * f - c * e) * id},
                        {(f * g - d * i) * id, (a * i - c * g) * id, (c * d - a * f) * id},
                        {(d * h - e * g) * id, (b * g - a * h) * id, (a * e - b * d) * id}};
    for (int r = 0; r < 3; r++) {
        for (int k = 0; k < 3; k++) in.w2o[r * 4 + k] = (float)inv[r][k];
        in.w2o[r * 4 + 3] = (float)(-(inv[r][0] * M[0][3] + inv[r][1] * M[1][3] + inv[r][2] * M[2][3]));
    }
    // transformInvT = glm::inverse(glm::transpose(M)) in float32, PathTracer.cpp:62.  glm's 4x4 inverse is the cofactor
    // expansion of external/glm/glm/detail/func_matrix.inl:294-352, restated here operation for operation so that the
    // normal matrix carries the same bits as the reference's host code computes; rows 0..2 of its 3x3 part are kept.
    {
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
        float inv4[4][4];   // inv4[c][r]
        for (int k = 0; k < 4; k++) {
            inv4[0][k] = ((v1[k] * f0[k] - v2[k] * f1[k]) + v3[k] * f2[k]) * sa[k];
            inv4[1][k] = ((v0[k] * f0[k] - v2[k] * f3[k]) + v3[k] * f4[k]) * sb[k];
            inv4[2][k] = ((v0[k] * f1[k] - v1[k] * f3[k]) + v3[k] * f5[k]) * sa[k];
            inv4[3][k] = ((v0[k] * f2[k] - v1[k] * f4[k]) + v2[k] * f5[k]) * sb[k];
        }
        const float d0 = m[0][0] * inv4[0][0], d1 = m[0][1] * inv4[1][0], d2 = m[0][2] * inv4[2][0], d3 = m[0][3] * inv4[3][0];
        const float oneOverDet = 1.0f / ((d0 + d1) + (d2 + d3));
        // normal' = transformInvT * vec4(n, 0): row r of the 3x3 part is (inv4[0][r], inv4[1][r], inv4[2][r])
        for (int r = 0; r < 3; r++)
            for (int k = 0; k < 3; k++) in.nrmM[r * 3 + k] = inv4[k][r] * oneOverDet;
    }
}

void instanceBounds(const or_scene& s, Instance& in) {
    const Mesh& m = s.meshes[in.src.mesh];
    double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int corner = 0; corner < 8; corner++) {
        double p[3] = {(corner & 1) ? m.hi[0] : m.lo[0], (corner & 2) ? m.hi[1] : m.lo[1], (corner & 4) ? m.hi[2] : m.lo[2]};
        for (int r = 0; r < 3; r++) {
            const float* row = in.o2w + 4 * r;
            double w = (double)row[0] * p[0] + (double)row[1] * p[1] + (double)row[2] * p[2] + (double)row[3];
            lo[r] = std::min(lo[r], w);
            hi[r] = std::max(hi[r], w);
        }
    }
    double scale = 0.0;
    for (int a = 0; a < 3; a++) scale = std::max(scale, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
    const double pad = scale * (double)kPadRel;
    for (int a = 0; a < 3; a++) { in.wlo[a] = lo[a] - pad; in.whi[a] = hi[a] + pad; }
}

inline float srgbEncode(float c) {
    if (!(c > 0.0f)) return 0.0f;
    if (c >= 1.0f) return 1.0f;
    return c <= 0.0031308f ? 12.92f * c : 1.055f * std::pow(c, 1.0f / 2.4f) - 0.055f;
}

}  // namespace

// =========================================================================================
// C interface
// =========================================================================================
extern "C" {

or_scene* or_scene_create(void) { return new or_scene(); }
void or_scene_destroy(or_scene* s) { delete s; }

uint32_t or_scene_add_mesh(or_scene* s, const float* positions3, const float* normals3, uint32_t nverts) {
    return or_scene_add_mesh_uv(s, positions3, normals3, nullptr, nverts);
}

int or_scene_set_environment(or_scene* s, const gsb_environment* env) {
    s->env = or_scene::Environment();
    if (!env || env->type == GSB_ENVIRONMENT_NONE) return 0;
    s->env.desc = *env;
    s->env.desc.texels = nullptr;
    if (env->type != GSB_ENVIRONMENT_LATLONG) return 0;
    if (!env->texels || !env->width || !env->height) return 1;
    const uint32_t W = env->width, H = env->height;
    s->env.texels.assign(env->texels, env->texels + (size_t)W * H * 3);
    // the sampling tables, by the rule stated at buildEnvironmentTables (gpuspectral_b200/csrc/api.cu): double sums, float tables
    s->env.rowCdf.assign((size_t)H + 1, 0.0f);
    s->env.colCdf.assign((size_t)H * (W + 1), 0.0f);
    std::vector<double> rowW(H);
    double total = 0.0;
    for (uint32_t y = 0; y < H; y++) {
        std::vector<double> c((size_t)W + 1, 0.0);
        for (uint32_t x = 0; x < W; x++) {
            const float* p = env->texels + 3 * ((size_t)y * W + x);
            const double lum = 0.2126 * (double)(p[0] * env->scale[0]) + 0.7152 * (double)(p[1] * env->scale[1]) + 0.0722 * (double)(p[2] * env->scale[2]);
            c[x + 1] = c[x] + (lum > 0.0 ? lum : 0.0);
        }
        for (uint32_t x = 0; x <= W; x++) s->env.colCdf[(size_t)y * (W + 1) + x] = (float)(c[W] > 0.0 ? c[x] / c[W] : (double)x / W);
        s->env.colCdf[(size_t)y * (W + 1) + W] = 1.0f;
        rowW[y] = c[W] * std::sin(3.14159265358979323846 * (y + 0.5) / H);
        total += rowW[y];
    }
    double acc = 0.0;
    for (uint32_t y = 0; y < H; y++) {
        acc += rowW[y];
        s->env.rowCdf[y + 1] = (float)(total > 0.0 ? acc / total : (double)(y + 1) / H);
    }
    s->env.rowCdf[H] = 1.0f;
    return 0;
}

// probes of the environment emitter: direction / radiance / solid-angle density for pairs of uniform numbers, and the
// radiance looked up for given directions
void or_env_sample(const or_scene* s, const float* e1, const float* e2, uint32_t n, float* dirs3, float* radiance3, float* pdf) {
    for (uint32_t i = 0; i < n; i++) {
        vec3 rad;
        float p = 0.0f;
        const vec3 d = envSample(s->env, e1[i], e2[i], rad, p);
        dirs3[3 * i] = d.x; dirs3[3 * i + 1] = d.y; dirs3[3 * i + 2] = d.z;
        radiance3[3 * i] = rad.x; radiance3[3 * i + 1] = rad.y; radiance3[3 * i + 2] = rad.z;
        pdf[i] = p;
    }
}
void or_env_radiance(const or_scene* s, const float* dirs3, uint32_t n, float* radiance3) {
    for (uint32_t i = 0; i < n; i++) {
        const vec3 r = envRadiance(s->env, vec3(dirs3[3 * i], dirs3[3 * i + 1], dirs3[3 * i + 2]));
        radiance3[3 * i] = r.x; radiance3[3 * i + 1] = r.y; radiance3[3 * i + 2] = r.z;
    }
}

int or_scene_set_textures(or_scene* s, const gsb_texture* textures, uint32_t count) {
    s->textures.clear();
    for (uint32_t i = 0; i < count; i++) {
        or_scene::Texture t;
        t.desc = textures[i];
        if (t.desc.type == GSB_TEXTURE_BITMAP) {
            if (!t.desc.texels || !t.desc.width || !t.desc.height) return 1;
            t.texels.assign(t.desc.texels, t.desc.texels + (size_t)t.desc.width * t.desc.height * 3);
        }
        t.desc.texels = nullptr;
        s->textures.push_back(std::move(t));
    }
    return 0;
}

uint32_t or_scene_add_mesh_uv(or_scene* s, const float* positions3, const float* normals3, const float* uv2, uint32_t nverts) {
    Mesh m;
    m.pos.assign(positions3, positions3 + 3 * (size_t)nverts);
    m.nrm.assign(normals3, normals3 + 3 * (size_t)nverts);
    if (uv2) m.uv.assign(uv2, uv2 + 2 * (size_t)nverts);
    m.ntris = nverts / 3;
    buildMeshBvh(m);
    s->meshes.push_back(std::move(m));
    return (uint32_t)s->meshes.size() - 1;
}

int or_scene_set_instances(or_scene* s, const gsb_instance* instances, uint32_t count) {
    s->instances.clear();
    uint64_t baked = 0;
    for (uint32_t i = 0; i < count; i++) {
        if (instances[i].mesh >= s->meshes.size()) return 1;
        Instance in;
        in.src = instances[i];
        deriveInstance(in);
        instanceBounds(*s, in);
        const Mesh& m = s->meshes[in.src.mesh];
        if (m.ntris > 0 && m.ntris <= kBakeMaxTris && baked + m.ntris <= kBakeBudget) {
            baked += m.ntris;
            in.world = std::make_shared<Mesh>();
            in.world->ntris = m.ntris;
            in.world->pos.resize(m.pos.size());
            for (size_t v = 0; v < m.pos.size() / 3; v++) {
                const float* p = &m.pos[3 * v];
                for (int r = 0; r < 3; r++) {
                    const float* row = in.o2w + 4 * r;
                    in.world->pos[3 * v + r] = ((row[0] * p[0] + row[1] * p[1]) + row[2] * p[2]) + row[3];
                }
            }
            buildMeshBvh(*in.world);
            for (int a = 0; a < 3; a++) { in.wlo[a] = in.world->lo[a]; in.whi[a] = in.world->hi[a]; }
        }
        s->instances.push_back(in);
    }
    return 0;
}

uint32_t or_scene_baked_instances(const or_scene* s) {
    uint32_t n = 0;
    for (const Instance& in : s->instances) n += in.world ? 1u : 0u;
    return n;
}

int or_scene_set_bsdfs(or_scene* s, uint32_t type, const void* params, uint32_t count) {
#define OR_ASSIGN(vec, T) { const T* p = (const T*)params; s->vec.assign(p, p + count); return 0; }
    switch (type) {
        case GSB_BSDF_DIFFUSE: OR_ASSIGN(diffuse, gsb_diffuse_bsdf)
        case GSB_BSDF_SMOOTH_DIELECTRIC: OR_ASSIGN(smoothDielectric, gsb_smooth_dielectric_bsdf)
        case GSB_BSDF_SMOOTH_CONDUCTOR: OR_ASSIGN(smoothConductor, gsb_smooth_conductor_bsdf)
        case GSB_BSDF_SMOOTH_PLASTIC: OR_ASSIGN(smoothPlastic, gsb_smooth_plastic_bsdf)
        case GSB_BSDF_ROUGH_CONDUCTOR: OR_ASSIGN(roughConductor, gsb_rough_conductor_bsdf)
        case GSB_BSDF_SMOOTH_FLOOR: OR_ASSIGN(smoothFloor, gsb_smooth_floor_bsdf)
        case GSB_BSDF_ROUGH_FLOOR: OR_ASSIGN(roughFloor, gsb_rough_floor_bsdf)
        case GSB_BSDF_ROUGH_PLASTIC: OR_ASSIGN(roughPlastic, gsb_rough_plastic_bsdf)
        case GSB_BSDF_ROUGH_DIELECTRIC: OR_ASSIGN(roughDielectric, gsb_rough_dielectric_bsdf)
    }
#undef OR_ASSIGN
    return 1;
}

int or_scene_set_lights(or_scene* s, const gsb_triangle_light* lights, uint32_t count) {
    s->lights.assign(lights, lights + count);
    return 0;
}

void or_scene_set_camera(or_scene* s, const float eye[3], const float to_world[16], float fov_radians) {
    std::memcpy(s->eye, eye, sizeof(float) * 3);
    std::memcpy(s->view, to_world, sizeof(float) * 16);
    s->fov = fov_radians;
}

void or_scene_set_params(or_scene* s, const gsb_params* p) { s->params = *p; }

int or_render(or_scene* s, uint32_t width, uint32_t height, uint32_t first, uint32_t n, uint32_t stride,
              float* image, int sum_mode, int threads, or_counters* counters) {
    int used = 1;
#ifdef _OPENMP
    used = threads > 0 ? threads : omp_get_max_threads();
#endif
    uint64_t ext = 0, shadow = 0, draws = 0;
    for (uint32_t i = 0; i < (uint32_t)s->instances.size(); i++) {
        const uint32_t h = s->instances[i].src.bsdf;
        if ((h >> 16) >= GSB_BSDF_TYPE_COUNT || (h & 0xffffu) >= bsdfCount(*s, h >> 16)) return -1;
    }
#pragma omp parallel for schedule(dynamic, 4) num_threads(used) reduction(+ : ext, shadow, draws)
    for (int64_t y = 0; y < (int64_t)height; y++) {
        PathStats st;
        for (uint32_t x = 0; x < width; x++) {
            float* px = image + 4 * ((size_t)y * width + x);
            for (uint32_t k = 0; k < n; k++) {
                const uint32_t ts = first + k * stride;
                vec3 c = pathSample(*s, width, height, x, (uint32_t)y, ts, st, nullptr, 0, nullptr);
                if (sum_mode) {
                    if (!(std::isnan(c.x) || std::isnan(c.y) || std::isnan(c.z))) {
                        px[0] += c.x; px[1] += c.y; px[2] += c.z; px[3] += 1.0f;
                    }
                } else {
                    vec3 accum = c;                                   // raygen.rgen:84-108
                    if (ts > 0) {
                        const float a = 1.0f / (float)(ts + 1);
                        vec3 prev(px[0], px[1], px[2]);
                        accum = prev * (1.0f - a) + accum * a;        // mix(prev, accum, a)
                    }
                    if (!(std::isnan(accum.x) || std::isnan(accum.y) || std::isnan(accum.z))) {
                        px[0] = accum.x; px[1] = accum.y; px[2] = accum.z; px[3] = 1.0f;
                    }
                }
            }
        }
        ext += st.ext; shadow += st.shadow; draws += st.draws;
    }
    if (counters) {
        counters->paths = (uint64_t)width * height * n;
        counters->extension_rays = ext;
        counters->shadow_rays = shadow;
        counters->rng_draws = draws;
    }
    return used;
}

int or_path_probe(or_scene* s, uint32_t width, uint32_t height, uint32_t x, uint32_t y, uint32_t timestamp,
                  float out_rgb[3], or_bounce* bounces, uint32_t max_bounces) {
    PathStats st;
    uint32_t nb = 0;
    vec3 c = pathSample(*s, width, height, x, y, timestamp, st, bounces, max_bounces, &nb);
    out_rgb[0] = c.x; out_rgb[1] = c.y; out_rgb[2] = c.z;
    return (int)nb;
}

// One ray query in the shape oracle/ref/ref_glue.cpp's traceRayEXT takes as its callback (the ray/triangle search has no
// source in the reference; the compiled reference shaders therefore run on top of this intersector).
int or_intersect(void* scene, const float* origin, const float* dir, float tmin, float tmax, int any_hit, gsb_hit* out) {
    const or_scene* s = static_cast<const or_scene*>(scene);
    Hit h = traceF32(*s, origin, dir, tmin, tmax, any_hit != 0, true);
    out->t = h.t; out->u = h.u; out->v = h.v; out->inst = h.inst; out->prim = h.prim;
    return h.inst != GSB_NO_HIT ? 1 : 0;
}

int or_trace_batch(or_scene* s, const gsb_ray* rays, uint32_t count, gsb_hit* hits, int any_hit, int mode, int threads) {
    int used = 1;
#ifdef _OPENMP
    used = threads > 0 ? threads : omp_get_max_threads();
#endif
#pragma omp parallel for schedule(dynamic, 64) num_threads(used)
    for (int64_t i = 0; i < (int64_t)count; i++) {
        const gsb_ray& r = rays[i];
        Hit h;
        if (mode == 2) h = traceF64(*s, r.origin, r.dir, r.tmin, r.tmax, any_hit != 0);
        else h = traceF32(*s, r.origin, r.dir, r.tmin, r.tmax, any_hit != 0, mode == 1);
        gsb_hit& o = hits[i];
        if (any_hit) {
            o.t = 0.0f; o.u = 0.0f; o.v = 0.0f;
            o.inst = h.inst == GSB_NO_HIT ? GSB_NO_HIT : 0u;
            o.prim = o.inst;
        } else {
            o.t = h.t; o.u = h.u; o.v = h.v; o.inst = h.inst; o.prim = h.prim;
        }
    }
    return used;
}

uint32_t or_tea(uint32_t v0, uint32_t v1) { return tea(v0, v1); }
uint32_t or_pcg_hash(uint32_t v) { return pcgHash(v); }
float or_rand_to_float(uint32_t word) { return (float)word * (1.0f / (float)0xffffffffu); }

void or_rng(const uint32_t* idx, const uint32_t* timestamp, uint32_t n, uint32_t n_draws, uint32_t* out) {
    for (uint32_t i = 0; i < n; i++) {
        uint32_t* o = out + (size_t)i * (1 + n_draws);
        Rng rng(pcgHash(tea(idx[i], timestamp[i])));
        o[0] = rng.state;
        for (uint32_t k = 0; k < n_draws; k++) o[1 + k] = rng.randPcg();
    }
}

void or_camera_rays(or_scene* s, uint32_t width, uint32_t height, const uint32_t* px, const uint32_t* py,
                    uint32_t n, gsb_ray* out) {
    for (uint32_t i = 0; i < n; i++) {
        vec3 o, d;
        primaryRay(*s, width, height, px[i], py[i], o, d);
        out[i].origin[0] = o.x; out[i].origin[1] = o.y; out[i].origin[2] = o.z;
        out[i].dir[0] = d.x; out[i].dir[1] = d.y; out[i].dir[2] = d.z;
        out[i].tmin = 0.0f;
        out[i].tmax = 1e10f;
    }
}

int or_bsdf_sample(or_scene* s, uint32_t handle, const float* wo, const uint32_t* seed, uint32_t n,
                   float* wi, float* f, float* pdf, uint32_t* is_delta, uint32_t* seed_out) {
    if ((handle >> 16) >= GSB_BSDF_TYPE_COUNT || (handle & 0xffffu) >= bsdfCount(*s, handle >> 16)) return 1;
    for (uint32_t i = 0; i < n; i++) {
        Rng rng(seed[i]);
        vec3 w;
        BSDFOutput res;
        res.bsdf = vec3(0.0f); res.pdf = 0.0f; res.isDelta = false;
        sampleBSDF(*s, handle, rng, v3(wo + 3 * (size_t)i), w, res);
        wi[3 * (size_t)i] = w.x; wi[3 * (size_t)i + 1] = w.y; wi[3 * (size_t)i + 2] = w.z;
        f[3 * (size_t)i] = res.bsdf.x; f[3 * (size_t)i + 1] = res.bsdf.y; f[3 * (size_t)i + 2] = res.bsdf.z;
        pdf[i] = res.pdf;
        is_delta[i] = res.isDelta ? 1u : 0u;
        seed_out[i] = rng.state;
    }
    return 0;
}

int or_bsdf_eval(or_scene* s, uint32_t handle, const float* wo, const float* wi, uint32_t n,
                 float* f, float* pdf, uint32_t* is_delta) {
    if ((handle >> 16) >= GSB_BSDF_TYPE_COUNT || (handle & 0xffffu) >= bsdfCount(*s, handle >> 16)) return 1;
    for (uint32_t i = 0; i < n; i++) {
        BSDFOutput res;
        res.bsdf = vec3(0.0f); res.pdf = 0.0f; res.isDelta = false;
        evalBSDF(*s, handle, v3(wo + 3 * (size_t)i), v3(wi + 3 * (size_t)i), res);
        f[3 * (size_t)i] = res.bsdf.x; f[3 * (size_t)i + 1] = res.bsdf.y; f[3 * (size_t)i + 2] = res.bsdf.z;
        pdf[i] = res.pdf;
        is_delta[i] = res.isDelta ? 1u : 0u;
    }
    return 0;
}

int or_sample_light(or_scene* s, const float* pos, const uint32_t* seed, uint32_t n,
                    float* lpos, float* emission, float* pdf, uint32_t* seed_out) {
    for (uint32_t i = 0; i < n; i++) {
        Rng rng(seed[i]);
        bool have = false;
        LightOutput r = sampleLight(*s, rng, v3(pos + 3 * (size_t)i), have);
        lpos[3 * (size_t)i] = r.position.x; lpos[3 * (size_t)i + 1] = r.position.y; lpos[3 * (size_t)i + 2] = r.position.z;
        emission[3 * (size_t)i] = r.emission.x; emission[3 * (size_t)i + 1] = r.emission.y; emission[3 * (size_t)i + 2] = r.emission.z;
        pdf[i] = r.pdf;
        seed_out[i] = rng.state;
    }
    return 0;
}

void or_encode_srgb8(const float* rgba, uint32_t npixels, uint8_t* out) {
    for (uint32_t i = 0; i < npixels; i++) {
        for (int c = 0; c < 3; c++) out[4 * (size_t)i + c] = (uint8_t)std::floor(srgbEncode(rgba[4 * (size_t)i + c]) * 255.0f + 0.5f);
        out[4 * (size_t)i + 3] = 255;
    }
}

}  // extern "C"

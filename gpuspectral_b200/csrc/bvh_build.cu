// bvh_build.cu — GPU construction of the two-level 8-wide BVH (see bvh.h / traverse.cuh).
//
// Pipeline per level (bottom: triangles of one mesh; top: instances):
//   1. primitive boxes (padded) + bounds of the centroids            k_triBoxes / upload
//   2. 63-bit Morton codes of the centroids                          k_morton
//   3. radix sort of (code, primitive)                               cub::DeviceRadixSort
//   4. binary radix tree over the sorted codes (Karras 2012)         k_radixTree
//   5. bottom-up box fit                                             k_fitBoxes
//   6. SAH treelet refinement of the binary tree (Karras & Aila 2013) treelet.cuh
//   7. top-down collapse into 8-wide nodes + octant slot assignment + 8-bit quantisation
//      (Ylitie, Karras & Laine 2017), level by level                 k_collapse
//   8. leaf records gathered in node order                           k_packTriangles / k_packInstances
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <vector>

#include "bvh.h"

namespace gsb {

namespace {

bool g_treelets = true;

// ---- helpers ---------------------------------------------------------------------------------
__host__ __device__ inline uint32_t orderedFromFloat(float f) {
    uint32_t u;
#ifdef __CUDA_ARCH__
    u = __float_as_uint(f);
#else
    memcpy(&u, &f, 4);
#endif
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ inline float floatFromOrdered(uint32_t u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
    float f;
#ifdef __CUDA_ARCH__
    f = __uint_as_float(u);
#else
    memcpy(&f, &u, 4);
#endif
    return f;
}

struct BoundsAcc { uint32_t lo[3], hi[3]; };  // ordered-uint encoded

__global__ void k_initBounds(BoundsAcc* b) {
    if (threadIdx.x < 3) {
        b->lo[threadIdx.x] = 0xffffffffu;
        b->hi[threadIdx.x] = 0u;
    }
}

// Everything one build keeps on the device for the host to read ONCE, when the build has run: the build itself never
// waits for the host (round 2 synchronised 6 times per mesh: bounds, two collapse batches, SAH cost, final sizes, timer).
constexpr int kMaxLevels = 64;   // of 8-wide nodes
struct BuildFrame {
    float cmin[3], cinv[3];   // Morton frame: centroid -> [0, 1)
    float pad;                // absolute padding of the primitive boxes
    float lo[3], hi[3];       // padded bounds of the level
};
struct BuildControl {
    uint32_t counters[4];     // [0] nodes allocated, [1] leaf positions allocated, [2] blocks of the running wide launch that have ended
    uint32_t level;           // the level the next collapse launch works on
    uint32_t reserved;
    double cost[2];           // SAH cost of the binary tree before / after the treelet passes
    float wideCost;           // SAH cost of the 8-wide tree the cut tables promise (root entry of k_collapseCost), 0 without them
    float reserved2;
    BuildFrame frame;
    uint32_t levelCount[kMaxLevels + 2];   // items (= nodes) per level
};

// the frame of a triangle level from its vertex bounds, with the float operations the host used to do
__global__ void k_blasFrame(const BoundsAcc* acc, BuildFrame* f) {
    float scale = 0.0f;
    float lo[3], hi[3];
    for (int a = 0; a < 3; a++) {
        lo[a] = floatFromOrdered(acc->lo[a]);
        hi[a] = floatFromOrdered(acc->hi[a]);
        scale = fmaxf(scale, fmaxf(fabsf(lo[a]), fabsf(hi[a])));
    }
    const float pad = __fmul_rn(scale, kBoxPadRel);
    f->pad = pad;
    for (int a = 0; a < 3; a++) {
        f->cmin[a] = lo[a];
        f->cinv[a] = hi[a] > lo[a] ? __fdiv_rn(1.0f, __fsub_rn(hi[a], lo[a])) : 0.0f;
        f->lo[a] = __fsub_rn(lo[a], pad);
        f->hi[a] = __fadd_rn(hi[a], pad);
    }
}

// bounds of all vertices of a triangle soup
__global__ void k_vertexBounds(const float4* __restrict__ v, uint32_t nverts, BoundsAcc* acc) {
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < nverts; i += gridDim.x * blockDim.x) {
        const float4 p = __ldg(v + i);
        lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
        hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
    }
#pragma unroll
    for (int a = 0; a < 3; a++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; a++) {
            atomicMin(&acc->lo[a], orderedFromFloat(lo[a]));
            atomicMax(&acc->hi[a], orderedFromFloat(hi[a]));
        }
    }
}

// padded primitive boxes of triangles
__global__ void k_triBoxes(const float4* __restrict__ v, uint32_t ntris, const BuildFrame* __restrict__ frame, float4* __restrict__ plo,
                           float4* __restrict__ phi) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= ntris) return;
    const float pad = frame->pad;
    const float4 a = __ldg(v + 3 * (size_t)i), b = __ldg(v + 3 * (size_t)i + 1), c = __ldg(v + 3 * (size_t)i + 2);
    plo[i] = make_float4(fminf(a.x, fminf(b.x, c.x)) - pad, fminf(a.y, fminf(b.y, c.y)) - pad, fminf(a.z, fminf(b.z, c.z)) - pad, 0.0f);
    phi[i] = make_float4(fmaxf(a.x, fmaxf(b.x, c.x)) + pad, fmaxf(a.y, fmaxf(b.y, c.y)) + pad, fmaxf(a.z, fmaxf(b.z, c.z)) + pad, 0.0f);
}

__device__ inline uint64_t spread21(uint32_t x) {
    uint64_t v = x & 0x1fffffu;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

__global__ void k_morton(const float4* __restrict__ plo, const float4* __restrict__ phi, uint32_t n, const BuildFrame* __restrict__ frame,
                         uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 cmin = make_float3(frame->cmin[0], frame->cmin[1], frame->cmin[2]);
    const float3 cinv = make_float3(frame->cinv[0], frame->cinv[1], frame->cinv[2]);
    const float4 lo = plo[i], hi = phi[i];
    const float cx = (0.5f * (lo.x + hi.x) - cmin.x) * cinv.x;
    const float cy = (0.5f * (lo.y + hi.y) - cmin.y) * cinv.y;
    const float cz = (0.5f * (lo.z + hi.z) - cmin.z) * cinv.z;
    const float s = 2097152.0f;  // 2^21
    const uint32_t qx = (uint32_t)fminf(fmaxf(cx * s, 0.0f), s - 1.0f);
    const uint32_t qy = (uint32_t)fminf(fmaxf(cy * s, 0.0f), s - 1.0f);
    const uint32_t qz = (uint32_t)fminf(fmaxf(cz * s, 0.0f), s - 1.0f);
    keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    vals[i] = i;
}

// Binary tree arrays.  Node references: >= 0 internal node index, < 0 leaf at sorted position ~ref.
struct BinTree {
    int* left;        // [n-1]
    int* right;       // [n-1]
    int* parent;      // [n-1] parent of internal nodes (root: -1)
    int* leafParent;  // [n]
    uint32_t* first;  // [n-1] sorted range covered
    uint32_t* last;   // [n-1]
    float4* lo;       // [n-1] boxes of internal nodes
    float4* hi;       // [n-1]
    uint32_t* flags;  // [n-1] arrival counters
};

__device__ inline int deltaKeys(const uint64_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

__global__ void k_radixTree(const uint64_t* __restrict__ keys, int n, BinTree t) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (deltaKeys(keys, n, i, i + 1) - deltaKeys(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = deltaKeys(keys, n, i, i - d);
    int lmax = 2;
    while (deltaKeys(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int s = lmax / 2; s >= 1; s /= 2)
        if (deltaKeys(keys, n, i, i + (l + s) * d) > dmin) l += s;
    const int j = i + l * d;
    const int dnode = deltaKeys(keys, n, i, j);
    int s = 0, tt = l;
    do {
        tt = (tt + 1) >> 1;
        if (deltaKeys(keys, n, i, i + (s + tt) * d) > dnode) s += tt;
    } while (tt > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    int l_ref, r_ref;
    if (lo == gamma) { l_ref = ~gamma; t.leafParent[gamma] = i; } else { l_ref = gamma; t.parent[gamma] = i; }
    if (hi == gamma + 1) { r_ref = ~(gamma + 1); t.leafParent[gamma + 1] = i; } else { r_ref = gamma + 1; t.parent[gamma + 1] = i; }
    t.left[i] = l_ref;
    t.right[i] = r_ref;
    t.first[i] = (uint32_t)lo;
    t.last[i] = (uint32_t)hi;
    if (i == 0) t.parent[0] = -1;
}

__device__ inline void refBox(const BinTree& t, const float4* plo, const float4* phi, const uint32_t* sorted, int ref, float4& lo, float4& hi) {
    if (ref >= 0) {
        lo = __ldcg(t.lo + ref);
        hi = __ldcg(t.hi + ref);
    } else {
        const uint32_t p = sorted[~ref];
        lo = plo[p];
        hi = phi[p];
    }
}

__global__ void k_fitBoxes(BinTree t, const float4* __restrict__ plo, const float4* __restrict__ phi,
                           const uint32_t* __restrict__ sorted, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int node = t.leafParent[i];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&t.flags[node], 1u) == 0u) return;  // first arrival: the sibling will finish
        float4 alo, ahi, blo, bhi;
        refBox(t, plo, phi, sorted, t.left[node], alo, ahi);
        refBox(t, plo, phi, sorted, t.right[node], blo, bhi);
        t.lo[node] = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.0f);
        t.hi[node] = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
        node = t.parent[node];
    }
}

__device__ inline float boxArea(float4 lo, float4 hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}

// SAH cost of the binary tree: sum of internal-node areas (C_i = 1) + leaf areas (C_t = 1), over root area
__global__ void k_sahCost(BinTree t, const float4* __restrict__ plo, const float4* __restrict__ phi, int n, double* out) {
    double acc = 0.0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        if (i < n - 1) acc += 1.2 * (double)boxArea(t.lo[i], t.hi[i]);
        acc += (double)boxArea(plo[i], phi[i]);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

}  // namespace

}  // namespace gsb

#include "treelet.cuh"

namespace gsb {
namespace {

// ---- collapse to 8-wide nodes ---------------------------------------------------------------------
struct CollapseItem { int binNode; uint32_t outNode; };

struct CollapseArgs {
    BinTree t;
    const float4* plo;
    const float4* phi;
    const uint32_t* sorted;
    uint4* nodes;          // output nodes, 5 x uint4 each
    uint32_t* leafPrim;    // output: primitive index per leaf position
    BuildControl* ctl;     // node / leaf allocation counters, the current level and the items per level
    uint32_t maxLeaf;      // primitives per leaf child (one box per leaf child): 1..3
    // SAH-aware cut (k_collapseCost): per binary node, cd[i-1] = least cost of standing for its subtree with at most i
    // children of an 8-wide node, ch[i-1] = how (see k_collapseCost); nullptr = greedy cut by surface area
    const float* cutCost;  // [n-1][7]
    const uint8_t* cutHow; // [n-1][8]
    float costNode, costPrim;
};

// Quantisation frame of one axis: the smallest exponent e with 255 * 2^e >= ext (e >= -126), from the float's own exponent
// and mantissa (ext = m * 2^x, m in [0.5, 1): e = x - 8, one more when m > 255/256) - it used to be ceil(log2(ext / 255)) in
// double plus a fix-up loop, ~150 instructions per axis.  inv = 2^-e: the grid coordinates are (v - origin) * inv, exact like
// the division by 2^e they replace (48 double divisions per node, the bulk of the collapse and refit kernels).
__device__ inline int quantExponent(float ext, double& scale, double& inv) {
    int e = -126;
    if (ext > 0.0f) {
        int x;
        const float m = frexpf(ext, &x);
        e = max(x - 8 + (m > 0.99609375f ? 1 : 0), -126);
    }
    scale = __hiloint2double((1023 + e) << 20, 0);
    inv = __hiloint2double((1023 - e) << 20, 0);
    return e;
}

__device__ inline uint32_t refCount(const BinTree& t, int ref) { return ref >= 0 ? t.last[ref] - t.first[ref] + 1u : 1u; }
__device__ inline uint32_t refFirst(const BinTree& t, int ref) { return ref >= 0 ? t.first[ref] : (uint32_t)~ref; }

// leafMask: bit k set = child k becomes a leaf slot (its primitives tested in place) even though it is a binary node;
// single primitives always are, and without the cost tables every reference of at most maxLeaf primitives is.
__device__ void emitNode(const CollapseArgs& a, uint32_t level, uint32_t outNode, float4 nlo, float4 nhi, const int* child, int nc,
                         CollapseItem* next, uint32_t leafMask = 0xffu) {
    // classify + boxes
    float4 clo[8], chi[8];
    bool internal[8];
    for (int k = 0; k < nc; k++) {
        refBox(a.t, a.plo, a.phi, a.sorted, child[k], clo[k], chi[k]);
        internal[k] = child[k] >= 0 && (refCount(a.t, child[k]) > a.maxLeaf || !(leafMask >> k & 1u));
    }
    // octant slot assignment: slot s (bits x,y,z = 4,2,1) should hold the child lying furthest
    // towards (+ if bit set, - if clear) on each axis; traversal visits slot^octant in descending order.
    const float ccx = 0.5f * (nlo.x + nhi.x), ccy = 0.5f * (nlo.y + nhi.y), ccz = 0.5f * (nlo.z + nhi.z);
    float cost[8][8];
    for (int k = 0; k < nc; k++) {
        const float px = 0.5f * (clo[k].x + chi[k].x) - ccx, py = 0.5f * (clo[k].y + chi[k].y) - ccy, pz = 0.5f * (clo[k].z + chi[k].z) - ccz;
        for (int s = 0; s < 8; s++)
            cost[k][s] = ((s & 4) ? px : -px) + ((s & 2) ? py : -py) + ((s & 1) ? pz : -pz);
    }
    int slotOf[8], childAt[8];
    for (int s = 0; s < 8; s++) childAt[s] = -1;
    for (int k = 0; k < nc; k++) slotOf[k] = -1;
    for (int round = 0; round < nc; round++) {
        float best = -INFINITY;
        int bk = -1, bs = -1;
        for (int k = 0; k < nc; k++) {
            if (slotOf[k] >= 0) continue;
            for (int s = 0; s < 8; s++) {
                if (childAt[s] >= 0) continue;
                if (cost[k][s] > best) { best = cost[k][s]; bk = k; bs = s; }
            }
        }
        slotOf[bk] = bs;
        childAt[bs] = bk;
    }
    // allocation
    uint32_t nInternal = 0, nLeafPrims = 0;
    for (int s = 0; s < 8; s++) {
        const int k = childAt[s];
        if (k < 0) continue;
        if (internal[k]) nInternal++;
        else nLeafPrims += refCount(a.t, child[k]);
    }
    const uint32_t childBase = nInternal ? atomicAdd(&a.ctl->counters[0], nInternal) : 0u;
    const uint32_t leafBase = nLeafPrims ? atomicAdd(&a.ctl->counters[1], nLeafPrims) : 0u;
    const uint32_t itemBase = nInternal ? atomicAdd(&a.ctl->levelCount[level + 1], nInternal) : 0u;

    // quantisation frame
    const float p[3] = {nlo.x, nlo.y, nlo.z};
    const float ext[3] = {nhi.x - nlo.x, nhi.y - nlo.y, nhi.z - nlo.z};
    int e[3];
    double scale[3], inv[3];
    for (int ax = 0; ax < 3; ax++) e[ax] = quantExponent(ext[ax], scale[ax], inv[ax]);
    uint32_t meta[8], qlo[3][8], qhi[3][8];
    uint32_t imask = 0, rank = 0, leafOff = 0;
    for (int s = 0; s < 8; s++) {
        meta[s] = 0;
        for (int ax = 0; ax < 3; ax++) { qlo[ax][s] = 0; qhi[ax][s] = 0; }
        const int k = childAt[s];
        if (k < 0) continue;
        const float lo3[3] = {clo[k].x, clo[k].y, clo[k].z}, hi3[3] = {chi[k].x, chi[k].y, chi[k].z};
        for (int ax = 0; ax < 3; ax++) {
            double ql = floor(((double)lo3[ax] - (double)p[ax]) * inv[ax]);
            double qh = ceil(((double)hi3[ax] - (double)p[ax]) * inv[ax]);
            ql = fmin(fmax(ql, 0.0), 255.0);
            qh = fmin(fmax(qh, 0.0), 255.0);
            qlo[ax][s] = (uint32_t)ql;
            qhi[ax][s] = (uint32_t)qh;
        }
        if (internal[k]) {
            imask |= 1u << s;
            meta[s] = 0x20u | (24u + (uint32_t)s);
            next[itemBase + rank] = CollapseItem{child[k], childBase + rank};
            rank++;
        } else {
            const uint32_t cnt = refCount(a.t, child[k]);
            const uint32_t first = refFirst(a.t, child[k]);
            meta[s] = (((1u << cnt) - 1u) << 5) | leafOff;
            for (uint32_t j = 0; j < cnt; j++) a.leafPrim[leafBase + leafOff + j] = a.sorted[first + j];
            leafOff += cnt;
        }
    }
    auto pack4 = [](const uint32_t* b) { return b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24); };
    uint4* out = a.nodes + (size_t)outNode * 5;
    out[0] = make_uint4(__float_as_uint(p[0]), __float_as_uint(p[1]), __float_as_uint(p[2]),
                        (uint32_t)(e[0] + 127) | ((uint32_t)(e[1] + 127) << 8) | ((uint32_t)(e[2] + 127) << 16) | (imask << 24));
    out[1] = make_uint4(childBase, leafBase, pack4(meta), pack4(meta + 4));
    out[2] = make_uint4(pack4(qlo[0]), pack4(qlo[0] + 4), pack4(qlo[1]), pack4(qlo[1] + 4));
    out[3] = make_uint4(pack4(qlo[2]), pack4(qlo[2] + 4), pack4(qhi[0]), pack4(qhi[0] + 4));
    out[4] = make_uint4(pack4(qhi[1]), pack4(qhi[1] + 4), pack4(qhi[2]), pack4(qhi[2] + 4));
}

// SAH cost tables for the cut (Ylitie, Karras & Laine 2017, section 3.1, on the binary tree the treelet pass left).
// For a binary node b with P primitives and box area A:
//   cd[0] = min( A * P * costPrim  if P <= maxLeaf            (b is a leaf slot),
//                A * costNode + min_k L[k] + R[8 - k] )        (b is an 8-wide node; k children come from its left subtree)
//   cd[i] = min( cd[i - 1], min_k L[k] + R[i + 1 - k] )        (b's subtree shows up as at most i + 1 children of an ancestor)
// with L / R the tables of its two children (a single primitive: its area * costPrim for every i).  ch[0] = 0 for the leaf
// slot, else the k of the 8-way split; ch[i] = 0 for "as with one child fewer", else the k of the split.
// Bottom-up with arrival counters like k_fitBoxes: the second thread to reach a node finds both tables written.
__global__ void k_collapseCost(BinTree t, const float4* __restrict__ plo, const float4* __restrict__ phi, const uint32_t* __restrict__ sorted,
                               int n, uint32_t maxLeaf, float costNode, float costPrim, float* cd, uint8_t* ch) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int node = t.leafParent[i];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&t.flags[node], 1u) == 0u) return;
        float tab[2][8];
        const int refs[2] = {t.left[node], t.right[node]};
#pragma unroll
        for (int c = 0; c < 2; c++) {
            if (refs[c] >= 0) {
#pragma unroll
                for (int k = 0; k < 7; k++) tab[c][k + 1] = __ldcg(cd + 7 * (size_t)refs[c] + k);
            } else {
                const uint32_t p = sorted[~refs[c]];
                const float leaf = boxArea(plo[p], phi[p]) * costPrim;
#pragma unroll
                for (int k = 1; k < 8; k++) tab[c][k] = leaf;
            }
        }
        const float area = boxArea(t.lo[node], t.hi[node]);
        const uint32_t prims = t.last[node] - t.first[node] + 1u;
        float best = INFINITY;
        int how = 1;
#pragma unroll
        for (int k = 1; k < 8; k++) {
            const float c = tab[0][k] + tab[1][8 - k];
            if (c < best) { best = c; how = k; }
        }
        best += area * costNode;
        if (prims <= maxLeaf) {
            const float leaf = area * (float)prims * costPrim;
            if (leaf <= best) { best = leaf; how = 0; }
        }
        float out[7];
        uint8_t outHow[8];
        out[0] = best;
        outHow[0] = (uint8_t)how;
#pragma unroll
        for (int m = 2; m < 8; m++) {      // at most m children
            float bm = out[m - 2];
            int hm = 0;
#pragma unroll
            for (int k = 1; k < m; k++) {
                const float c = tab[0][k] + tab[1][m - k];
                if (c < bm) { bm = c; hm = k; }
            }
            out[m - 1] = bm;
            outHow[m - 1] = (uint8_t)hm;
        }
        outHow[7] = 0;
#pragma unroll
        for (int k = 0; k < 7; k++) __stcg(cd + 7 * (size_t)node + k, out[k]);
        uint2 packed;
        packed.x = outHow[0] | (outHow[1] << 8) | (outHow[2] << 16) | ((uint32_t)outHow[3] << 24);
        packed.y = outHow[4] | (outHow[5] << 8) | (outHow[6] << 16);
        __stcg(reinterpret_cast<uint2*>(ch + 8 * (size_t)node), packed);
        node = t.parent[node];
    }
}

// The children of the 8-wide node that stands for binary node b, by the cost tables: `budget` children for the subtree of
// `ref`, split as the tables say until a piece has a budget of one (or is a single primitive).
__device__ int cutBySah(const CollapseArgs& a, int b, int* child, uint32_t& leafMask) {
    // a node's eight "how" bytes come in one load, together with its two child references: a piece of the cut costs one
    // round trip to L2, not one per byte looked at
    auto howOf = [&](int ref) {
        const uint2 h = __ldg(reinterpret_cast<const uint2*>(a.cutHow + 8 * (size_t)ref));
        return (unsigned long long)h.x | ((unsigned long long)h.y << 32);
    };
    int stackRef[8], stackBudget[8];
    int sp = 0, nc = 0;
    leafMask = 0u;
    int k8 = (int)(howOf(b) & 0xffull);
    if (k8 == 0) k8 = 1;   // the root of a tree of at most maxLeaf primitives still becomes a node: one child per side
    stackRef[sp] = a.t.right[b]; stackBudget[sp++] = 8 - k8;
    stackRef[sp] = a.t.left[b]; stackBudget[sp++] = k8;
    while (sp > 0) {
        int ref = stackRef[--sp], budget = stackBudget[sp];
        for (;;) {
            if (ref < 0) { child[nc++] = ref; break; }
            const unsigned long long how = howOf(ref);
            const int l = a.t.left[ref], r = a.t.right[ref];   // asked for with the bytes: used when the piece is split
            while (budget > 1 && ((how >> (8 * (budget - 1))) & 0xffull) == 0ull) budget--;
            if (budget == 1) {
                if ((how & 0xffull) == 0ull) leafMask |= 1u << nc;
                child[nc++] = ref;
                break;
            }
            const int k = (int)((how >> (8 * (budget - 1))) & 0xffull);
            stackRef[sp] = r; stackBudget[sp++] = budget - k;
            ref = l;
            budget = k;
        }
    }
    return nc;
}

// One item of a level: the binary node it stands for is opened - by the SAH cost tables when the build made them, else
// greedily (the internal child with the largest surface area first) until 8 slots are used - and the 8-wide node is emitted.
__device__ void collapseItem(const CollapseArgs& a, uint32_t level, const CollapseItem it, CollapseItem* next) {
    const int b = it.binNode;
    int child[8];
    if (a.cutHow) {
        uint32_t leafMask;
        const int n = cutBySah(a, b, child, leafMask);
        emitNode(a, level, it.outNode, a.t.lo[b], a.t.hi[b], child, n, next, leafMask);
        return;
    }
    int nc = 2;
    child[0] = a.t.left[b];
    child[1] = a.t.right[b];
    while (nc < 8) {
        float bestArea = -1.0f;
        int bk = -1;
        for (int k = 0; k < nc; k++) {
            if (child[k] < 0) continue;
            const float ar = boxArea(a.t.lo[child[k]], a.t.hi[child[k]]);
            if (ar > bestArea) { bestArea = ar; bk = k; }
        }
        if (bk < 0) break;
        const int x = child[bk];
        child[bk] = a.t.left[x];
        child[nc++] = a.t.right[x];
    }
    emitNode(a, level, it.outNode, a.t.lo[b], a.t.hi[b], child, nc, next);
}

// The collapse runs level by level (node indices of a level are contiguous: the refit walks them bottom-up), and the host
// does not know how many items a level holds or how deep the tree is.  Two kernels read both from the control block:
// k_collapseNarrow, ONE block, works through consecutive levels while they hold at most `limit` items (limit 0: to the end
// of the tree) - the first levels (1, 8, 64, 512 items) and whatever a deep, thin tree has left; k_collapseWide, a fixed
// grid, does one level with a grid-stride loop, and the last of its blocks to end moves the control block to the next level.
// The items of level l are read from buffer l & 1 and those of level l + 1 appended to the other one.
__global__ void __launch_bounds__(256) k_collapseNarrow(CollapseArgs a, CollapseItem* bufA, CollapseItem* bufB, uint32_t limit) {
    for (;;) {
        const uint32_t level = *(volatile uint32_t*)&a.ctl->level;
        if (level >= (uint32_t)kMaxLevels) return;
        const uint32_t cnt = *(volatile uint32_t*)&a.ctl->levelCount[level];
        if (cnt == 0u || (limit != 0u && cnt > limit)) return;
        const CollapseItem* items = (level & 1u) ? bufB : bufA;
        CollapseItem* next = (level & 1u) ? bufA : bufB;
        for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) {
            CollapseItem it;
            it.binNode = __ldcg(&items[i].binNode);     // written by this kernel one level earlier: not through the read-only path
            it.outNode = __ldcg(&items[i].outNode);
            collapseItem(a, level, it, next);
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) *(volatile uint32_t*)&a.ctl->level = level + 1u;
        __threadfence();
        __syncthreads();
    }
}

__global__ void __launch_bounds__(128) k_collapseWide(CollapseArgs a, CollapseItem* bufA, CollapseItem* bufB) {
    // every block reads the same level: it only changes when all blocks of this launch have ended
    const uint32_t level = *(volatile uint32_t*)&a.ctl->level;
    const uint32_t cnt = level < (uint32_t)kMaxLevels ? *(volatile uint32_t*)&a.ctl->levelCount[level] : 0u;
    const CollapseItem* items = (level & 1u) ? bufB : bufA;
    CollapseItem* next = (level & 1u) ? bufA : bufB;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += gridDim.x * blockDim.x) collapseItem(a, level, items[i], next);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const uint32_t ended = atomicAdd(&a.ctl->counters[2], 1u);
        if (ended == gridDim.x - 1u) {
            a.ctl->counters[2] = 0u;
            if (cnt) a.ctl->level = level + 1u;
            __threadfence();
        }
    }
}

// start of a collapse: the root item, one node allocated, one item on level 0
__global__ void k_collapseStart(BuildControl* ctl, CollapseItem* bufA, const float* cutCost) {
    ctl->wideCost = cutCost ? cutCost[0] : 0.0f;
    ctl->counters[0] = 1u;
    ctl->counters[1] = 0u;
    ctl->counters[2] = 0u;
    ctl->level = 0u;
    ctl->levelCount[0] = 1u;
    bufA[0] = CollapseItem{0, 0u};
}

// n == 1: a root with a single leaf slot
__global__ void k_singleLeafRoot(CollapseArgs a) {
    const int child[1] = {~0};
    a.ctl->counters[0] = 1u;
    a.ctl->counters[1] = 0u;
    a.ctl->levelCount[0] = 1u;
    emitNode(a, 0u, 0u, a.plo[a.sorted[0]], a.phi[a.sorted[0]], child, 1, nullptr);
}

__global__ void k_emptyRoot(uint4* nodes) {
    nodes[0] = make_uint4(0u, 0u, 0u, 127u | (127u << 8) | (127u << 16));
    nodes[1] = make_uint4(0u, 0u, 0u, 0u);
    nodes[2] = nodes[3] = nodes[4] = make_uint4(0u, 0u, 0u, 0u);
}

// Leaf record of a triangle: three copies, one per dominant ray axis kz, each rotated so that kz comes last
// (traverse.cuh: permute) — the traversal fetches the copy that matches its ray and needs no per-vertex selects.
// Copy kz of leaf position k sits at float4 index 9 k + 3 kz; v0.w carries the primitive id in every copy.
// idsInVertices (world level): v0.w / v1.w are copied from the input (primitive id, instance id of a baked triangle).
__global__ void k_packTriangles(const uint32_t* __restrict__ leafPrim, uint32_t n, const float4* __restrict__ v, float4* __restrict__ out,
                                bool idsInVertices) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t p = leafPrim[k];
    const float4 a = __ldg(v + 3 * (size_t)p), b = __ldg(v + 3 * (size_t)p + 1), c = __ldg(v + 3 * (size_t)p + 2);
    const float pw = idsInVertices ? a.w : __uint_as_float(p), iw = idsInVertices ? b.w : 0.0f;
    float4* o = out + 9 * (size_t)k;
    o[0] = make_float4(a.y, a.z, a.x, pw); o[1] = make_float4(b.y, b.z, b.x, iw); o[2] = make_float4(c.y, c.z, c.x, 0.0f);   // kz = 0
    o[3] = make_float4(a.z, a.x, a.y, pw); o[4] = make_float4(b.z, b.x, b.y, iw); o[5] = make_float4(c.z, c.x, c.y, 0.0f);   // kz = 1
    o[6] = make_float4(a.x, a.y, a.z, pw); o[7] = make_float4(b.x, b.y, b.z, iw); o[8] = make_float4(c.x, c.y, c.z, 0.0f);   // kz = 2
}

__global__ void k_packInstances(const uint32_t* __restrict__ leafPrim, uint32_t n, const uint4* __restrict__ rec, uint4* __restrict__ out) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t p = leafPrim[k];
#pragma unroll
    for (int w = 0; w < 4; w++) out[4 * (size_t)k + w] = rec[4 * (size_t)p + w];
}

template <class T>
T* dalloc(size_t count, cudaStream_t s) {
    void* p = nullptr;
    GSB_CUDA(cudaMallocAsync(&p, std::max<size_t>(count, 1) * sizeof(T), s));
    return (T*)p;
}

// Temporaries of a build, stream-ordered.  With a capacity the pool is ONE cudaMallocAsync block that get() carves up
// (a build makes ~30 requests, and at ~20 us each the requests of the 17 meshes of the field scene cost more host time
// than the GPU needs for the builds); what does not fit, and every request of a pool without capacity, is its own block.
struct TempPool {
    cudaStream_t s;
    std::vector<void*> ptrs;
    uint8_t* arena = nullptr;
    size_t capacity = 0, used = 0;
    explicit TempPool(cudaStream_t st, size_t arenaBytes = 0) : s(st), capacity(arenaBytes) {
        if (capacity) {
            arena = dalloc<uint8_t>(capacity, s);
            ptrs.push_back(arena);
        }
    }
    template <class T>
    T* get(size_t count) {
        const size_t bytes = (std::max<size_t>(count, 1) * sizeof(T) + 255) & ~(size_t)255;
        if (used + bytes <= capacity) {
            T* p = reinterpret_cast<T*>(arena + used);
            used += bytes;
            return p;
        }
        T* p = dalloc<T>(count, s);
        ptrs.push_back(p);
        return p;
    }
    ~TempPool() {
        for (void* p : ptrs) cudaFreeAsync(p, s);
    }
};

// One build in flight.  start() queues the whole pipeline on the job's stream without waiting for the host; collect()
// reads the control block back (the one wait of a build); finish() allocates the final buffer and copies the nodes.
struct BuildJob {
    TempPool pool;
    cudaStream_t s;
    uint32_t n = 0;
    bool topLevel = false;
    bool refine = false;
    uint4* nodes = nullptr;        // temp, worst-case sized
    uint32_t* leafPrim = nullptr;  // temp: primitive index per leaf position
    uint32_t numNodes = 1;
    BuildControl* ctl = nullptr;   // device
    BuildControl host;             // its copy after collect()
    int passesOverride = -1, gammaOverride = 0;   // a candidate build (buildBlasBatch): treelet settings of its own
    // 16 + 16 primitive boxes, 80 nodes, 4 leaf order, 24 keys / values twice, 60 binary tree, 16 collapse items, 36 cut tables, 16 + 32 treelet
    // arrays per primitive, the sort's own temporaries (a second key / value buffer, 12 per primitive, and its histograms)
    // and the 256-byte rounding of ~35 requests; a request that does not fit falls back to a block of its own
    static size_t arenaBytes(uint32_t n) { return (size_t)std::max<uint32_t>(n, 1) * 328 + (128u << 10); }
    BuildJob(cudaStream_t st, uint32_t count) : pool(st, arenaBytes(count)), s(st) { memset(&host, 0, sizeof(host)); }

    // the control block, zeroed; the caller then fills ctl->frame (k_blasFrame, or a copy of a host frame)
    void begin(uint32_t count, bool top) {
        n = count;
        topLevel = top;
        ctl = pool.get<BuildControl>(1);
        GSB_CUDA(cudaMemsetAsync(ctl, 0, sizeof(BuildControl), s));
    }
    void start(const float4* plo, const float4* phi);
    void collect(DeviceBvh& out);
};

void BuildJob::start(const float4* plo, const float4* phi) {
    nodes = pool.get<uint4>((size_t)std::max<uint32_t>(n, 1) * 5);
    leafPrim = pool.get<uint32_t>(n);
    if (n == 0) {
        k_emptyRoot<<<1, 1, 0, s>>>(nodes);
        return;
    }
    const int B = 256;
    uint64_t* keys = pool.get<uint64_t>(n);
    uint64_t* keysSorted = pool.get<uint64_t>(n);
    uint32_t* vals = pool.get<uint32_t>(n);
    uint32_t* sorted = pool.get<uint32_t>(n);
    k_morton<<<(n + B - 1) / B, B, 0, s>>>(plo, phi, n, &ctl->frame, keys, vals);
    size_t tempBytes = 0;
    GSB_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tempBytes, keys, keysSorted, vals, sorted, (int)n, 0, 63, s));
    void* temp = pool.get<uint8_t>(tempBytes);
    GSB_CUDA(cub::DeviceRadixSort::SortPairs(temp, tempBytes, keys, keysSorted, vals, sorted, (int)n, 0, 63, s));

    CollapseArgs ca;
    ca.plo = plo;
    ca.phi = phi;
    ca.sorted = sorted;
    ca.nodes = nodes;
    ca.leafPrim = leafPrim;
    ca.ctl = ctl;
    // Top level: one instance per leaf child, so that every instance is culled by a box of its own before the ray pays
    // for its transform and root fetch (with 3 per leaf the gallery entered 5.8 and the field 4.7 - 6.9 instances per ray).
    // Bottom level: up to GSB_MAX_LEAF (default 3) triangles share a leaf box.
    ca.maxLeaf = topLevel ? 1u : 3u;
    if (const char* e = getenv(topLevel ? "GSB_MAX_LEAF_TLAS" : "GSB_MAX_LEAF")) ca.maxLeaf = (uint32_t)std::min(3, std::max(1, atoi(e)));
    if (n == 1) {
        memset(&ca.t, 0, sizeof(ca.t));
        k_singleLeafRoot<<<1, 1, 0, s>>>(ca);
        GSB_CUDA(cudaGetLastError());
        return;
    }
    BinTree t;
    t.left = pool.get<int>(n - 1);
    t.right = pool.get<int>(n - 1);
    t.parent = pool.get<int>(n - 1);
    t.leafParent = pool.get<int>(n);
    t.first = pool.get<uint32_t>(n - 1);
    t.last = pool.get<uint32_t>(n - 1);
    t.lo = pool.get<float4>(n - 1);
    t.hi = pool.get<float4>(n - 1);
    t.flags = pool.get<uint32_t>(n - 1);
    GSB_CUDA(cudaMemsetAsync(t.flags, 0, (size_t)(n - 1) * sizeof(uint32_t), s));
    k_radixTree<<<(n - 1 + B - 1) / B, B, 0, s>>>(keysSorted, (int)n, t);
    k_fitBoxes<<<(n + B - 1) / B, B, 0, s>>>(t, plo, phi, sorted, (int)n);
    GSB_CUDA(cudaGetLastError());

    const int costBlocks = (int)std::min<uint32_t>((n + B - 1) / B, 1024u);
    k_sahCost<<<costBlocks, B, 0, s>>>(t, plo, phi, (int)n, &ctl->cost[0]);
    // Bottom level: one pass (on one B200 the first pass buys the gallery +10 % paths/s and the field +1 %, a second one
    // +-0.5 % on both at 4 ms of build time for the field's 5.2 M triangles: profiles/README.md, r2w; GSB_TREELET_PASSES=2
    // brings it back).  Top level: off by default — on the instanced test scene the refined TLAS lowered the SAH
    // cost but raised the instances entered per ray (4.7 -> 6.1): the Morton order of the instance centroids gives a
    // better front-to-back visiting order than the SAH-optimal grouping (profiles/README.md, r1g).
    int passes = (g_treelets && !topLevel) ? 1 : 0;
    if (const char* e = getenv(topLevel ? "GSB_TREELET_PASSES_TLAS" : "GSB_TREELET_PASSES")) passes = atoi(e);
    if (passesOverride >= 0) passes = passesOverride;
    refine = passes > 0 && n >= 16;
    if (refine) {
        refineTreelets(s, t, plo, phi, sorted, (int)n, passes, [&](size_t bytes) { return (void*)pool.get<uint8_t>(bytes); }, nullptr, gammaOverride);
        ca.sorted = sorted;  // the refinement re-derives the leaf order
        k_sahCost<<<costBlocks, B, 0, s>>>(t, plo, phi, (int)n, &ctl->cost[1]);
    }

    // collapse: the first levels in one block, the wide ones one launch each, the rest (if any) in one block again
    CollapseItem* itemsA = pool.get<CollapseItem>(n);
    CollapseItem* itemsB = pool.get<CollapseItem>(n);
    ca.t = t;
    // Which binary nodes become 8-wide nodes, which leaf slots: by SAH cost tables over the binary tree (default), or
    // greedily by surface area (GSB_COLLAPSE_SAH=0; top level: GSB_COLLAPSE_SAH_TLAS).  costNode : costPrim = what one
    // 8-wide node test costs against one triangle test (top level: one instance entry).
    ca.cutCost = nullptr;
    ca.cutHow = nullptr;
    ca.costNode = 1.0f;
    ca.costPrim = topLevel ? 2.0f : 1.0f;
    bool sahCut = true;
    if (const char* e = getenv(topLevel ? "GSB_COLLAPSE_SAH_TLAS" : "GSB_COLLAPSE_SAH")) sahCut = atoi(e) != 0;
    if (const char* e = getenv(topLevel ? "GSB_SAH_COST_INST" : "GSB_SAH_COST_PRIM")) ca.costPrim = (float)atof(e);
    if (sahCut) {
        float* cd = pool.get<float>((size_t)(n - 1) * 7);
        uint8_t* ch = pool.get<uint8_t>((size_t)(n - 1) * 8);
        GSB_CUDA(cudaMemsetAsync(t.flags, 0, (size_t)(n - 1) * sizeof(uint32_t), s));
        k_collapseCost<<<(n + B - 1) / B, B, 0, s>>>(t, plo, phi, ca.sorted, (int)n, ca.maxLeaf, ca.costNode, ca.costPrim, cd, ch);
        ca.cutCost = cd;
        ca.cutHow = ch;
    }
    k_collapseStart<<<1, 1, 0, s>>>(ctl, itemsA, ca.cutCost);
    constexpr uint32_t kNarrowLimit = 512;
    k_collapseNarrow<<<1, 256, 0, s>>>(ca, itemsA, itemsB, kNarrowLimit);
    if (n > kNarrowLimit) {
        int wide = 4;   // levels past the narrow ones that a balanced tree of n leaves needs, and a margin of four
        for (double cap = (double)kNarrowLimit; cap < (double)n; cap *= 8.0) wide++;
        static int maxBlocks = 0;
        if (!maxBlocks) {
            int dev = 0, sms = 148, perSm = 4;
            cudaGetDevice(&dev);
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, (const void*)k_collapseWide, 128, 0);
            maxBlocks = sms * std::max(perSm, 1);
        }
        const int blocks = (int)std::min<uint32_t>((n + 127) / 128, (uint32_t)maxBlocks);
        for (int k = 0; k < wide; k++) k_collapseWide<<<blocks, 128, 0, s>>>(ca, itemsA, itemsB);
        k_collapseNarrow<<<1, 256, 0, s>>>(ca, itemsA, itemsB, 0u);
    }
    GSB_CUDA(cudaGetLastError());
}

// Waits for the build and takes what the host needs from its control block: node count, level ranges, bounds, SAH cost.
void BuildJob::collect(DeviceBvh& out) {
    GSB_CUDA(cudaMemcpyAsync(&host, ctl, sizeof(BuildControl), cudaMemcpyDeviceToHost, s));
    GSB_CUDA(cudaStreamSynchronize(s));
    out.levelEnd.clear();
    numNodes = 1;
    if (n <= 1) {
        out.levelEnd.assign(1, 1u);
        out.sahBefore = out.sahAfter = 0.0;
        return;
    }
    if (host.level >= (uint32_t)kMaxLevels || host.levelCount[host.level] != 0u)
        throw std::runtime_error("BVH collapse: tree deeper than 64 levels of 8-wide nodes");
    // node index ranges of the levels: one node per item, levels emitted one after the other
    uint32_t levelEnd = 0;
    for (int l = 0; l < kMaxLevels && host.levelCount[l] > 0u; l++) {
        levelEnd += host.levelCount[l];
        out.levelEnd.push_back(levelEnd);
    }
    numNodes = host.counters[0];
    if (levelEnd != numNodes) throw std::runtime_error("BVH collapse: level counts do not add up to the node count");
    out.sahBefore = host.cost[0];
    out.sahAfter = refine ? host.cost[1] : host.cost[0];
    if (getenv("GSB_BUILD_TRACE") && n <= 100000)
        fprintf(stderr, "[gsb build] %u primitives: %u nodes on %zu levels, binary SAH %.1f -> %.1f, wide SAH (cut tables) %.2f\n", n, numNodes,
                out.levelEnd.size(), out.sahBefore, out.sahAfter, (double)host.wideCost);
}

struct GenericResult {
    uint4* nodes;
    uint32_t numNodes;
    uint32_t* leafPrim;
};

// The final buffer: nodes, leaf records and, behind them, the leaf order (kept for the refit) in ONE block from the
// stream-ordered pool (two cudaMalloc calls per mesh cost 0.6 ms of host time each on the field scene, with outliers of
// tens of ms when the driver went to the OS); freeBvh returns it with cudaFree, which waits for the device as before.
void finishBuffer(const GenericResult& r, uint32_t n, size_t leafBytes, cudaStream_t s, DeviceBvh& out) {
    const size_t nodeBytes = (size_t)r.numNodes * 80;
    out.numNodes = r.numNodes;
    out.numPrims = n;
    out.leafOffset16 = (uint32_t)(nodeBytes / 16);
    out.bytes = nodeBytes + (size_t)n * leafBytes;
    const size_t orderOffset = (out.bytes + 255) & ~(size_t)255;
    void* buf = nullptr;
    GSB_CUDA(cudaMallocAsync(&buf, std::max<size_t>(orderOffset + (size_t)n * sizeof(uint32_t), 16), s));
    out.buffer = (uint4*)buf;
    GSB_CUDA(cudaMemcpyAsync(out.buffer, r.nodes, nodeBytes, cudaMemcpyDeviceToDevice, s));
    out.leafPrim = nullptr;
    if (n > 0) {
        out.leafPrim = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(buf) + orderOffset);
        GSB_CUDA(cudaMemcpyAsync(out.leafPrim, r.leafPrim, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
    }
}

}  // namespace

void setTreeletRefinement(bool enabled) { g_treelets = enabled; }

void freeBvh(DeviceBvh& bvh) {
    if (bvh.buffer) cudaFree(bvh.buffer);   // leafPrim lives in the same block
    bvh = DeviceBvh();
}

// Bottom levels of several meshes at once.  Every build is a chain of dependent launches whose tail (the last levels of the
// bottom-up treelet and box passes, the first levels of the collapse) leaves most of the GPU idle, so the builds of a batch
// run on up to kBuildStreams streams side by side and the host waits once per phase, not per mesh:
//   phase 1  every job queued on its stream: bounds, boxes, Morton codes, sort, tree, treelets, collapse
//   phase 2  control blocks read back (node counts), final buffers allocated, nodes copied, triangles packed
// Jobs are taken in waves of at most kWaveTriangles so that the temporaries (~230 B per triangle) stay bounded.
void buildBlasBatch(const BlasBuildRequest* requests, size_t count, cudaStream_t s) {
    if (count == 0) return;
    // candidate treelet settings {passes, gamma} of small levels, beside the default build
    static const int kCandidates[5][2] = {{1, 7}, {1, 10}, {1, 24}, {2, 14}, {2, 7}};
    constexpr uint32_t kCandidateMaxTris = 16384;
    bool candidates = true;
    if (const char* e = getenv("GSB_BUILD_CANDIDATES")) candidates = atoi(e) != 0;
    if (getenv("GSB_TREELET_PASSES") || getenv("GSB_TREELET_GAMMA") || getenv("GSB_COLLAPSE_SAH") || !g_treelets) candidates = false;   // explicit settings are taken as given
    size_t kBuildStreams = 8;
    if (const char* e = getenv("GSB_BUILD_STREAMS")) kBuildStreams = (size_t)std::max(1, atoi(e));
    constexpr uint64_t kWaveTriangles = 48ull << 20;
    cudaEvent_t e0, e1;
    GSB_CUDA(cudaEventCreate(&e0));
    GSB_CUDA(cudaEventCreate(&e1));
    GSB_CUDA(cudaEventRecord(e0, s));
    std::vector<cudaStream_t> streams;
    cudaEvent_t joined = nullptr;
    bool anySmall = false;
    for (size_t i = 0; i < count; i++)
        anySmall = anySmall || (requests[i].ntris >= 16 && requests[i].ntris <= kCandidateMaxTris && !(requests[i].setting && *requests[i].setting >= 0));
    if (count > 1 || (candidates && anySmall)) {
        streams.resize(candidates && anySmall ? kBuildStreams : std::min(count, kBuildStreams));
        for (cudaStream_t& st : streams) {
            GSB_CUDA(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
            GSB_CUDA(cudaStreamWaitEvent(st, e0, 0));
        }
        GSB_CUDA(cudaEventCreateWithFlags(&joined, cudaEventDisableTiming));
    }
    const int B = 256;
    const bool trace = getenv("GSB_BUILD_TRACE") != nullptr;
    uint64_t totalTris = 0;
    for (size_t i = 0; i < count; i++) totalTris += requests[i].ntris;
    size_t first = 0;
    while (first < count) {
        size_t end = first;
        uint64_t waveTris = 0;
        while (end < count && (end == first || waveTris + requests[end].ntris <= kWaveTriangles)) waveTris += requests[end++].ntris;
        std::vector<std::unique_ptr<BuildJob>> jobs;
        std::vector<std::vector<std::unique_ptr<BuildJob>>> rivals(end - first);   // candidate builds of small levels
        std::vector<int> winner(end - first, 0);
        const auto tWave = std::chrono::steady_clock::now();
        for (size_t i = first; i < end; i++) {
            const BlasBuildRequest& q = requests[i];
            size_t turn = i - first;
            auto launch = [&](int passes, int gamma) {
                cudaStream_t js = streams.empty() ? s : streams[turn++ % streams.size()];   // candidates of a level side by side
                std::unique_ptr<BuildJob> job(new BuildJob(js, q.ntris));
                BuildJob& j = *job;
                j.passesOverride = passes;
                j.gammaOverride = gamma;
                j.begin(q.ntris, false);
                float4* plo = j.pool.get<float4>(q.ntris);
                float4* phi = j.pool.get<float4>(q.ntris);
                if (q.ntris > 0) {
                    BoundsAcc* acc = j.pool.get<BoundsAcc>(1);
                    k_initBounds<<<1, 32, 0, js>>>(acc);
                    const uint32_t nverts = q.ntris * 3;
                    k_vertexBounds<<<(int)std::min<uint32_t>((nverts + B - 1) / B, 296u), B, 0, js>>>(q.vertices, nverts, acc);
                    k_blasFrame<<<1, 1, 0, js>>>(acc, &j.ctl->frame);
                    k_triBoxes<<<(q.ntris + B - 1) / B, B, 0, js>>>(q.vertices, q.ntris, &j.ctl->frame, plo, phi);
                }
                j.start(plo, phi);
                return job;
            };
            const int fixed = (q.setting && *q.setting > 0 && *q.setting <= 5) ? *q.setting : 0;
            jobs.push_back(fixed ? launch(kCandidates[fixed - 1][0], kCandidates[fixed - 1][1]) : launch(-1, 0));
            winner[i - first] = fixed;
            // A small level is built several times, with different treelet settings, and the tree with the least SAH cost
            // AS AN 8-WIDE TREE (the root entry of the cut tables) is kept: on a few dozen triangles which treelets happen
            // to be formed decides whether the level has three nodes or four, and the binary SAH cost does not see it
            // (Cornell box, r3c: 386.3 against 386.8, but 144.7 against 149.1 as wide trees, 1096 against 1003 Mpaths/s).
            if (candidates && q.ntris >= 16 && q.ntris <= kCandidateMaxTris && !(q.setting && *q.setting >= 0))
                for (const int* c : {kCandidates[0], kCandidates[1], kCandidates[2], kCandidates[3], kCandidates[4]})
                    rivals[i - first].push_back(launch(c[0], c[1]));
        }
        const auto tIssued = std::chrono::steady_clock::now();
        double msCollect = 0.0, msFinish = 0.0;
        for (size_t i = first; i < end; i++) {
            const BlasBuildRequest& q = requests[i];
            DeviceBvh& out = *q.out;
            const auto tc0 = std::chrono::steady_clock::now();
            jobs[i - first]->collect(out);
            for (size_t c = 0; c < rivals[i - first].size(); c++) {
                std::unique_ptr<BuildJob>& r = rivals[i - first][c];
                DeviceBvh other;
                r->collect(other);
                if (r->host.wideCost > 0.0f && jobs[i - first]->host.wideCost > 0.0f && r->host.wideCost < jobs[i - first]->host.wideCost) {
                    out.levelEnd = other.levelEnd;
                    out.sahBefore = other.sahBefore;
                    out.sahAfter = other.sahAfter;
                    jobs[i - first].swap(r);
                    winner[i - first] = (int)c + 1;
                }
            }
            if (q.setting && (!rivals[i - first].empty() || *q.setting < 0)) *q.setting = winner[i - first];
            rivals[i - first].clear();
            BuildJob& j = *jobs[i - first];
            const auto tc1 = std::chrono::steady_clock::now();
            msCollect += std::chrono::duration<double, std::milli>(tc1 - tc0).count();
            const GenericResult r = {j.nodes, j.numNodes, j.leafPrim};
            finishBuffer(r, q.ntris, 144, j.s, out);
            if (q.ntris > 0)
                k_packTriangles<<<(q.ntris + B - 1) / B, B, 0, j.s>>>(j.leafPrim, q.ntris, q.vertices,
                                                                       reinterpret_cast<float4*>(out.buffer + out.leafOffset16), q.idsInVertices);
            GSB_CUDA(cudaGetLastError());
            for (int a = 0; a < 3; a++) {
                out.lo[a] = q.ntris ? j.host.frame.lo[a] : 0.0f;
                out.hi[a] = q.ntris ? j.host.frame.hi[a] : 0.0f;
            }
            msFinish += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tc1).count();
        }
        if (trace)
            fprintf(stderr, "[gsb build] wave of %zu meshes, %llu triangles: queued in %.2f ms, waited %.2f ms for the control blocks, %.2f ms in allocation + pack launches\n",
                    end - first, (unsigned long long)waveTris, std::chrono::duration<double, std::milli>(tIssued - tWave).count(), msCollect, msFinish);
        jobs.clear();   // the temporaries go back to the pool in stream order
        first = end;
    }
    for (cudaStream_t st : streams) {
        GSB_CUDA(cudaEventRecord(joined, st));
        GSB_CUDA(cudaStreamWaitEvent(s, joined, 0));
    }
    GSB_CUDA(cudaEventRecord(e1, s));
    GSB_CUDA(cudaEventSynchronize(e1));
    float ms = 0.0f;
    GSB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    // the batch time is shared out by triangle count, so that the per-mesh figures still add up to the time that passed
    for (size_t i = 0; i < count; i++)
        requests[i].out->buildMs = count == 1 ? ms : (totalTris ? ms * (float)((double)requests[i].ntris / (double)totalTris) : ms / (float)count);
    for (cudaStream_t st : streams) cudaStreamDestroy(st);
    if (joined) cudaEventDestroy(joined);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

void buildBlas(const float4* dVertices, uint32_t ntris, cudaStream_t s, DeviceBvh& out, bool idsInVertices, int* setting) {
    const BlasBuildRequest q = {dVertices, ntris, &out, idsInVertices, setting};
    buildBlasBatch(&q, 1, s);
}

// ---- refit ---------------------------------------------------------------------------------------------------------
namespace {

// leaf boxes of a triangle level from its own packed records (copy kz = 2 holds x, y, z in order), padded like k_triBoxes
__global__ void k_leafTriBoxes(const float4* __restrict__ leaves, uint32_t n, float pad, float4* __restrict__ lo, float4* __restrict__ hi) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const float4 a = leaves[9 * (size_t)k + 6], b = leaves[9 * (size_t)k + 7], c = leaves[9 * (size_t)k + 8];
    lo[k] = make_float4(fminf(a.x, fminf(b.x, c.x)) - pad, fminf(a.y, fminf(b.y, c.y)) - pad, fminf(a.z, fminf(b.z, c.z)) - pad, 0.0f);
    hi[k] = make_float4(fmaxf(a.x, fmaxf(b.x, c.x)) + pad, fmaxf(a.y, fmaxf(b.y, c.y)) + pad, fmaxf(a.z, fmaxf(b.z, c.z)) + pad, 0.0f);
}

__global__ void k_gatherBoxes(const uint32_t* __restrict__ leafPrim, uint32_t n, const float4* __restrict__ inLo, const float4* __restrict__ inHi,
                              float4* __restrict__ lo, float4* __restrict__ hi) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    lo[k] = inLo[leafPrim[k]];
    hi[k] = inHi[leafPrim[k]];
}

// One level of the bottom-up refit: nodes [first, end).  The topology (child / leaf ranges, slot assignment, meta bytes)
// stays; every node gets the union of its children's new boxes, a new quantisation frame and new quantised child boxes
// (same outward rounding as emitNode).  nodeLo / nodeHi: per-node boxes, written here, read by the parents' launch.
__global__ void k_refitLevel(uint4* __restrict__ nodes, uint32_t first, uint32_t end, const float4* __restrict__ leafLo,
                             const float4* __restrict__ leafHi, float4* __restrict__ nodeLo, float4* __restrict__ nodeHi, double* area, bool dry) {
    const uint32_t node = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= end) return;
    uint4* w = nodes + (size_t)node * 5;
    const uint4 w0 = w[0], w1 = w[1];
    const uint32_t imask = w0.w >> 24;
    float clo[8][3], chi[8][3];
    bool used[8];
    float nlo[3] = {INFINITY, INFINITY, INFINITY}, nhi[3] = {-INFINITY, -INFINITY, -INFINITY};
    for (int s = 0; s < 8; s++) {
        const uint32_t m = ((s < 4 ? w1.z : w1.w) >> (8 * (s & 3))) & 0xffu;
        used[s] = m != 0u;
        if (!used[s]) continue;
        float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        if ((m & 0x1fu) >= 24u) {
            const uint32_t c = w1.x + (uint32_t)__popc(imask & ((1u << s) - 1u));
            const float4 a = nodeLo[c], b = nodeHi[c];
            lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = b.x; hi[1] = b.y; hi[2] = b.z;
        } else {
            const uint32_t cnt = (uint32_t)__popc(m >> 5), off = m & 0x1fu;
            for (uint32_t j = 0; j < cnt; j++) {
                const float4 a = leafLo[w1.y + off + j], b = leafHi[w1.y + off + j];
                lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
                hi[0] = fmaxf(hi[0], b.x); hi[1] = fmaxf(hi[1], b.y); hi[2] = fmaxf(hi[2], b.z);
            }
        }
        for (int ax = 0; ax < 3; ax++) {
            clo[s][ax] = lo[ax]; chi[s][ax] = hi[ax];
            nlo[ax] = fminf(nlo[ax], lo[ax]); nhi[ax] = fmaxf(nhi[ax], hi[ax]);
        }
    }
    if (!(nhi[0] >= nlo[0])) { nlo[0] = nlo[1] = nlo[2] = 0.0f; nhi[0] = nhi[1] = nhi[2] = 0.0f; }   // an empty root
    uint32_t e[3];
    double scale[3], inv[3];
    for (int ax = 0; ax < 3; ax++) e[ax] = (uint32_t)(quantExponent(nhi[ax] - nlo[ax], scale[ax], inv[ax]) + 127);
    uint32_t qlo[3][8], qhi[3][8];
    for (int s = 0; s < 8; s++)
        for (int ax = 0; ax < 3; ax++) {
            qlo[ax][s] = 0; qhi[ax][s] = 0;
            if (!used[s]) continue;
            double ql = floor(((double)clo[s][ax] - (double)nlo[ax]) * inv[ax]);
            double qh = ceil(((double)chi[s][ax] - (double)nlo[ax]) * inv[ax]);
            qlo[ax][s] = (uint32_t)fmin(fmax(ql, 0.0), 255.0);
            qhi[ax][s] = (uint32_t)fmin(fmax(qh, 0.0), 255.0);
        }
    auto pack4 = [](const uint32_t* b) { return b[0] | (b[1] << 8) | (b[2] << 16) | (b[3] << 24); };
    if (!dry) {
        w[0] = make_uint4(__float_as_uint(nlo[0]), __float_as_uint(nlo[1]), __float_as_uint(nlo[2]), e[0] | (e[1] << 8) | (e[2] << 16) | (imask << 24));
        w[2] = make_uint4(pack4(qlo[0]), pack4(qlo[0] + 4), pack4(qlo[1]), pack4(qlo[1] + 4));
        w[3] = make_uint4(pack4(qlo[2]), pack4(qlo[2] + 4), pack4(qhi[0]), pack4(qhi[0] + 4));
        w[4] = make_uint4(pack4(qhi[1]), pack4(qhi[1] + 4), pack4(qhi[2]), pack4(qhi[2] + 4));
    }
    // what the parents (and the area figure) see is the box the traversal will really test: the dequantised union
    float dlo[3], dhi[3];
    for (int ax = 0; ax < 3; ax++) {
        uint32_t mx = 0;
        for (int s = 0; s < 8; s++) mx = max(mx, used[s] ? qhi[ax][s] : 0u);
        dlo[ax] = nlo[ax];
        dhi[ax] = (float)((double)nlo[ax] + (double)mx * scale[ax]);
        dhi[ax] = fmaxf(dhi[ax], nhi[ax]);
    }
    nodeLo[node] = make_float4(dlo[0], dlo[1], dlo[2], 0.0f);
    nodeHi[node] = make_float4(dhi[0], dhi[1], dhi[2], 0.0f);
    const double dx = (double)dhi[0] - dlo[0], dy = (double)dhi[1] - dlo[1], dz = (double)dhi[2] - dlo[2];
    atomicAdd(area, 2.0 * (dx * dy + dy * dz + dz * dx));
}

double refitLevels(DeviceBvh& bvh, const float4* leafLo, const float4* leafHi, TempPool& pool, cudaStream_t s, bool dry) {
    float4* nodeLo = pool.get<float4>(bvh.numNodes);
    float4* nodeHi = pool.get<float4>(bvh.numNodes);
    double* dArea = pool.get<double>(1);
    GSB_CUDA(cudaMemsetAsync(dArea, 0, sizeof(double), s));
    for (int l = (int)bvh.levelEnd.size() - 1; l >= 0; l--) {
        const uint32_t first = l ? bvh.levelEnd[l - 1] : 0u, end = bvh.levelEnd[l];
        if (end > first) k_refitLevel<<<(end - first + 127) / 128, 128, 0, s>>>(bvh.buffer, first, end, leafLo, leafHi, nodeLo, nodeHi, dArea, dry);
    }
    double area = 0.0;
    float4 rootLo, rootHi;
    GSB_CUDA(cudaMemcpyAsync(&area, dArea, sizeof(double), cudaMemcpyDeviceToHost, s));
    GSB_CUDA(cudaMemcpyAsync(&rootLo, nodeLo, sizeof(float4), cudaMemcpyDeviceToHost, s));
    GSB_CUDA(cudaMemcpyAsync(&rootHi, nodeHi, sizeof(float4), cudaMemcpyDeviceToHost, s));
    GSB_CUDA(cudaStreamSynchronize(s));
    GSB_CUDA(cudaGetLastError());
    if (!dry) {
        bvh.lo[0] = rootLo.x; bvh.lo[1] = rootLo.y; bvh.lo[2] = rootLo.z;
        bvh.hi[0] = rootHi.x; bvh.hi[1] = rootHi.y; bvh.hi[2] = rootHi.z;
    }
    return area;
}

struct ScopedTimer {
    cudaEvent_t e0, e1;
    cudaStream_t s;
    float* out;
    ScopedTimer(cudaStream_t st, float* o) : s(st), out(o) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, s); }
    ~ScopedTimer() { cudaEventRecord(e1, s); cudaEventSynchronize(e1); cudaEventElapsedTime(out, e0, e1); cudaEventDestroy(e0); cudaEventDestroy(e1); }
};

}  // namespace

float soupScale(const float4* dVertices, uint32_t nverts, cudaStream_t s) {
    if (nverts == 0) return 0.0f;
    TempPool pool(s);
    BoundsAcc* acc = pool.get<BoundsAcc>(1);
    k_initBounds<<<1, 32, 0, s>>>(acc);
    k_vertexBounds<<<(int)std::min<uint32_t>((nverts + 255) / 256, 2048u), 256, 0, s>>>(dVertices, nverts, acc);
    BoundsAcc h;
    GSB_CUDA(cudaMemcpyAsync(&h, acc, sizeof(h), cudaMemcpyDeviceToHost, s));
    GSB_CUDA(cudaStreamSynchronize(s));
    float scale = 0.0f;
    for (int a = 0; a < 3; a++) scale = std::fmax(scale, std::fmax(std::fabs(floatFromOrdered(h.lo[a])), std::fabs(floatFromOrdered(h.hi[a]))));
    return scale;
}

void repackTriangles(DeviceBvh& bvh, const float4* dVertices, bool idsInVertices, cudaStream_t s) {
    if (bvh.numPrims == 0) return;
    k_packTriangles<<<(bvh.numPrims + 255) / 256, 256, 0, s>>>(bvh.leafPrim, bvh.numPrims, dVertices,
                                                               reinterpret_cast<float4*>(bvh.buffer + bvh.leafOffset16), idsInVertices);
}

double refitTriangles(DeviceBvh& bvh, float worldScale, cudaStream_t s, bool dry) {
    if (bvh.numPrims == 0) return 0.0;
    float scratch = 0.0f;
    ScopedTimer timer(s, dry ? &scratch : &bvh.refitMs);
    TempPool pool(s);
    const uint32_t n = bvh.numPrims;
    float4* lo = pool.get<float4>(n);
    float4* hi = pool.get<float4>(n);
    k_leafTriBoxes<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const float4*>(bvh.buffer + bvh.leafOffset16), n, worldScale * kBoxPadRel, lo, hi);
    return refitLevels(bvh, lo, hi, pool, s, dry);
}

double refitTlas(DeviceBvh& bvh, const float* hostLo, const float* hostHi, const TraceInstance* hostRecords, uint32_t n, cudaStream_t s, bool dry) {
    if (n == 0 || n != bvh.numPrims) return 0.0;
    float scratch = 0.0f;
    ScopedTimer timer(s, dry ? &scratch : &bvh.refitMs);
    TempPool pool(s);
    std::vector<float4> lo4(n), hi4(n);
    for (uint32_t i = 0; i < n; i++) {
        lo4[i] = make_float4(hostLo[3 * i], hostLo[3 * i + 1], hostLo[3 * i + 2], 0.0f);
        hi4[i] = make_float4(hostHi[3 * i], hostHi[3 * i + 1], hostHi[3 * i + 2], 0.0f);
    }
    float4* inLo = pool.get<float4>(n);
    float4* inHi = pool.get<float4>(n);
    float4* lo = pool.get<float4>(n);
    float4* hi = pool.get<float4>(n);
    uint4* rec = pool.get<uint4>((size_t)n * 4);
    GSB_CUDA(cudaMemcpyAsync(inLo, lo4.data(), (size_t)n * sizeof(float4), cudaMemcpyHostToDevice, s));
    GSB_CUDA(cudaMemcpyAsync(inHi, hi4.data(), (size_t)n * sizeof(float4), cudaMemcpyHostToDevice, s));
    GSB_CUDA(cudaMemcpyAsync(rec, hostRecords, (size_t)n * sizeof(TraceInstance), cudaMemcpyHostToDevice, s));
    if (!dry) k_packInstances<<<(n + 255) / 256, 256, 0, s>>>(bvh.leafPrim, n, rec, bvh.buffer + bvh.leafOffset16);
    k_gatherBoxes<<<(n + 255) / 256, 256, 0, s>>>(bvh.leafPrim, n, inLo, inHi, lo, hi);
    return refitLevels(bvh, lo, hi, pool, s, dry);   // synchronises: the host vectors above outlive their copies
}

void buildTlas(const float* hostLo, const float* hostHi, const TraceInstance* hostRecords, uint32_t n, cudaStream_t s, DeviceBvh& out) {
    cudaEvent_t e0, e1;
    GSB_CUDA(cudaEventCreate(&e0));
    GSB_CUDA(cudaEventCreate(&e1));
    GSB_CUDA(cudaEventRecord(e0, s));
    {
        TempPool pool(s);
        std::vector<float4> lo4(n), hi4(n);
        float clo[3] = {INFINITY, INFINITY, INFINITY}, chi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (uint32_t i = 0; i < n; i++) {
            lo4[i] = make_float4(hostLo[3 * i], hostLo[3 * i + 1], hostLo[3 * i + 2], 0.0f);
            hi4[i] = make_float4(hostHi[3 * i], hostHi[3 * i + 1], hostHi[3 * i + 2], 0.0f);
            for (int a = 0; a < 3; a++) {
                const float c = 0.5f * (hostLo[3 * i + a] + hostHi[3 * i + a]);
                clo[a] = std::fmin(clo[a], c);
                chi[a] = std::fmax(chi[a], c);
                out.lo[a] = i == 0 ? hostLo[a] : std::fmin(out.lo[a], hostLo[3 * i + a]);
                out.hi[a] = i == 0 ? hostHi[a] : std::fmax(out.hi[a], hostHi[3 * i + a]);
            }
        }
        float4* plo = pool.get<float4>(n);
        float4* phi = pool.get<float4>(n);
        uint4* rec = pool.get<uint4>((size_t)n * 4);
        if (n > 0) {
            GSB_CUDA(cudaMemcpyAsync(plo, lo4.data(), (size_t)n * sizeof(float4), cudaMemcpyHostToDevice, s));
            GSB_CUDA(cudaMemcpyAsync(phi, hi4.data(), (size_t)n * sizeof(float4), cudaMemcpyHostToDevice, s));
            GSB_CUDA(cudaMemcpyAsync(rec, hostRecords, (size_t)n * sizeof(TraceInstance), cudaMemcpyHostToDevice, s));
        }
        BuildJob job(s, n);
        job.begin(n, true);
        BuildFrame frame;
        memset(&frame, 0, sizeof(frame));
        for (int a = 0; a < 3; a++) {
            frame.cmin[a] = clo[a];
            frame.cinv[a] = chi[a] > clo[a] ? 1.0f / (chi[a] - clo[a]) : 0.0f;
        }
        GSB_CUDA(cudaMemcpyAsync(&job.ctl->frame, &frame, sizeof(frame), cudaMemcpyHostToDevice, s));
        job.start(plo, phi);
        job.collect(out);
        const GenericResult r = {job.nodes, job.numNodes, job.leafPrim};
        finishBuffer(r, n, sizeof(TraceInstance), s, out);
        if (n > 0) k_packInstances<<<(n + 255) / 256, 256, 0, s>>>(r.leafPrim, n, rec, out.buffer + out.leafOffset16);
        GSB_CUDA(cudaGetLastError());
        GSB_CUDA(cudaStreamSynchronize(s));  // lo4/hi4 host vectors must outlive the copies
    }
    GSB_CUDA(cudaEventRecord(e1, s));
    GSB_CUDA(cudaEventSynchronize(e1));
    GSB_CUDA(cudaEventElapsedTime(&out.buildMs, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

}  // namespace gsb

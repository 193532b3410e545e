// treelet.cuh — SAH refinement of the binary LBVH by treelet restructuring (Karras & Aila 2013,
// "Fast parallel construction of high-quality bounding volume hierarchies", section 3).
// (included by bvh_build.cu after the BinTree definition)
//
// One bottom-up pass: every lane starts at a leaf and climbs; the second lane to arrive at a node owns
// it (its two subtrees are final).  An owned node whose subtree holds >= gamma leaves becomes a treelet
// root: the warp grows a treelet of 7 leaves under it (always expanding the treelet leaf of largest area),
// finds the topology over those 7 leaves with the least summed internal-node area by dynamic programming
// over the 127 leaf subsets (the subtrees below the treelet leaves and the root box do not change, so
// that sum is the whole SAH difference), and rewrites the 6 internal nodes in place when that is
// cheaper.  Nothing waits on another warp, so there is no forward-progress requirement.
//
// Restructuring breaks the "subtree = contiguous range of the sorted order" property that the collapse
// relies on (first/last, leaf groups copied as sorted[first..last]); relineariseTree() restores it by
// re-deriving the leaf order from the final topology.
#pragma once

namespace gsb {
namespace {

constexpr int kTreeletWarps = 4;
constexpr unsigned kFullMask = 0xffffffffu;

struct TreeletScratch {   // per warp
    float area[128];      // surface area of the union of the leaves in subset s
    float copt[128];      // least summed internal-node area of a tree over subset s
    uint8_t part[128];    // the left side of the best partition of s
    float4 lo[7], hi[7];  // treelet leaves
    int leafRef[7];
    uint32_t leafCnt[7];
    int nodeId[6];        // internal nodes available for reuse; [0] = treelet root
    int newLeft[6], newRight[6];
    uint8_t subsetOf[6];
};

// slo / shi: the primitive boxes in sorted order (k_sortedBoxes), so that the box of a leaf reference is one load away
__device__ inline void treeletRefBox(const BinTree& t, const float4* __restrict__ slo, const float4* __restrict__ shi, int ref, float4& lo,
                                     float4& hi) {
    if (ref >= 0) {
        lo = __ldcg(t.lo + ref);
        hi = __ldcg(t.hi + ref);
    } else {
        lo = __ldg(slo + ~ref);
        hi = __ldg(shi + ~ref);
    }
}

__global__ void k_sortedBoxes(const float4* __restrict__ plo, const float4* __restrict__ phi, const uint32_t* __restrict__ sorted, int n,
                              float4* __restrict__ slo, float4* __restrict__ shi) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t p = sorted[i];
    slo[i] = plo[p];
    shi[i] = phi[p];
}

__device__ void optimiseTreelet(const BinTree& t, uint32_t* cnt, const float4* __restrict__ slo, const float4* __restrict__ shi, int root,
                                TreeletScratch& S, int lane, const uint8_t* __restrict__ subsetsBySize, uint32_t* stats) {
    // ---- 1. grow the treelet: lanes 0..6 each hold one treelet leaf -------------------------------------
    // A lane that holds an internal node also fetches that node's two child references at once (myL / myR): when the node
    // is the one expanded next, its children's boxes are one load away instead of two dependent ones.  The whole pass is
    // bound by these chains of dependent L2 loads, not by arithmetic.
    int myRef = 0, myL = 0, myR = 0;
    float myArea = -2.0f;  // -2: unused lane, -1: primitive leaf (cannot be expanded)
    float4 mylo = make_float4(0, 0, 0, 0), myhi = mylo;
    if (lane < 2) {
        myRef = lane == 0 ? __ldcg(t.left + root) : __ldcg(t.right + root);
        if (myRef >= 0) {
            myL = __ldcg(t.left + myRef);
            myR = __ldcg(t.right + myRef);
        }
        treeletRefBox(t, slo, shi, myRef, mylo, myhi);
        myArea = myRef >= 0 ? boxArea(mylo, myhi) : -1.0f;
    }
    if (lane == 0) S.nodeId[0] = root;
    float curCost = 0.0f;
    for (int nLeaves = 2; nLeaves < 7; nLeaves++) {
        float best = myArea;
        int bl = lane;
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) {  // lanes 0..7 hold the candidates
            const float oa = __shfl_xor_sync(kFullMask, best, o);
            const int ol = __shfl_xor_sync(kFullMask, bl, o);
            if (oa > best || (oa == best && ol < bl)) { best = oa; bl = ol; }
        }
        best = __shfl_sync(kFullMask, best, 0);
        bl = __shfl_sync(kFullMask, bl, 0);
        if (best < 0.0f) return;  // fewer than 7 leaves below the root (cannot happen while cnt >= 7); warp-uniform
        const int x = __shfl_sync(kFullMask, myRef, bl);
        const int xl = __shfl_sync(kFullMask, myL, bl), xr = __shfl_sync(kFullMask, myR, bl);
        curCost += best;
        if (lane == 0) S.nodeId[nLeaves - 1] = x;
        if (lane == bl || lane == nLeaves) {
            myRef = lane == bl ? xl : xr;
            if (myRef >= 0) {
                myL = __ldcg(t.left + myRef);
                myR = __ldcg(t.right + myRef);
            }
            treeletRefBox(t, slo, shi, myRef, mylo, myhi);
            myArea = myRef >= 0 ? boxArea(mylo, myhi) : -1.0f;
        }
    }
    if (lane < 7) {
        S.lo[lane] = mylo;
        S.hi[lane] = myhi;
        S.leafRef[lane] = myRef;
        S.leafCnt[lane] = myRef >= 0 ? __ldcg(cnt + myRef) : 1u;
    }
    __syncwarp();

    // ---- 2. area of every subset, cost of the singletons ------------------------------------------------------
    for (int s = lane; s < 128; s += 32) {
        float a = 0.0f;
        if (s) {
            float4 lo = make_float4(INFINITY, INFINITY, INFINITY, 0), hi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0);
#pragma unroll
            for (int k = 0; k < 7; k++)
                if (s >> k & 1) {
                    const float4 l = S.lo[k], h = S.hi[k];
                    lo.x = fminf(lo.x, l.x); lo.y = fminf(lo.y, l.y); lo.z = fminf(lo.z, l.z);
                    hi.x = fmaxf(hi.x, h.x); hi.y = fmaxf(hi.y, h.y); hi.z = fmaxf(hi.z, h.z);
                }
            a = boxArea(lo, hi);
        }
        S.area[s] = a;
        S.copt[s] = 0.0f;
    }
    __syncwarp();

    // ---- 3. dynamic programming over subsets of 2..6 leaves --------------------------------------------------
    const int base[8] = {0, 0, 7, 28, 63, 98, 119, 126};
    for (int k = 2; k <= 6; k++) {
        const int count = base[k + 1] - base[k];
        for (int idx = lane; idx < count; idx += 32) {
            const int s = subsetsBySize[base[k] + idx];
            float bestC = INFINITY;
            int bestP = 0;
            const int delta = (s - 1) & s;
            int p = (-delta) & s;
            do {
                const float c = S.copt[p] + S.copt[s ^ p];
                if (c < bestC) { bestC = c; bestP = p; }
                p = (p - delta) & s;
            } while (p != 0);
            S.copt[s] = S.area[s] + bestC;
            S.part[s] = (uint8_t)bestP;
        }
        __syncwarp();
    }
    // the full set: 63 partitions spread over the warp (the side without leaf 0 is 2j, j = 1..63)
    float bestC = INFINITY;
    int bestP = 0;
    for (int j = lane + 1; j < 64; j += 32) {
        const int p = j << 1;
        const float c = S.copt[p] + S.copt[127 ^ p];
        if (c < bestC) { bestC = c; bestP = p; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float oc = __shfl_xor_sync(kFullMask, bestC, o);
        const int op = __shfl_xor_sync(kFullMask, bestP, o);
        if (oc < bestC || (oc == bestC && op < bestP)) { bestC = oc; bestP = op; }
    }
    if (lane == 0 && stats) atomicAdd(&stats[0], 1u);
    if (!(bestC < curCost * 0.9999f)) return;  // warp-uniform: nothing cheaper than the present topology
    if (lane == 0 && stats) atomicAdd(&stats[1], 1u);

    // ---- 4. rebuild the six internal nodes from the partition table ------------------------------------
    if (lane == 0) {
        S.subsetOf[0] = 127;
        int na = 1;
        for (int j = 0; j < 6; j++) {
            const int s = S.subsetOf[j];
            const int p = j == 0 ? bestP : (int)S.part[s];
            const int side[2] = {p, s ^ p};
            int ref[2];
            for (int c = 0; c < 2; c++) {
                if (__popc(side[c]) == 1) {
                    ref[c] = S.leafRef[__ffs(side[c]) - 1];
                } else {
                    S.subsetOf[na] = (uint8_t)side[c];
                    ref[c] = S.nodeId[na];
                    na++;
                }
            }
            S.newLeft[j] = ref[0];
            S.newRight[j] = ref[1];
        }
    }
    __syncwarp();
    if (lane < 6) {
        const int nd = S.nodeId[lane];
        const int s = S.subsetOf[lane];
        const int l = S.newLeft[lane], r = S.newRight[lane];
        __stcg(t.left + nd, l);
        __stcg(t.right + nd, r);
        if (l >= 0) __stcg(t.parent + l, nd); else __stcg(t.leafParent + ~l, nd);
        if (r >= 0) __stcg(t.parent + r, nd); else __stcg(t.leafParent + ~r, nd);
        if (lane > 0) {  // the root keeps its box and leaf count
            float4 lo = make_float4(INFINITY, INFINITY, INFINITY, 0), hi = make_float4(-INFINITY, -INFINITY, -INFINITY, 0);
            uint32_t c = 0;
#pragma unroll
            for (int k = 0; k < 7; k++)
                if (s >> k & 1) {
                    const float4 a = S.lo[k], b = S.hi[k];
                    lo.x = fminf(lo.x, a.x); lo.y = fminf(lo.y, a.y); lo.z = fminf(lo.z, a.z);
                    hi.x = fmaxf(hi.x, b.x); hi.y = fmaxf(hi.y, b.y); hi.z = fmaxf(hi.z, b.z);
                    c += S.leafCnt[k];
                }
            __stcg(t.lo + nd, lo);
            __stcg(t.hi + nd, hi);
            __stcg(cnt + nd, c);
        }
        __threadfence();
    }
    __syncwarp();
}

__global__ void __launch_bounds__(32 * kTreeletWarps, 16)
k_treeletPass(BinTree t, uint32_t* cnt, const float4* __restrict__ slo, const float4* __restrict__ shi, int n, uint32_t gamma,
              uint32_t* stats) {
    __shared__ TreeletScratch scratch[kTreeletWarps];
    __shared__ uint8_t subsetsBySize[128];  // the 127 non-empty subsets ordered by leaf count
    for (int s = threadIdx.x + 1; s < 128; s += blockDim.x) {
        const int k = __popc(s);
        int rank = 0;
        for (int q = 1; q < s; q++) rank += __popc(q) == k;
        const int base[8] = {0, 0, 7, 28, 63, 98, 119, 126};
        subsetsBySize[base[k] + rank] = (uint8_t)s;
    }
    __syncthreads();
    const int lane = threadIdx.x & 31;
    TreeletScratch& S = scratch[threadIdx.x >> 5];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    int node = i < n ? __ldcg(t.leafParent + i) : -1;
    bool active = node >= 0;
    while (__any_sync(kFullMask, active)) {
        bool own = false;
        if (active) {
            __threadfence();
            own = atomicAdd(&t.flags[node], 1u) != 0u;  // the first arrival stops; the second finds both subtrees final
            active = own;
        }
        unsigned m = __ballot_sync(kFullMask, own && __ldcg(cnt + node) >= gamma);
        while (m) {
            const int src = __ffs(m) - 1;
            m &= m - 1;
            optimiseTreelet(t, cnt, slo, shi, __shfl_sync(kFullMask, node, src), S, lane, subsetsBySize, stats);
        }
        if (active) {
            node = __ldcg(t.parent + node);
            active = node >= 0;
        }
    }
}

__global__ void k_initCounts(BinTree t, uint32_t* cnt, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n - 1) cnt[i] = t.last[i] - t.first[i] + 1u;
}

// number of leaves left of `ref` in a depth-first walk of the whole tree
__device__ inline uint32_t dfsPosition(const BinTree& t, const uint32_t* __restrict__ cnt, int ref, int node) {
    uint32_t pos = 0;
    while (node >= 0) {
        const int l = t.left[node];
        if (l != ref) pos += l >= 0 ? cnt[l] : 1u;
        ref = node;
        node = t.parent[node];
    }
    return pos;
}

__global__ void k_newLeafOrder(BinTree t, const uint32_t* __restrict__ cnt, const uint32_t* __restrict__ sorted, int n,
                               uint32_t* __restrict__ newSorted, uint32_t* __restrict__ newPos) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t pos = dfsPosition(t, cnt, ~i, t.leafParent[i]);
    newSorted[pos] = sorted[i];
    newPos[i] = pos;
}

__global__ void k_newRanges(BinTree t, const uint32_t* __restrict__ cnt, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const uint32_t f = dfsPosition(t, cnt, i, t.parent[i]);
    t.first[i] = f;
    t.last[i] = f + cnt[i] - 1u;
}

__global__ void k_remapLeafRefs(BinTree t, const uint32_t* __restrict__ newPos, const int* __restrict__ oldLeafParent, int n,
                                int* __restrict__ newLeafParent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n - 1) {
        const int l = t.left[i], r = t.right[i];
        if (l < 0) t.left[i] = ~(int)newPos[~l];
        if (r < 0) t.right[i] = ~(int)newPos[~r];
    }
    if (i < n) newLeafParent[newPos[i]] = oldLeafParent[i];
}

// `passes` bottom-up treelet passes, then the leaf order is re-derived so that every subtree again covers a
// contiguous range of `sorted` (which is replaced by the new order).  stats (host, optional): treelets
// examined / rewritten over all passes.
template <class Alloc>
void refineTreelets(cudaStream_t s, BinTree& t, const float4* plo, const float4* phi, uint32_t*& sorted, int n, int passes,
                    Alloc alloc, uint32_t* hostStats = nullptr, int gammaOverride = 0) {
    const int B = 256;
    uint32_t* cnt = (uint32_t*)alloc((size_t)n * sizeof(uint32_t));
    uint32_t* stats = (uint32_t*)alloc(2 * sizeof(uint32_t));
    GSB_CUDA(cudaMemsetAsync(stats, 0, 2 * sizeof(uint32_t), s));
    k_initCounts<<<(n + B - 1) / B, B, 0, s>>>(t, cnt, n);
    float4* slo = (float4*)alloc((size_t)n * sizeof(float4));
    float4* shi = (float4*)alloc((size_t)n * sizeof(float4));
    k_sortedBoxes<<<(n + B - 1) / B, B, 0, s>>>(plo, phi, sorted, n, slo, shi);
    const int TB = 32 * kTreeletWarps;
    // gamma: a node becomes a treelet root when its subtree holds at least that many leaves (7 = every node a treelet can
    // be grown under; larger values skip the small subtrees, which the treelets of their ancestors still reach into)
    // Default 14: on the field scene the pass costs 2.4 ms instead of 4 and the SAH cost ends 0.4 % higher (66 982 against
    // 66 726; paths/s equal within noise on field and gallery, profiles/README.md r3c).
    uint32_t gamma = 14u;
    if (const char* e = getenv("GSB_TREELET_GAMMA")) gamma = (uint32_t)std::max(7, atoi(e));
    if (gammaOverride >= 7) gamma = (uint32_t)gammaOverride;   // a candidate build of a small level (bvh_build.cu: kCandidates)
    for (int p = 0; p < passes; p++) {
        GSB_CUDA(cudaMemsetAsync(t.flags, 0, (size_t)(n - 1) * sizeof(uint32_t), s));
        k_treeletPass<<<(n + TB - 1) / TB, TB, 0, s>>>(t, cnt, slo, shi, n, gamma, stats);
    }
    uint32_t* newSorted = (uint32_t*)alloc((size_t)n * sizeof(uint32_t));
    uint32_t* newPos = (uint32_t*)alloc((size_t)n * sizeof(uint32_t));
    int* newLeafParent = (int*)alloc((size_t)n * sizeof(int));
    k_newLeafOrder<<<(n + B - 1) / B, B, 0, s>>>(t, cnt, sorted, n, newSorted, newPos);
    k_newRanges<<<(n + B - 1) / B, B, 0, s>>>(t, cnt, n);
    k_remapLeafRefs<<<(n + B - 1) / B, B, 0, s>>>(t, newPos, t.leafParent, n, newLeafParent);
    GSB_CUDA(cudaGetLastError());
    t.leafParent = newLeafParent;
    sorted = newSorted;
    if (hostStats) {
        GSB_CUDA(cudaMemcpyAsync(hostStats, stats, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        GSB_CUDA(cudaStreamSynchronize(s));
    }
}

}  // namespace
}  // namespace gsb

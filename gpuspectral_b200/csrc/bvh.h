// bvh.h — GPU BVH builder interface (internal to the CUDA core).
//
// Stands where the reference calls the Vulkan driver's acceleration-structure builds:
//   VulkanRays::buildBLAS  backend/vulkan/VulkanRays.cpp:6-86    (one per unique mesh)
//   VulkanRays::buildTLAS  backend/vulkan/VulkanRays.cpp:91-181  (over {blas, mat4} instances)
// B200 has no RT cores, so both levels are built here in CUDA: Morton-code LBVH
// (Karras 2012) -> SAH treelet refinement -> collapse to 8-wide quantised nodes (traverse.cuh).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <vector>

#include "common.cuh"

namespace gsb {

struct DeviceBvh {
    uint4* buffer = nullptr;   // [numNodes x 80 B nodes | leaf records]; root = node 0
    size_t bytes = 0;
    uint32_t numNodes = 0;
    uint32_t numPrims = 0;
    uint32_t leafOffset16 = 0; // offset of the leaf records, in 16-byte units
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0};  // padded root bounds
    double sahBefore = 0.0, sahAfter = 0.0;      // binary-tree SAH cost before/after treelet refinement
    float buildMs = 0.0f;
    // kept for refitting (animated instances): the input primitive of every leaf position, the node index range of every
    // level of the 8-wide tree (levels are emitted one after the other, children always behind their parents), and the
    // summed surface area of the node boxes right after the build (the refit's quality yardstick)
    uint32_t* leafPrim = nullptr;
    std::vector<uint32_t> levelEnd;
    double areaAtBuild = 0.0;
    float refitMs = 0.0f;
};

// Padding of every leaf box relative to the coordinate scale of the mesh (shared with the
// checker): keeps the box tests conservative with respect to the float32 triangle test.
constexpr float kBoxPadRel = 1.0f / 32768.0f;

// Bottom level over a triangle soup (3 float4 per triangle, object space).  Leaf record =
// 3 copies (one per dominant ray axis, pre-rotated) of 3 x float4 (v0.xyz | primId bits, v1.xyz, v2.xyz) = 144 B.
// idsInVertices: the leaf records take v0.w / v1.w from the input vertices (world level: primitive and instance id of
// every baked triangle) instead of the position of the triangle in the soup.
// out.leafPrim keeps the input triangle index of every leaf position.
void buildBlas(const float4* dVertices, uint32_t numTriangles, cudaStream_t stream, DeviceBvh& out, bool idsInVertices = false, int* setting = nullptr);
// The bottom levels of several meshes in one call: the builds run side by side on a few streams and the host waits once
// per phase instead of several times per mesh.  buildMs of every `out` = its share (by triangle count) of the batch time.
// setting (optional, in / out): which treelet setting to build a small level with; -1 = try them all (see kCandidates in
// bvh_build.cu) and report the one that won, so that a level that is rebuilt again and again (the world level, at every
// gsb_scene_set_instances) pays for the comparison once.
struct BlasBuildRequest { const float4* vertices; uint32_t ntris; DeviceBvh* out; bool idsInVertices; int* setting = nullptr; };
void buildBlasBatch(const BlasBuildRequest* requests, size_t count, cudaStream_t stream);

// Bake rule (shared with the checker): in gl_InstanceID order an instance whose mesh has at most
// kBakeMaxTris triangles is stored in world space at the world level, while the baked total stays within kBakeBudget.
constexpr uint32_t kBakeMaxTris = 16384;
constexpr uint64_t kBakeBudget = 1ull << 22;

// Top level over instances.  hostLo/hostHi: padded world bounds (3 floats each) per instance;
// hostRecords: TraceInstance per instance in gl_InstanceID order.  Leaf record = TraceInstance.
void buildTlas(const float* hostLo, const float* hostHi, const TraceInstance* hostRecords, uint32_t numInstances,
               cudaStream_t stream, DeviceBvh& out);

void freeBvh(DeviceBvh& bvh);

// Refit (VulkanRays.cpp:92-118 rebuilds its TLAS every frame; here moved primitives keep the tree's topology): the leaf
// records have already been rewritten in place; every node's quantisation frame and child boxes are recomputed bottom-up
// from the new leaf boxes (conservatively: child boxes are the dequantised ones).  Returns the summed surface area of the
// node boxes — compare with DeviceBvh::areaAtBuild to decide when a rebuild is due.
// Bottom level / world level: boxes from the packed triangle records of the buffer itself; worldScale = largest absolute coordinate.
// dry = true only measures the area of the tree as it is (nothing is written).
double refitTriangles(DeviceBvh& bvh, float worldScale, cudaStream_t stream, bool dry = false);
// Top level: hostLo / hostHi = padded world bounds per instance in the order given to buildTlas, hostRecords likewise.
double refitTlas(DeviceBvh& bvh, const float* hostLo, const float* hostHi, const TraceInstance* hostRecords, uint32_t numInstances,
                 cudaStream_t stream, bool dry = false);
// re-packs the leaf records of a triangle level from its vertex soup (after the soup changed), in the kept leaf order
void repackTriangles(DeviceBvh& bvh, const float4* dVertices, bool idsInVertices, cudaStream_t stream);
// largest absolute coordinate of a vertex array (the scale the box padding is relative to)
float soupScale(const float4* dVertices, uint32_t numVertices, cudaStream_t stream);

// Treelet refinement switch (on by default); exposed for A/B measurements.
void setTreeletRefinement(bool enabled);

}  // namespace gsb

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
};

// Padding of every leaf box relative to the coordinate scale of the mesh (shared with the
// checker): keeps the box tests conservative with respect to the float32 triangle test.
constexpr float kBoxPadRel = 1.0f / 32768.0f;

// Bottom level over a triangle soup (3 float4 per triangle, object space).  Leaf record =
// 3 copies (one per dominant ray axis, pre-rotated) of 3 x float4 (v0.xyz | primId bits, v1.xyz, v2.xyz) = 144 B.
// idsInVertices: the leaf records take v0.w / v1.w from the input vertices (world level: primitive and instance id of
// every baked triangle) instead of the position of the triangle in the soup.
// outLeafPrim (optional): receives a cudaMalloc'ed array with the input triangle index of every leaf position (the caller frees it).
void buildBlas(const float4* dVertices, uint32_t numTriangles, cudaStream_t stream, DeviceBvh& out, bool idsInVertices = false,
               uint32_t** outLeafPrim = nullptr);

// Bake rule (shared with the checker): in gl_InstanceID order an instance whose mesh has at most
// kBakeMaxTris triangles is stored in world space at the world level, while the baked total stays within kBakeBudget.
constexpr uint32_t kBakeMaxTris = 16384;
constexpr uint64_t kBakeBudget = 1ull << 22;

// Top level over instances.  hostLo/hostHi: padded world bounds (3 floats each) per instance;
// hostRecords: TraceInstance per instance in gl_InstanceID order.  Leaf record = TraceInstance.
void buildTlas(const float* hostLo, const float* hostHi, const TraceInstance* hostRecords, uint32_t numInstances,
               cudaStream_t stream, DeviceBvh& out);

void freeBvh(DeviceBvh& bvh);

// Treelet refinement switch (on by default); exposed for A/B measurements.
void setTreeletRefinement(bool enabled);

}  // namespace gsb

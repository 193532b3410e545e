// treelet.cuh — SAH refinement of the binary LBVH by treelet restructuring.
// (included by bvh_build.cu after the BinTree definition)
#pragma once

namespace gsb {
namespace {

// Placeholder until the restructuring kernel lands: leaves the tree untouched.
template <class Alloc>
void refineTreelets(cudaStream_t, BinTree&, const float4*, const float4*, const uint32_t*, int, Alloc) {}

}  // namespace
}  // namespace gsb

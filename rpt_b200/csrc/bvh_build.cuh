// GPU LBVH build (Morton sort + Karras 2012 hierarchy + bottom-up refit), replacing the
// reference's serial median-split kd-tree construction (src/kdtree.rs:238-348).  Results need
// not match the kd-tree, only the hits do.  Output: BVH2 nodes of 64 B with both children's
// padded f32 boxes (4 x float4 loads per visit), leaves of <= kLeafMax primitives that are
// contiguous in Morton order.
#pragma once
#include "common.cuh"

namespace rptx {

struct PrimBox {
    float mn[3], mx[3];
};

struct LbvhScratch;  // opaque

// d_boxes: n primitive boxes (device).  d_nodes: room for max(n-1,1) nodes (device).
// d_perm: n entries (device), perm[k] = original index of the k-th primitive in leaf order.
// pad: absolute outward padding baked into every child box.  leaf_max: primitives per leaf (<= 8).  Returns cudaError_t as int.
int build_lbvh(const PrimBox* d_boxes, uint32_t n, float pad, int leaf_max, const double cmin[3], const double cmax[3],
               BvhNode* d_nodes, uint32_t* d_perm, int* root_ref, cudaStream_t stream, uint64_t* launches);

// Same hierarchy over items already sorted by 63-bit Morton keys (d_keys[k], d_boxes[k] = item in
// slot k; equal keys allowed).  d_nodes: room for max(n-1,1) nodes.
int build_lbvh_sorted(const uint64_t* d_keys, const PrimBox* d_boxes, uint32_t n, float pad, int leaf_max, BvhNode* d_nodes, int* root_ref,
                      cudaStream_t stream, uint64_t* launches);

// SAH-quality tree on the GPU: PLOC (parallel locally-ordered clustering over the Morton order, surface-area merge criterion),
// same contract and node layout as build_lbvh.  Replaces the reference's serial SAH-less median split (src/kdtree.rs:238-348)
// for object meshes and the instance level.  cudaErrorNotSupported = deeper than max_depth (build the LBVH instead).
int build_ploc(const PrimBox* d_boxes, uint32_t n, float pad, int leaf_max, const double cmin[3], const double cmax[3], BvhNode* d_nodes,
               uint32_t* d_perm, int* root_ref, int max_depth, cudaStream_t stream, uint64_t* launches);

}  // namespace rptx

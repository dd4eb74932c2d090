// GPU LBVH build — see bvh_build.cuh.
#include <cub/cub.cuh>

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "bvh_build.cuh"
#include "host_types.hpp"

namespace rptx {

namespace {

__device__ __forceinline__ uint32_t expand_bits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void k_morton(const PrimBox* __restrict__ boxes, uint32_t n, float3 cmin, float3 cinv, uint64_t* __restrict__ keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PrimBox b = boxes[i];
    const float cx = (0.5f * (b.mn[0] + b.mx[0]) - cmin.x) * cinv.x;
    const float cy = (0.5f * (b.mn[1] + b.mx[1]) - cmin.y) * cinv.y;
    const float cz = (0.5f * (b.mn[2] + b.mx[2]) - cmin.z) * cinv.z;
    const uint32_t x = (uint32_t)fminf(fmaxf(cx * 1024.0f, 0.0f), 1023.0f);
    const uint32_t y = (uint32_t)fminf(fmaxf(cy * 1024.0f, 0.0f), 1023.0f);
    const uint32_t z = (uint32_t)fminf(fmaxf(cz * 1024.0f, 0.0f), 1023.0f);
    const uint32_t code = (expand_bits10(x) << 2) | (expand_bits10(y) << 1) | expand_bits10(z);
    keys[i] = ((uint64_t)code << 32) | (uint64_t)i;  // index makes every key unique
}

// SORTED = false: keys are (30-bit Morton << 32 | primitive index), unique, and a primitive's box is
// boxes[key & 0xffffffff].  SORTED = true: keys are 63-bit Morton codes already sorted together with
// their boxes (boxes[k] belongs to position k); equal keys are split by position (Karras 2012, sec. 4).
template <bool SORTED>
__device__ __forceinline__ int delta(const uint64_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t x = keys[i] ^ keys[j];
    if (SORTED && x == 0) return 64 + __clz(i ^ j);
    return __clzll((long long)x);
}
template <bool SORTED>
__device__ __forceinline__ uint32_t box_of(const uint64_t* __restrict__ keys, int k) {
    return SORTED ? (uint32_t)k : (uint32_t)(keys[k] & 0xffffffffu);
}

// Karras 2012: one thread per internal node
template <bool SORTED>
__global__ void k_hierarchy(const uint64_t* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right,
                            int* __restrict__ first, int* __restrict__ last, int* __restrict__ parent_i,
                            int* __restrict__ parent_l) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta<SORTED>(keys, n, i, i + 1) - delta<SORTED>(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta<SORTED>(keys, n, i, i - d);
    int lmax = 2;
    while (delta<SORTED>(keys, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (delta<SORTED>(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta<SORTED>(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) / 2;
        if (delta<SORTED>(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    // children encoded: >= 0 internal, < 0 leaf ~index
    const int lc = (lo == gamma) ? ~gamma : gamma;
    const int rc = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    left[i] = lc; right[i] = rc; first[i] = lo; last[i] = hi;
    if (lc >= 0) parent_i[lc] = i; else parent_l[~lc] = i;
    if (rc >= 0) parent_i[rc] = i; else parent_l[~rc] = i;
    if (i == 0) parent_i[0] = -1;
}

__device__ __forceinline__ PrimBox merge(const PrimBox& a, const PrimBox& b) {
    PrimBox r;
#pragma unroll
    for (int k = 0; k < 3; k++) { r.mn[k] = fminf(a.mn[k], b.mn[k]); r.mx[k] = fmaxf(a.mx[k], b.mx[k]); }
    return r;
}

// L2-coherent read of a box written by another thread of the same launch
__device__ __forceinline__ PrimBox load_cg(const PrimBox* p) {
    PrimBox r;
#pragma unroll
    for (int k = 0; k < 3; k++) { r.mn[k] = __ldcg(&p->mn[k]); r.mx[k] = __ldcg(&p->mx[k]); }
    return r;
}

// bottom-up refit: one thread per leaf, second arrival at a node merges its children
template <bool SORTED>
__global__ void k_refit(const PrimBox* __restrict__ boxes, const uint64_t* __restrict__ keys, int n, const int* __restrict__ left,
                        const int* __restrict__ right, const int* __restrict__ parent_i, const int* __restrict__ parent_l,
                        int* __restrict__ flags, PrimBox* __restrict__ ibox) {
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int p = parent_l[k];
    while (p >= 0) {
        __threadfence();
        if (atomicAdd(&flags[p], 1) == 0) return;
        __threadfence();
        const int lc = left[p], rc = right[p];
        const PrimBox a = lc >= 0 ? load_cg(&ibox[lc]) : boxes[box_of<SORTED>(keys, ~lc)];
        const PrimBox b = rc >= 0 ? load_cg(&ibox[rc]) : boxes[box_of<SORTED>(keys, ~rc)];
        ibox[p] = merge(a, b);
        p = parent_i[p];
    }
}

template <bool SORTED>
__device__ __forceinline__ void child_out(int c, const PrimBox* __restrict__ boxes, const uint64_t* __restrict__ keys,
                                          const PrimBox* __restrict__ ibox, const int* __restrict__ first,
                                          const int* __restrict__ last, float pad, int leaf_max, float* mn, float* mx, int* ref) {
    PrimBox b;
    if (c >= 0) {
        b = ibox[c];
        const int sz = last[c] - first[c] + 1;
        *ref = (sz <= leaf_max) ? make_leaf_ref((uint32_t)first[c], (uint32_t)sz) : c;
    } else {
        b = boxes[box_of<SORTED>(keys, ~c)];
        *ref = make_leaf_ref((uint32_t)(~c), 1u);
    }
#pragma unroll
    for (int k = 0; k < 3; k++) { mn[k] = __fsub_rd(b.mn[k], pad); mx[k] = __fadd_ru(b.mx[k], pad); }
}

template <bool SORTED>
__global__ void k_emit(const PrimBox* __restrict__ boxes, const uint64_t* __restrict__ keys, int n, const int* __restrict__ left,
                       const int* __restrict__ right, const int* __restrict__ first, const int* __restrict__ last,
                       const PrimBox* __restrict__ ibox, float pad, int leaf_max, BvhNode* __restrict__ nodes, uint32_t* __restrict__ perm) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (!SORTED && i < n) perm[i] = (uint32_t)(keys[i] & 0xffffffffu);
    if (i >= n - 1) return;
    BvhNode nd;
    child_out<SORTED>(left[i], boxes, keys, ibox, first, last, pad, leaf_max, nd.lmin, nd.lmax, &nd.left);
    child_out<SORTED>(right[i], boxes, keys, ibox, first, last, pad, leaf_max, nd.rmin, nd.rmax, &nd.right);
    nd.pad0 = 0; nd.pad1 = 0;
    nodes[i] = nd;
}

}  // namespace

int build_lbvh(const PrimBox* d_boxes, uint32_t n, float pad, int leaf_max, const double cmin[3], const double cmax[3], BvhNode* d_nodes,
               uint32_t* d_perm, int* root_ref, cudaStream_t stream, uint64_t* launches) {
    if (n == 0) { *root_ref = 0; return 0; }
    cudaError_t e;
    uint64_t *keys_in = nullptr, *keys_out = nullptr;
    int* ints = nullptr;
    PrimBox* ibox = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0;
#define RPTX_CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    RPTX_CK(rptx_alloc_async((void**)&keys_in, sizeof(uint64_t) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&keys_out, sizeof(uint64_t) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&ints, sizeof(int) * (size_t)n * 7, stream));
    RPTX_CK(rptx_alloc_async((void**)&ibox, sizeof(PrimBox) * n, stream));
    {
        float3 fmin = make_float3((float)cmin[0], (float)cmin[1], (float)cmin[2]);
        float3 finv;
        finv.x = cmax[0] > cmin[0] ? (float)(1.0 / (cmax[0] - cmin[0])) : 0.f;
        finv.y = cmax[1] > cmin[1] ? (float)(1.0 / (cmax[1] - cmin[1])) : 0.f;
        finv.z = cmax[2] > cmin[2] ? (float)(1.0 / (cmax[2] - cmin[2])) : 0.f;
        const int T = 128;
        const int B = (int)((n + T - 1) / T);
        k_morton<<<B, T, 0, stream>>>(d_boxes, n, fmin, finv, keys_in);
        (*launches)++;
        RPTX_CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys_in, keys_out, (int)n, 0, 64, stream));
        RPTX_CK(rptx_alloc_async((void**)&tmp, tmp_bytes ? tmp_bytes : 16, stream));
        RPTX_CK(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, keys_in, keys_out, (int)n, 0, 64, stream));
        (*launches) += 1;
        int* left = ints;
        int* right = ints + n;
        int* first = ints + 2 * (size_t)n;
        int* last = ints + 3 * (size_t)n;
        int* parent_i = ints + 4 * (size_t)n;
        int* parent_l = ints + 5 * (size_t)n;
        int* flags = ints + 6 * (size_t)n;
        RPTX_CK(cudaMemsetAsync(flags, 0, sizeof(int) * n, stream));
        if (n > 1) {
            k_hierarchy<false><<<B, T, 0, stream>>>(keys_out, (int)n, left, right, first, last, parent_i, parent_l);
            k_refit<false><<<B, T, 0, stream>>>(d_boxes, keys_out, (int)n, left, right, parent_i, parent_l, flags, ibox);
            (*launches) += 2;
        }
        k_emit<false><<<B, T, 0, stream>>>(d_boxes, keys_out, (int)n, left, right, first, last, ibox, pad, leaf_max, d_nodes, d_perm);
        (*launches)++;
        RPTX_CK(cudaGetLastError());
        RPTX_CK(cudaStreamSynchronize(stream));
        *root_ref = (n <= (uint32_t)leaf_max) ? make_leaf_ref(0u, n) : 0;
    }
fail:
#undef RPTX_CK
    rptx_free_async(keys_in, stream); rptx_free_async(keys_out, stream); rptx_free_async(ints, stream); rptx_free_async(ibox, stream); rptx_free_async(tmp, stream);
    return (int)e;
}

// LBVH over items that are ALREADY sorted by a 63-bit Morton key (photon point maps).
int build_lbvh_sorted(const uint64_t* d_keys, const PrimBox* d_boxes, uint32_t n, float pad, int leaf_max, BvhNode* d_nodes, int* root_ref,
                      cudaStream_t stream, uint64_t* launches) {
    if (n == 0) { *root_ref = 0; return 0; }
    cudaError_t e;
    int* ints = nullptr;
    PrimBox* ibox = nullptr;
#define RPTX_CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    RPTX_CK(rptx_alloc_async((void**)&ints, sizeof(int) * (size_t)n * 7, stream));
    RPTX_CK(rptx_alloc_async((void**)&ibox, sizeof(PrimBox) * n, stream));
    {
        const int T = 128;
        const int B = (int)((n + T - 1) / T);
        int* left = ints;
        int* right = ints + n;
        int* first = ints + 2 * (size_t)n;
        int* last = ints + 3 * (size_t)n;
        int* parent_i = ints + 4 * (size_t)n;
        int* parent_l = ints + 5 * (size_t)n;
        int* flags = ints + 6 * (size_t)n;
        RPTX_CK(cudaMemsetAsync(flags, 0, sizeof(int) * n, stream));
        if (n > 1) {
            k_hierarchy<true><<<B, T, 0, stream>>>(d_keys, (int)n, left, right, first, last, parent_i, parent_l);
            k_refit<true><<<B, T, 0, stream>>>(d_boxes, d_keys, (int)n, left, right, parent_i, parent_l, flags, ibox);
            (*launches) += 2;
        }
        k_emit<true><<<B, T, 0, stream>>>(d_boxes, d_keys, (int)n, left, right, first, last, ibox, pad, leaf_max, d_nodes, nullptr);
        (*launches)++;
        RPTX_CK(cudaGetLastError());
        RPTX_CK(cudaStreamSynchronize(stream));
        *root_ref = (n <= (uint32_t)leaf_max) ? make_leaf_ref(0u, n) : 0;
    }
fail:
#undef RPTX_CK
    rptx_free_async(ints, stream); rptx_free_async(ibox, stream);
    return (int)e;
}

// ---- PLOC: SAH-quality trees on the GPU ----------------------------------------------------------------------------
// Parallel locally-ordered clustering (Meister & Bittner 2018) over the Morton order: every cluster looks kPlocRadius
// neighbours to each side for the partner whose union has the smallest surface area; mutual nearest neighbours merge into a
// new node; the survivors are compacted in order; repeat until one cluster is left.  The agglomeration follows the surface-area
// heuristic instead of the Morton bits, which is what the LBVH lacks (photon maps keep the LBVH: points have no area).
// The finished topology is laid out like the LBVH's: a node covers a contiguous range of the leaf order, subtrees of
// <= leaf_max primitives collapse into leaf references, node 0 is the root.
namespace {

constexpr int kPlocRadius = 64;  // neighbours searched to each side (RPTX_PLOC_RADIUS; measured sweep in DESIGN.md section 5)

struct PlocNode {
    PrimBox box;
    int left, right, parent;  // children: node ids (leaves are 0..n-1, inner nodes n..2n-2); parent -1 at the root
    int size;                 // primitives below
};

__device__ __forceinline__ float box_area(const PrimBox& b) {
    const float dx = b.mx[0] - b.mn[0], dy = b.mx[1] - b.mn[1], dz = b.mx[2] - b.mn[2];
    return dx * dy + dy * dz + dz * dx;
}

__global__ void k_iota(uint32_t* __restrict__ perm, uint32_t n) {
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) perm[i] = i;
}

__global__ void k_ploc_leaves(const PrimBox* __restrict__ boxes, const uint64_t* __restrict__ keys, uint32_t n, PlocNode* __restrict__ nodes,
                              int* __restrict__ cluster) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    PlocNode nd;
    nd.box = boxes[(uint32_t)(keys[i] & 0xffffffffu)];
    nd.left = nd.right = -1; nd.parent = -1; nd.size = 1;
    nodes[i] = nd;
    cluster[i] = (int)i;
}

__global__ void k_ploc_nearest(const PlocNode* __restrict__ nodes, const int* __restrict__ cluster, int m, int radius, int* __restrict__ nn) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const PrimBox a = nodes[cluster[i]].box;
    float best = CUDART_INF_F;
    int arg = -1;
    const int lo = max(0, i - radius), hi = min(m - 1, i + radius);
    // Ties on the cost are common (instances of one mesh on a regular pattern: equal boxes, equal gaps).  Broken by index
    // they form chains - everybody prefers its left neighbour and only the pair at the end is mutual: one merge per pass,
    // 800 passes for the 937 instances of fractal_teapots.  A symmetric hash of the pair breaks them instead: the pair with
    // the smallest hash among both partners' ties is mutual, about a third of the tied pairs per pass.
    unsigned best_h = 0xffffffffu;
    for (int j = lo; j <= hi; j++) {
        if (j == i) continue;
        const float c = box_area(merge(a, nodes[cluster[j]].box));
        const unsigned p = (unsigned)min(i, j), q = (unsigned)max(i, j);
        unsigned h = p * 0x9E3779B1u ^ (q + 0x7F4A7C15u) * 0x85EBCA77u;
        h ^= h >> 15; h *= 0x2C1B3C6Du; h ^= h >> 12;
        if (c < best || (c == best && h < best_h)) { best = c; best_h = h; arg = j; }
    }
    nn[i] = arg;
}

__global__ void k_ploc_merge(PlocNode* __restrict__ nodes, const int* __restrict__ cluster, const int* __restrict__ nn, int m, int* __restrict__ next_node,
                             int* __restrict__ out, uint8_t* __restrict__ keep) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int j = nn[i];
    int c = cluster[i];
    uint8_t k = 1;
    if (j >= 0 && nn[j] == i) {
        if (i < j) {
            const int cj = cluster[j];
            const int id = atomicAdd(next_node, 1);
            PlocNode nd;
            nd.box = merge(nodes[c].box, nodes[cj].box);
            nd.left = c; nd.right = cj; nd.parent = -1;
            nd.size = nodes[c].size + nodes[cj].size;
            nodes[id] = nd;
            nodes[c].parent = id;
            nodes[cj].parent = id;
            c = id;
        } else {
            k = 0;
        }
    }
    out[i] = c;
    keep[i] = k;
}

// position of a node's first primitive in the leaf order: the sizes of the left siblings on the way to the root
__global__ void k_ploc_first(const PlocNode* __restrict__ nodes, int total, int* __restrict__ first, int* __restrict__ max_depth) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= total) return;
    int pos = 0, depth = 0, c = v, p = nodes[v].parent;
    while (p >= 0) {
        if (nodes[p].right == c) pos += nodes[nodes[p].left].size;
        c = p;
        p = nodes[p].parent;
        depth++;
    }
    first[v] = pos;
    atomicMax(max_depth, depth);
}

__global__ void k_ploc_emit(const PlocNode* __restrict__ nodes, const int* __restrict__ first, const uint64_t* __restrict__ keys, int n, float pad,
                            int leaf_max, BvhNode* __restrict__ out, uint32_t* __restrict__ perm) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= 2 * n - 1) return;
    if (v < n) { perm[first[v]] = (uint32_t)(keys[v] & 0xffffffffu); return; }
    const PlocNode nd = nodes[v];
    BvhNode o;
    auto child = [&](int c, float* mn, float* mx, int* ref) {
        const PlocNode ch = nodes[c];
        *ref = ch.size <= leaf_max ? make_leaf_ref((uint32_t)first[c], (uint32_t)ch.size) : (2 * n - 2 - c);
#pragma unroll
        for (int k = 0; k < 3; k++) { mn[k] = __fsub_rd(ch.box.mn[k], pad); mx[k] = __fadd_ru(ch.box.mx[k], pad); }
    };
    child(nd.left, o.lmin, o.lmax, &o.left);
    child(nd.right, o.rmin, o.rmax, &o.right);
    o.pad0 = 0; o.pad1 = 0;
    out[2 * n - 2 - v] = o;  // the root (created last) is node 0
}

}  // namespace

// Same contract as build_lbvh.  Returns cudaErrorNotSupported when the tree came out deeper than max_depth (the caller then
// builds the LBVH, whose depth is bounded by its key length).
int build_ploc(const PrimBox* d_boxes, uint32_t n, float pad, int leaf_max, const double cmin[3], const double cmax[3], BvhNode* d_nodes,
               uint32_t* d_perm, int* root_ref, int max_depth, cudaStream_t stream, uint64_t* launches) {
    if (n == 0) { *root_ref = 0; return 0; }
    if (n <= (uint32_t)leaf_max) {  // the whole mesh is one leaf (a polygon of the Cornell box): no tree, identity order
        k_iota<<<1, 32, 0, stream>>>(d_perm, n);
        (*launches)++;
        *root_ref = make_leaf_ref(0u, n);
        return (int)cudaGetLastError();
    }
    cudaError_t e = cudaSuccess;
    uint64_t *keys_in = nullptr, *keys_out = nullptr;
    PlocNode* nodes = nullptr;
    int *cl_a = nullptr, *cl_b = nullptr, *cl_t = nullptr, *nn = nullptr, *first = nullptr, *ctr = nullptr;
    uint8_t* keep = nullptr;
    void* tmp = nullptr;
    size_t tmp_bytes = 0, sel_bytes = 0;
    int h_ctr[3] = {0, 0, 0};  // [0] next node id, [1] clusters after the compaction, [2] max depth
#define RPTX_CK(x) do { e = (x); if (e != cudaSuccess) goto fail; } while (0)
    RPTX_CK(rptx_alloc_async((void**)&keys_in, sizeof(uint64_t) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&keys_out, sizeof(uint64_t) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&nodes, sizeof(PlocNode) * (2 * (size_t)n), stream));
    RPTX_CK(rptx_alloc_async((void**)&cl_a, sizeof(int) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&cl_b, sizeof(int) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&cl_t, sizeof(int) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&nn, sizeof(int) * n, stream));
    RPTX_CK(rptx_alloc_async((void**)&first, sizeof(int) * (2 * (size_t)n), stream));
    RPTX_CK(rptx_alloc_async((void**)&keep, n, stream));
    RPTX_CK(rptx_alloc_async((void**)&ctr, sizeof(int) * 4, stream));
    {
        float3 fmin = make_float3((float)cmin[0], (float)cmin[1], (float)cmin[2]);
        float3 finv;
        finv.x = cmax[0] > cmin[0] ? (float)(1.0 / (cmax[0] - cmin[0])) : 0.f;
        finv.y = cmax[1] > cmin[1] ? (float)(1.0 / (cmax[1] - cmin[1])) : 0.f;
        finv.z = cmax[2] > cmin[2] ? (float)(1.0 / (cmax[2] - cmin[2])) : 0.f;
        const int T = 128;
        const int B = (int)((n + T - 1) / T);
        k_morton<<<B, T, 0, stream>>>(d_boxes, n, fmin, finv, keys_in);
        RPTX_CK(cub::DeviceRadixSort::SortKeys(nullptr, tmp_bytes, keys_in, keys_out, (int)n, 0, 64, stream));
        RPTX_CK(cub::DeviceSelect::Flagged(nullptr, sel_bytes, cl_t, keep, cl_b, ctr + 1, (int)n, stream));
        tmp_bytes = tmp_bytes > sel_bytes ? tmp_bytes : sel_bytes;
        RPTX_CK(rptx_alloc_async((void**)&tmp, tmp_bytes ? tmp_bytes : 16, stream));
        RPTX_CK(cub::DeviceRadixSort::SortKeys(tmp, tmp_bytes, keys_in, keys_out, (int)n, 0, 64, stream));
        k_ploc_leaves<<<B, T, 0, stream>>>(d_boxes, keys_out, n, nodes, cl_a);
        (*launches) += 3;
        h_ctr[0] = (int)n;
        RPTX_CK(cudaMemcpyAsync(ctr, h_ctr, sizeof(int) * 3, cudaMemcpyHostToDevice, stream));
        int m = (int)n, passes = 0;
        int radius = kPlocRadius;
        if (const char* ev = std::getenv("RPTX_PLOC_RADIUS")) radius = std::min(256, std::max(1, std::atoi(ev)));
        while (m > 1) {
            const int Bm = (m + T - 1) / T;
            k_ploc_nearest<<<Bm, T, 0, stream>>>(nodes, cl_a, m, radius, nn);
            k_ploc_merge<<<Bm, T, 0, stream>>>(nodes, cl_a, nn, m, ctr, cl_t, keep);
            size_t bytes = tmp_bytes;
            RPTX_CK(cub::DeviceSelect::Flagged(tmp, bytes, cl_t, keep, cl_b, ctr + 1, m, stream));
            (*launches) += 4;
            RPTX_CK(cudaMemcpyAsync(h_ctr + 1, ctr + 1, sizeof(int), cudaMemcpyDeviceToHost, stream));
            RPTX_CK(cudaStreamSynchronize(stream));
            if (h_ctr[1] >= m || h_ctr[1] < 1) { e = cudaErrorNotSupported; goto fail; }  // a pass without a merge (hash collision on a tie): the caller builds the LBVH
            passes++;
            m = h_ctr[1];
            int* t = cl_a; cl_a = cl_b; cl_b = t;
        }
        const int total = 2 * (int)n - 1;
        k_ploc_first<<<(total + T - 1) / T, T, 0, stream>>>(nodes, total, first, ctr + 2);
        k_ploc_emit<<<(total + T - 1) / T, T, 0, stream>>>(nodes, first, keys_out, (int)n, pad, leaf_max, d_nodes, d_perm);
        (*launches) += 2;
        RPTX_CK(cudaMemcpyAsync(h_ctr + 2, ctr + 2, sizeof(int), cudaMemcpyDeviceToHost, stream));
        RPTX_CK(cudaGetLastError());
        RPTX_CK(cudaStreamSynchronize(stream));
        *root_ref = (n <= (uint32_t)leaf_max) ? make_leaf_ref(0u, n) : 0;
        if (std::getenv("RPTX_TRACE_BUILD")) fprintf(stderr, "[rptx] PLOC: %u items, radius %d, %d passes, depth %d\n", n, radius, passes, h_ctr[2]);
        if (h_ctr[2] > max_depth) e = cudaErrorNotSupported;
    }
fail:
#undef RPTX_CK
    for (void* p : {(void*)keys_in, (void*)keys_out, (void*)nodes, (void*)cl_a, (void*)cl_b, (void*)cl_t, (void*)nn, (void*)first, (void*)keep, (void*)ctr, tmp})
        rptx_free_async(p, stream);
    return (int)e;
}

}  // namespace rptx

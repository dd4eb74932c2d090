// Host side of the photon integrators (C ABI of include/rptx.h, "photon" section): emission
// driver, photon-map builders and the camera pass.  Kernels: photon.cuh + the builders below.
#include <cub/cub.cuh>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include "../../include/rptx.h"
#include "bvh_build.cuh"
#include "host_types.hpp"
#include "photon.cuh"
#include "photon_host.hpp"

using namespace rptx;
using rptx_host::ctx_fail;

#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) return ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)


namespace {

struct PointMapHost {
    PointMapDev dev = {};
    std::vector<void*> owned;
};

struct BoxBvhHost {
    BoxBvhDev dev = {};
    std::vector<void*> owned;
};

int grid_for(size_t n, int block) { return (int)std::max<size_t>(1, (n + block - 1) / block); }

// developer knob RPTX_TRACE_BUILD: per-phase wall time of the photon passes on stderr (synchronises the stream)
struct PhaseTrace {
    rptx_ctx* ctx;
    bool on;
    double t0;
    static double now() {
        timespec ts;
        clock_gettime(CLOCK_MONOTONIC, &ts);
        return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
    }
    explicit PhaseTrace(rptx_ctx* c) : ctx(c), on(std::getenv("RPTX_TRACE_BUILD") != nullptr), t0(0) { if (on) { cudaStreamSynchronize(ctx->stream); t0 = now(); } }
    void mark(const char* what) {
        if (!on) return;
        cudaStreamSynchronize(ctx->stream);
        const double t = now();
        fprintf(stderr, "[rptx] %-28s %8.2f ms\n", what, t - t0);
        t0 = t;
    }
};

// Everything here comes from the stream-ordered pool on the library's stream: cudaMalloc / cudaFree
// next to a mapped multi-GB film slab or photon list stall for 20-460 ms at random (profiles/r1_v2_summary.md).
template <class T>
cudaError_t dev_alloc(rptx_ctx* ctx, T** p, size_t n, std::vector<void*>* owned) {
    *p = nullptr;
    cudaError_t e = rptx_alloc_async((void**)p, std::max<size_t>(n, 1) * sizeof(T), ctx->stream);
    if (e == cudaSuccess && owned) owned->push_back(*p);
    return e;
}

// ---- builders' kernels ----

__global__ void k_rows_keys(const PhotonRec* __restrict__ rows, uint32_t n, uint64_t* __restrict__ keys, uint32_t* __restrict__ idx) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    keys[i] = rows[i].key;
    idx[i] = i;
}

// per-block bounds of the photon positions (finalised on the host: <= 1184 x 6 doubles)
__global__ void __launch_bounds__(256) k_rows_bounds(const PhotonRec* __restrict__ rows, uint32_t n, double* __restrict__ out) {
    double mn[3] = {CUDART_INF, CUDART_INF, CUDART_INF}, mx[3] = {-CUDART_INF, -CUDART_INF, -CUDART_INF};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int k = 0; k < 3; k++) { mn[k] = fmin(mn[k], rows[i].pos[k]); mx[k] = fmax(mx[k], rows[i].pos[k]); }
    __shared__ double sh[6][8];
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_down_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_down_sync(0xffffffffu, mx[k], o));
        }
        if ((threadIdx.x & 31) == 0) { sh[k][threadIdx.x >> 5] = mn[k]; sh[3 + k][threadIdx.x >> 5] = mx[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = sh[threadIdx.x][0];
        for (int w = 1; w < 8; w++) v = threadIdx.x < 3 ? fmin(v, sh[threadIdx.x][w]) : fmax(v, sh[threadIdx.x][w]);
        out[6 * blockIdx.x + threadIdx.x] = v;
    }
}

__device__ __forceinline__ uint64_t expand_bits21(uint64_t v) {
    v &= 0x1fffffull;
    v = (v | v << 32) & 0x1f00000000ffffull;
    v = (v | v << 16) & 0x1f0000ff0000ffull;
    v = (v | v << 8) & 0x100f00f00f00f00full;
    v = (v | v << 4) & 0x10c30c30c30c30c3ull;
    v = (v | v << 2) & 0x1249249249249249ull;
    return v;
}

// 63-bit Morton key of row idx[i] (canonical order in, so that the stable sort breaks ties canonically)
__global__ void k_rows_morton(const PhotonRec* __restrict__ rows, const uint32_t* __restrict__ idx, uint32_t n, double3 mn, double3 inv,
                              uint64_t* __restrict__ keys) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PhotonRec& r = rows[idx[i]];
    const double s = 2097152.0;  // 2^21
    const uint64_t x = (uint64_t)fmin(fmax((r.pos[0] - mn.x) * inv.x * s, 0.0), s - 1.0);
    const uint64_t y = (uint64_t)fmin(fmax((r.pos[1] - mn.y) * inv.y * s, 0.0), s - 1.0);
    const uint64_t z = (uint64_t)fmin(fmax((r.pos[2] - mn.z) * inv.z * s, 0.0), s - 1.0);
    keys[i] = (expand_bits21(x) << 2) | (expand_bits21(y) << 1) | expand_bits21(z);
}

__global__ void k_rows_gather(const PhotonRec* __restrict__ rows, const uint32_t* __restrict__ idx, uint32_t n, double* __restrict__ px,
                              double* __restrict__ py, double* __restrict__ pz, double* __restrict__ dir, double* __restrict__ power,
                              uint64_t* __restrict__ key) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PhotonRec r = rows[idx[i]];
    px[i] = r.pos[0]; py[i] = r.pos[1]; pz[i] = r.pos[2];
    for (int k = 0; k < 3; k++) { dir[3 * (size_t)i + k] = r.dir[k]; power[3 * (size_t)i + k] = r.power[k]; }
    key[i] = r.key;
}

// f32 boxes (rounded outward) of the Morton-ordered photons, for the LBVH refit
__global__ void k_point_boxes(const double* __restrict__ px, const double* __restrict__ py, const double* __restrict__ pz, uint32_t n,
                              PrimBox* __restrict__ boxes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    PrimBox b;
    b.mn[0] = __double2float_rd(px[i]); b.mx[0] = __double2float_ru(px[i]);
    b.mn[1] = __double2float_rd(py[i]); b.mx[1] = __double2float_ru(py[i]);
    b.mn[2] = __double2float_rd(pz[i]); b.mx[2] = __double2float_ru(pz[i]);
    boxes[i] = b;
}

// radius of a photon sphere = distance to its 10th nearest volume photon, itself included
// (src/photon.rs:212-236), + its f32 box (:119-128)
__global__ void __launch_bounds__(128) k_sphere_radius(const __grid_constant__ PointMapDev vm, double* __restrict__ radius,
                                                       Aabb32* __restrict__ boxes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= vm.n) return;
    const d3 p = mk3(vm.px[i], vm.py[i], vm.pz[i]);
    KnnHeap<16> heap;
    heap.k = kSphereKnn;
    knn_query(vm, p, heap);
    double m = -1.0;
    for (uint32_t j = 0; j < heap.size; j++) m = fmax(m, heap.d2[j]);
    const double r = sqrt(m);
    radius[i] = r;
    Aabb32 b;
    b.mn[0] = (float)(p.x - r); b.mn[1] = (float)(p.y - r); b.mn[2] = (float)(p.z - r);
    b.mx[0] = (float)(p.x + r); b.mx[1] = (float)(p.y + r); b.mx[2] = (float)(p.z + r);
    boxes[i] = b;
}

__global__ void k_sphere_gather(const __grid_constant__ PointMapDev vm, const double* __restrict__ radius, const Aabb32* __restrict__ boxes,
                                const uint32_t* __restrict__ perm, uint32_t n, double* __restrict__ center, double* __restrict__ rad,
                                double* __restrict__ dir, double* __restrict__ power, Aabb32* __restrict__ obox) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = perm[i];
    center[3 * (size_t)i] = vm.px[s]; center[3 * (size_t)i + 1] = vm.py[s]; center[3 * (size_t)i + 2] = vm.pz[s];
    rad[i] = radius[s];
    for (int k = 0; k < 3; k++) { dir[3 * (size_t)i + k] = vm.dir[3 * (size_t)s + k]; power[3 * (size_t)i + k] = vm.power[3 * (size_t)s + k]; }
    obox[i] = boxes[s];
}

// PhotonBeam of a volume photon (src/photon.rs:264-288) + its f32 box (:76-105); canonical order
__global__ void k_beam_make(const PhotonRec* __restrict__ rows, const uint32_t* __restrict__ idx, uint32_t n, double* __restrict__ start,
                            double* __restrict__ dirn, double* __restrict__ len, double* __restrict__ power, Aabb32* __restrict__ boxes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PhotonRec r = rows[idx[i]];
    const d3 a = mk3(r.start[0], r.start[1], r.start[2]), b = mk3(r.pos[0], r.pos[1], r.pos[2]);
    const d3 dn = normalize(b - a);
    const double l = length(b - a);
    const double cx = (a.x - b.x) * (a.x - b.x), cy = (a.y - b.y) * (a.y - b.y), cz = (a.z - b.z) * (a.z - b.z);
    const double sq = cx + cy + cz;
    const double kx = sqrt((cy + cz) / sq), ky = sqrt((cx + cz) / sq), kz = sqrt((cx + cy) / sq);
    const d3 adj = mk3(kx * kBeamRadius, ky * kBeamRadius, kz * kBeamRadius);
    const d3 mn = mk3(fmin(a.x, b.x), fmin(a.y, b.y), fmin(a.z, b.z)) - adj;
    const d3 mx = mk3(fmax(a.x, b.x), fmax(a.y, b.y), fmax(a.z, b.z)) + adj;
    Aabb32 bx;
    bx.mn[0] = (float)mn.x; bx.mn[1] = (float)mn.y; bx.mn[2] = (float)mn.z;
    bx.mx[0] = (float)mx.x; bx.mx[1] = (float)mx.y; bx.mx[2] = (float)mx.z;
    boxes[i] = bx;
    start[3 * (size_t)i] = a.x; start[3 * (size_t)i + 1] = a.y; start[3 * (size_t)i + 2] = a.z;
    dirn[3 * (size_t)i] = dn.x; dirn[3 * (size_t)i + 1] = dn.y; dirn[3 * (size_t)i + 2] = dn.z;
    len[i] = l;
    for (int k = 0; k < 3; k++) power[3 * (size_t)i + k] = r.power[k];
}

__global__ void k_beam_gather(const double* __restrict__ start, const double* __restrict__ dirn, const double* __restrict__ len,
                              const double* __restrict__ power, const Aabb32* __restrict__ boxes, const uint32_t* __restrict__ perm, uint32_t n,
                              double* __restrict__ ostart, double* __restrict__ odirn, double* __restrict__ olen, double* __restrict__ opower,
                              Aabb32* __restrict__ obox) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = perm[i];
    for (int k = 0; k < 3; k++) {
        ostart[3 * (size_t)i + k] = start[3 * (size_t)s + k];
        odirn[3 * (size_t)i + k] = dirn[3 * (size_t)s + k];
        opower[3 * (size_t)i + k] = power[3 * (size_t)s + k];
    }
    olen[i] = len[s];
    obox[i] = boxes[s];
}

// ---- beam segments (BeamSegDev, photon.cuh) ----
__global__ void k_seg_count(const double* __restrict__ len, uint32_t n, double seg_len, uint32_t* __restrict__ count) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double l = len[i];
    uint32_t k = l > 0.0 && l < 1e300 ? (uint32_t)fmin(ceil(l / seg_len), (double)kBeamMaxSeg) : 1u;  // NaN / inf lengths: one piece
    count[i] = k < 1u ? 1u : k;
}

// piece k of K of beam i: the box of the points within the radius (plus a margin for the rounding of floor(beam_t K / len)) of
// the axis between len k / K and len (k + 1) / K, rounded outward to f32
__global__ void k_seg_make(const double* __restrict__ start, const double* __restrict__ dirn, const double* __restrict__ len,
                           const uint32_t* __restrict__ offset, const uint32_t* __restrict__ count, uint32_t n, Aabb32* __restrict__ boxes,
                           uint32_t* __restrict__ beam, uint16_t* __restrict__ kk, uint16_t* __restrict__ KK) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t K = count[i], o = offset[i];
    const d3 a = mk3(start[3 * (size_t)i], start[3 * (size_t)i + 1], start[3 * (size_t)i + 2]);
    const d3 d = mk3(dirn[3 * (size_t)i], dirn[3 * (size_t)i + 1], dirn[3 * (size_t)i + 2]);
    const double l = len[i];
    const double m = kBeamRadius * 1.0001 + 1e-9 * (fabs(a.x) + fabs(a.y) + fabs(a.z) + fabs(l));
    for (uint32_t k = 0; k < K; k++) {
        const d3 p = a + (l * (double)k / (double)K) * d, q = a + (l * (double)(k + 1) / (double)K) * d;
        Aabb32 b;
        b.mn[0] = __double2float_rd(fmin(p.x, q.x) - m); b.mx[0] = __double2float_ru(fmax(p.x, q.x) + m);
        b.mn[1] = __double2float_rd(fmin(p.y, q.y) - m); b.mx[1] = __double2float_ru(fmax(p.y, q.y) + m);
        b.mn[2] = __double2float_rd(fmin(p.z, q.z) - m); b.mx[2] = __double2float_ru(fmax(p.z, q.z) + m);
        if (K == 1u && !(l > 0.0 && l < 1e300)) {  // a degenerate beam keeps an unbounded piece: its own box decides
            for (int c = 0; c < 3; c++) { b.mn[c] = -3.0e38f; b.mx[c] = 3.0e38f; }
        }
        boxes[o + k] = b;
        beam[o + k] = i;
        kk[o + k] = (uint16_t)k;
        KK[o + k] = (uint16_t)K;
    }
}

__global__ void k_seg_gather(const Aabb32* __restrict__ boxes, const uint32_t* __restrict__ beam, const uint16_t* __restrict__ kk,
                             const uint16_t* __restrict__ KK, const uint32_t* __restrict__ perm, uint32_t n, Aabb32* __restrict__ obox,
                             uint32_t* __restrict__ obeam, uint16_t* __restrict__ okk, uint16_t* __restrict__ oKK) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t s = perm[i];
    obox[i] = boxes[s]; obeam[i] = beam[s]; okk[i] = kk[s]; oKK[i] = KK[s];
}

// per-block bounds of finite f32 boxes (Morton normalisation of the box BVH)
__global__ void __launch_bounds__(256) k_box_bounds(const Aabb32* __restrict__ boxes, uint32_t n, double* __restrict__ out) {
    double mn[3] = {CUDART_INF, CUDART_INF, CUDART_INF}, mx[3] = {-CUDART_INF, -CUDART_INF, -CUDART_INF};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        for (int k = 0; k < 3; k++) {
            const float a = boxes[i].mn[k], b = boxes[i].mx[k];
            if (isfinite(a)) mn[k] = fmin(mn[k], (double)a);
            if (isfinite(b)) mx[k] = fmax(mx[k], (double)b);
        }
    __shared__ double sh[6][8];
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_down_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_down_sync(0xffffffffu, mx[k], o));
        }
        if ((threadIdx.x & 31) == 0) { sh[k][threadIdx.x >> 5] = mn[k]; sh[3 + k][threadIdx.x >> 5] = mx[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        double v = sh[threadIdx.x][0];
        for (int w = 1; w < 8; w++) v = threadIdx.x < 3 ? fmin(v, sh[threadIdx.x][w]) : fmax(v, sh[threadIdx.x][w]);
        out[6 * blockIdx.x + threadIdx.x] = v;
    }
}

// batch_sum[pixel] = sum over lanes, fixed order (as k_batch_reduce of the path tracer)
__global__ void k_lane_reduce(const double* __restrict__ ax, const double* __restrict__ ay, const double* __restrict__ az, uint32_t n_pixels,
                              uint32_t n_lanes, double* __restrict__ batch_sum) {
    const uint32_t pixel = blockIdx.x * blockDim.x + threadIdx.x;
    if (pixel >= n_pixels) return;
    double x = 0, y = 0, z = 0;
    for (uint32_t l = 0; l < n_lanes; l++) {
        const size_t s = (size_t)l * n_pixels + pixel;
        x += ax[s]; y += ay[s]; z += az[s];
    }
    batch_sum[3 * (size_t)pixel] = x; batch_sum[3 * (size_t)pixel + 1] = y; batch_sum[3 * (size_t)pixel + 2] = z;
}

// kNN parity hook: one thread per query
template <int KMAX>
__global__ void __launch_bounds__(128) k_knn_batch(const __grid_constant__ PointMapDev m, const double* __restrict__ q, size_t n, uint32_t k,
                                                   uint64_t* __restrict__ keys, double* __restrict__ d2, uint32_t* __restrict__ counts) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    KnnHeap<KMAX> heap;
    heap.k = k;
    knn_query(m, mk3(q[3 * i], q[3 * i + 1], q[3 * i + 2]), heap);
    heap.sort_ascending();
    for (uint32_t j = 0; j < k; j++) {
        keys[i * k + j] = j < heap.size ? m.key[heap.idx[j]] : ~0ull;
        d2[i * k + j] = j < heap.size ? heap.d2[j] : CUDART_INF;
    }
    counts[i] = heap.size;
}

// ---- sharded build (one process per GPU; comm.cu) ----

// out[i] = rows[idx[i]]: this rank's rows in (Morton, canonical) order, ready to be cut into per-destination runs
__global__ void k_rows_permute(const PhotonRec* __restrict__ rows, const uint32_t* __restrict__ idx, uint32_t n, PhotonRec* __restrict__ out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = rows[idx[i]];
}

// lb[b] = number of (sorted) keys below b << shift, b = 0..bins (lb[bins] = n): summed over the ranks this is the global
// rank of every bin boundary, from which all ranks pick the same splitters
__global__ void k_bin_lower_bounds(const uint64_t* __restrict__ keys, uint32_t n, int shift, uint32_t bins, unsigned long long* __restrict__ lb) {
    const uint32_t b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b > bins) return;
    if (b == bins) { lb[b] = n; return; }
    const uint64_t bound = (uint64_t)b << shift;
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (keys[mid] < bound) lo = mid + 1; else hi = mid;
    }
    lb[b] = lo;
}

// cut[r] = this rank's first row that goes to rank r (cut[world] = n_local): rank r owns the bins [s_r, s_r+1), s_r the
// first bin whose global lower bound reaches r/world of all rows
__global__ void k_pick_splitters(const unsigned long long* __restrict__ lb_global, const unsigned long long* __restrict__ lb_local, uint32_t bins,
                                 int world, unsigned long long* __restrict__ cut) {
    const int r = threadIdx.x;
    if (r > world) return;
    if (r == 0) { cut[0] = 0; return; }
    if (r == world) { cut[r] = lb_local[bins]; return; }
    const unsigned long long total = lb_global[bins];
    const unsigned long long target = total * (unsigned long long)r / (unsigned long long)world;  // total < 2^32
    uint32_t lo = 0, hi = bins;
    while (lo < hi) {
        const uint32_t mid = lo + ((hi - lo) >> 1);
        if (lb_global[mid] < target) lo = mid + 1; else hi = mid;
    }
    cut[r] = lb_local[lo];
}

// k_sphere_radius over the photons [first, first + count) of the (complete) volume map: one rank's share of the radius pass
__global__ void __launch_bounds__(128) k_sphere_radius_range(const __grid_constant__ PointMapDev vm, uint32_t first, uint32_t count,
                                                             double* __restrict__ radius, Aabb32* __restrict__ boxes) {
    const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= count) return;
    const uint32_t i = first + j;
    const d3 p = mk3(vm.px[i], vm.py[i], vm.pz[i]);
    KnnHeap<16> heap;
    heap.k = kSphereKnn;
    knn_query(vm, p, heap);
    double m = -1.0;
    for (uint32_t q = 0; q < heap.size; q++) m = fmax(m, heap.d2[q]);
    const double r = sqrt(m);
    radius[i] = r;
    Aabb32 b;
    b.mn[0] = (float)(p.x - r); b.mn[1] = (float)(p.y - r); b.mn[2] = (float)(p.z - r);
    b.mx[0] = (float)(p.x + r); b.mx[1] = (float)(p.y + r); b.mx[2] = (float)(p.z + r);
    boxes[i] = b;
}

// ---- host helpers ----

int reduce_bounds(rptx_ctx* ctx, const double* d_blocks, int blocks, double mn[3], double mx[3]) {
    std::vector<double> h((size_t)blocks * 6);
    CK(cudaMemcpyAsync(h.data(), d_blocks, sizeof(double) * h.size(), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 3; k++) { mn[k] = INFINITY; mx[k] = -INFINITY; }
    for (int b = 0; b < blocks; b++)
        for (int k = 0; k < 3; k++) { mn[k] = std::fmin(mn[k], h[6 * b + k]); mx[k] = std::fmax(mx[k], h[6 * b + 3 + k]); }
    return 0;
}

// stable radix sort of (key, value) pairs; result in keys_out / vals_out
int sort_pairs(rptx_ctx* ctx, uint64_t* keys_in, uint64_t* keys_out, uint32_t* vals_in, uint32_t* vals_out, uint32_t n, int end_bit) {
    size_t tmp_bytes = 0;
    CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, end_bit, ctx->stream));
    void* tmp = nullptr;
    CK(rptx_alloc_async(&tmp, std::max<size_t>(tmp_bytes, 16), ctx->stream));
    cudaError_t e = cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys_in, keys_out, vals_in, vals_out, (int)n, 0, end_bit, ctx->stream);
    ctx->launches += 4;
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    rptx_free_async(tmp, ctx->stream);
    if (e != cudaSuccess) return ctx_fail(ctx, std::string("radix sort: ") + cudaGetErrorString(e));
    return 0;
}

// canonical order of a photon list: permutation sorting the rows by key
int canonical_order(rptx_ctx* ctx, const PhotonRec* rows, uint32_t n, uint32_t** d_idx_out) {
    uint64_t *k0 = nullptr, *k1 = nullptr;
    uint32_t *i0 = nullptr, *i1 = nullptr;
    CK(dev_alloc(ctx, &k0, n, nullptr));
    CK(dev_alloc(ctx, &k1, n, nullptr));
    CK(dev_alloc(ctx, &i0, n, nullptr));
    CK(dev_alloc(ctx, &i1, n, nullptr));
    k_rows_keys<<<grid_for(n, 256), 256, 0, ctx->stream>>>(rows, n, k0, i0);
    ctx->launches++;
    const int rc = sort_pairs(ctx, k0, k1, i0, i1, n, 64);
    rptx_free_async(k0, ctx->stream); rptx_free_async(k1, ctx->stream); rptx_free_async(i0, ctx->stream);
    if (rc) { rptx_free_async(i1, ctx->stream); return rc; }
    *d_idx_out = i1;
    return 0;
}

void free_owned(rptx_ctx* ctx, std::vector<void*>& owned) {
    for (void* p : owned) rptx_free_async(p, ctx->stream);
    owned.clear();
}

// KdTree::build_by (src/photon.rs:194-199): Morton-ordered structure-of-arrays + implicit box tree.
// d_canon: canonical permutation of the rows.
int build_point_map(rptx_ctx* ctx, const PhotonRec* rows, const uint32_t* d_canon, uint32_t n, PointMapHost* out) {
    cudaStream_t st = ctx->stream;
    out->dev = PointMapDev{};
    out->dev.n = n;
    if (n >= (1u << 28)) return ctx_fail(ctx, "photon map exceeds 2^28 records per list");
    double *px = nullptr, *py = nullptr, *pz = nullptr, *dir = nullptr, *power = nullptr;
    uint64_t* key = nullptr;
    BvhNode* nodes = nullptr;
    CK(dev_alloc(ctx, &px, n, &out->owned)); CK(dev_alloc(ctx, &py, n, &out->owned)); CK(dev_alloc(ctx, &pz, n, &out->owned));
    CK(dev_alloc(ctx, &dir, 3 * (size_t)n, &out->owned)); CK(dev_alloc(ctx, &power, 3 * (size_t)n, &out->owned));
    CK(dev_alloc(ctx, &key, n, &out->owned));
    CK(dev_alloc(ctx, &nodes, std::max<uint32_t>(n, 1), &out->owned));
    out->dev.nodes = nodes;
    out->dev.px = px; out->dev.py = py; out->dev.pz = pz; out->dev.dir = dir; out->dev.power = power; out->dev.key = key;
    if (n == 0) return 0;
    const int blocks = std::min(grid_for(n, 256), 1184);
    double* d_b = nullptr;
    CK(dev_alloc(ctx, &d_b, (size_t)blocks * 6, nullptr));
    k_rows_bounds<<<blocks, 256, 0, st>>>(rows, n, d_b);
    ctx->launches++;
    double mn[3], mx[3];
    const int rc0 = reduce_bounds(ctx, d_b, blocks, mn, mx);
    rptx_free_async(d_b, ctx->stream);
    if (rc0) return rc0;
    uint64_t *k0 = nullptr, *k1 = nullptr;
    uint32_t* i1 = nullptr;
    PrimBox* boxes = nullptr;
    int rc = 0;
    auto body = [&]() -> int {
        CK(dev_alloc(ctx, &k0, n, nullptr)); CK(dev_alloc(ctx, &k1, n, nullptr)); CK(dev_alloc(ctx, &i1, n, nullptr));
        const double3 dmn = make_double3(mn[0], mn[1], mn[2]);
        const double3 inv = make_double3(mx[0] > mn[0] ? 1.0 / (mx[0] - mn[0]) : 0.0, mx[1] > mn[1] ? 1.0 / (mx[1] - mn[1]) : 0.0,
                                         mx[2] > mn[2] ? 1.0 / (mx[2] - mn[2]) : 0.0);
        PhaseTrace trace(ctx);
        k_rows_morton<<<grid_for(n, 256), 256, 0, st>>>(rows, d_canon, n, dmn, inv, k0);
        ctx->launches++;
        trace.mark("  morton keys");
        if (int r = sort_pairs(ctx, k0, k1, const_cast<uint32_t*>(d_canon), i1, n, 63)) return r;
        trace.mark("  morton sort");
        k_rows_gather<<<grid_for(n, 256), 256, 0, st>>>(rows, i1, n, px, py, pz, dir, power, key);
        ctx->launches++;
        trace.mark("  rows -> SoA");
        rptx_free_async(k0, ctx->stream); k0 = nullptr;
        rptx_free_async(i1, ctx->stream); i1 = nullptr;
        CK(dev_alloc(ctx, &boxes, n, nullptr));
        k_point_boxes<<<grid_for(n, 256), 256, 0, st>>>(px, py, pz, n, boxes);
        ctx->launches++;
        const int e = build_lbvh_sorted(k1, boxes, n, 0.0f, kPointLeaf, nodes, &out->dev.root_ref, st, &ctx->launches);
        if (e != 0) return ctx_fail(ctx, std::string("point map LBVH: ") + cudaGetErrorString((cudaError_t)e));
        trace.mark("  boxes + LBVH");
        return 0;
    };
    rc = body();
    rptx_free_async(k0, ctx->stream); rptx_free_async(k1, ctx->stream); rptx_free_async(i1, ctx->stream); rptx_free_async(boxes, ctx->stream);
    return rc;
}

// The same map from lists that are spread over the ranks of the context's communicator (row f1 of SURVEY.md §8: the
// reference builds its kd-trees serially, src/photon.rs:193-199; here the N ranks split the work):
//   1. global bounds (all-reduce min / max) -> every rank computes the Morton keys the one-rank build would;
//   2. each rank sorts ITS rows by (Morton, canonical key) and counts them per Morton bin (2^20 bins = the key's top 20 bits);
//      the summed counts give every rank the same N - 1 splitters: rank r owns the r-th N-th of the key space by population;
//   3. personalised exchange (ncclSend / ncclRecv): a rank's sorted rows are already contiguous per destination;
//   4. rank r sorts what it received (1 / N of the rows) and writes its segment of the FINAL Morton-ordered SoA arrays;
//   5. ragged all-gather in place of the SoA segments and their Morton keys (88 B per photon instead of the 104-B rows);
//   6. the implicit box tree (Karras hierarchy + refit) over the complete sorted keys, on every rank.
// Sort, gather and key work are 1 / N per rank; step 6 is replicated.  The result is bit-identical to build_point_map over the
// concatenated lists: same keys, same (Morton, canonical) order, same tree.
constexpr int kShardBinBits = 20;

int build_point_map_sharded(rptx_ctx* ctx, const PhotonRec* rows, uint32_t n_local, PointMapHost* out) {
    using namespace rptx_host;
    cudaStream_t st = ctx->stream;
    const int world = comm_world(ctx), rank = comm_rank(ctx);
    out->dev = PointMapDev{};
    std::vector<void*> tmp;
    auto done = [&](int rc) { free_owned(ctx, tmp); return rc; };
#define CKT(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return done(ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_))); } while (0)
#define RCT(call) do { if (int r_ = (call)) return done(r_); } while (0)
    if (world > 63) return ctx_fail(ctx, "sharded photon-map build supports at most 63 ranks");  // k_pick_splitters: one block of 64
    PhaseTrace trace(ctx);
    // ---- 1. bounds ----
    double mn[3] = {INFINITY, INFINITY, INFINITY}, mx[3] = {-INFINITY, -INFINITY, -INFINITY};
    if (n_local) {
        const int blocks = std::min(grid_for(n_local, 256), 1184);
        double* d_b = nullptr;
        CKT(dev_alloc(ctx, &d_b, (size_t)blocks * 6, &tmp));
        k_rows_bounds<<<blocks, 256, 0, st>>>(rows, n_local, d_b);
        ctx->launches++;
        RCT(reduce_bounds(ctx, d_b, blocks, mn, mx));
    }
    double* d_mm = nullptr;
    CKT(dev_alloc(ctx, &d_mm, 6, &tmp));
    {
        const double h[6] = {mn[0], mn[1], mn[2], mx[0], mx[1], mx[2]};
        CKT(cudaMemcpyAsync(d_mm, h, sizeof(h), cudaMemcpyHostToDevice, st));
        CKT(cudaStreamSynchronize(st));  // h goes out of scope
        RCT(comm_allreduce_f64(ctx, d_mm, 3, 1));
        RCT(comm_allreduce_f64(ctx, d_mm + 3, 3, 2));
        double g[6];
        CKT(cudaMemcpyAsync(g, d_mm, sizeof(g), cudaMemcpyDeviceToHost, st));
        CKT(cudaStreamSynchronize(st));
        for (int k = 0; k < 3; k++) { mn[k] = g[k]; mx[k] = g[3 + k]; }
    }
    const double3 dmn = make_double3(mn[0], mn[1], mn[2]);
    const double3 inv = make_double3(mx[0] > mn[0] ? 1.0 / (mx[0] - mn[0]) : 0.0, mx[1] > mn[1] ? 1.0 / (mx[1] - mn[1]) : 0.0,
                                     mx[2] > mn[2] ? 1.0 / (mx[2] - mn[2]) : 0.0);
    // ---- 2. local (Morton, canonical) order and the splitters ----
    uint32_t *canon = nullptr, *i1 = nullptr;
    uint64_t *k0 = nullptr, *k1 = nullptr;
    RCT(canonical_order(ctx, rows, n_local, &canon));
    tmp.push_back(canon);
    CKT(dev_alloc(ctx, &k0, n_local, &tmp)); CKT(dev_alloc(ctx, &k1, n_local, &tmp)); CKT(dev_alloc(ctx, &i1, n_local, &tmp));
    if (n_local) {
        k_rows_morton<<<grid_for(n_local, 256), 256, 0, st>>>(rows, canon, n_local, dmn, inv, k0);
        ctx->launches++;
        RCT(sort_pairs(ctx, k0, k1, canon, i1, n_local, 63));
    }
    const uint32_t bins = 1u << kShardBinBits;
    unsigned long long *lb_local = nullptr, *lb_global = nullptr, *d_cut = nullptr, *d_mat = nullptr;
    CKT(dev_alloc(ctx, &lb_local, (size_t)bins + 1, &tmp)); CKT(dev_alloc(ctx, &lb_global, (size_t)bins + 1, &tmp));
    CKT(dev_alloc(ctx, &d_cut, (size_t)world + 1, &tmp)); CKT(dev_alloc(ctx, &d_mat, (size_t)world * world, &tmp));
    k_bin_lower_bounds<<<grid_for((size_t)bins + 1, 256), 256, 0, st>>>(k1, n_local, 63 - kShardBinBits, bins, lb_local);
    CKT(cudaMemcpyAsync(lb_global, lb_local, sizeof(unsigned long long) * ((size_t)bins + 1), cudaMemcpyDeviceToDevice, st));
    RCT(comm_allreduce_u64_sum(ctx, lb_global, (size_t)bins + 1));
    k_pick_splitters<<<1, 64, 0, st>>>(lb_global, lb_local, bins, world, d_cut);
    ctx->launches += 2;
    std::vector<unsigned long long> cut((size_t)world + 1), mat((size_t)world * world);
    CKT(cudaMemcpyAsync(cut.data(), d_cut, sizeof(unsigned long long) * cut.size(), cudaMemcpyDeviceToHost, st));
    CKT(cudaStreamSynchronize(st));
    {   // counts matrix: mat[s][r] = rows rank s sends to rank r
        std::vector<unsigned long long> mine((size_t)world);
        for (int r = 0; r < world; r++) mine[r] = cut[r + 1] - cut[r];
        CKT(cudaMemcpyAsync(d_mat + (size_t)rank * world, mine.data(), sizeof(unsigned long long) * world, cudaMemcpyHostToDevice, st));
        CKT(cudaStreamSynchronize(st));
        RCT(comm_allgather_inplace(ctx, d_mat, sizeof(unsigned long long) * world));
        CKT(cudaMemcpyAsync(mat.data(), d_mat, sizeof(unsigned long long) * mat.size(), cudaMemcpyDeviceToHost, st));
        CKT(cudaStreamSynchronize(st));
    }
    std::vector<unsigned long long> seg_cnt((size_t)world, 0), seg_off((size_t)world, 0);
    unsigned long long n_total = 0;
    for (int r = 0; r < world; r++) {
        for (int sr = 0; sr < world; sr++) seg_cnt[r] += mat[(size_t)sr * world + r];
        seg_off[r] = n_total;
        n_total += seg_cnt[r];
    }
    if (n_total >= (1ull << 28)) return done(ctx_fail(ctx, "photon map exceeds 2^28 records per list"));
    trace.mark("  shard: local sort + splitters");
    // ---- 3. exchange ----
    const uint32_t m = (uint32_t)seg_cnt[rank];
    PhotonRec *srows = nullptr, *rrows = nullptr;
    CKT(dev_alloc(ctx, &srows, n_local, &tmp)); CKT(dev_alloc(ctx, &rrows, m, &tmp));
    if (n_local) {
        k_rows_permute<<<grid_for(n_local, 256), 256, 0, st>>>(rows, i1, n_local, srows);
        ctx->launches++;
    }
    {
        std::vector<size_t> so((size_t)world), sc((size_t)world), ro((size_t)world), rc((size_t)world);
        size_t acc = 0;
        for (int r = 0; r < world; r++) {
            so[r] = (size_t)cut[r] * sizeof(PhotonRec);
            sc[r] = (size_t)(cut[r + 1] - cut[r]) * sizeof(PhotonRec);
            ro[r] = acc;
            rc[r] = (size_t)mat[(size_t)r * world + rank] * sizeof(PhotonRec);
            acc += rc[r];
        }
        RCT(comm_alltoallv(ctx, srows, so.data(), sc.data(), rrows, ro.data(), rc.data()));
        CKT(cudaStreamSynchronize(st));  // the offset vectors go out of scope
    }
    trace.mark("  shard: exchange");
    // ---- 4. this rank's segment of the final arrays ----
    const uint32_t n = (uint32_t)n_total;
    out->dev.n = n;
    double *px = nullptr, *py = nullptr, *pz = nullptr, *dir = nullptr, *power = nullptr;
    uint64_t *key = nullptr, *mkey = nullptr;
    BvhNode* nodes = nullptr;
    CKT(dev_alloc(ctx, &px, n, &out->owned)); CKT(dev_alloc(ctx, &py, n, &out->owned)); CKT(dev_alloc(ctx, &pz, n, &out->owned));
    CKT(dev_alloc(ctx, &dir, 3 * (size_t)n, &out->owned)); CKT(dev_alloc(ctx, &power, 3 * (size_t)n, &out->owned));
    CKT(dev_alloc(ctx, &key, n, &out->owned));
    CKT(dev_alloc(ctx, &nodes, std::max<uint32_t>(n, 1), &out->owned));
    CKT(dev_alloc(ctx, &mkey, n, &tmp));
    out->dev.nodes = nodes;
    out->dev.px = px; out->dev.py = py; out->dev.pz = pz; out->dev.dir = dir; out->dev.power = power; out->dev.key = key;
    if (n == 0) return done(0);
    if (m) {
        uint32_t *canon2 = nullptr, *i2 = nullptr;
        uint64_t* q0 = nullptr;
        RCT(canonical_order(ctx, rrows, m, &canon2));
        tmp.push_back(canon2);
        CKT(dev_alloc(ctx, &q0, m, &tmp)); CKT(dev_alloc(ctx, &i2, m, &tmp));
        const size_t off = (size_t)seg_off[rank];
        k_rows_morton<<<grid_for(m, 256), 256, 0, st>>>(rrows, canon2, m, dmn, inv, q0);
        ctx->launches++;
        RCT(sort_pairs(ctx, q0, mkey + off, canon2, i2, m, 63));
        k_rows_gather<<<grid_for(m, 256), 256, 0, st>>>(rrows, i2, m, px + off, py + off, pz + off, dir + 3 * off, power + 3 * off, key + off);
        ctx->launches++;
    }
    trace.mark("  shard: segment sort + SoA");
    // ---- 5. everybody gets every segment ----
    {
        void* bufs[7] = {px, py, pz, dir, power, key, mkey};
        const size_t elem[7] = {8, 8, 8, 24, 24, 8, 8};
        RCT(comm_allgatherv_inplace(ctx, bufs, elem, 7, seg_off.data(), seg_cnt.data()));
        CKT(cudaStreamSynchronize(st));
    }
    trace.mark("  shard: all-gather segments");
    // ---- 6. the tree ----
    PrimBox* boxes = nullptr;
    CKT(dev_alloc(ctx, &boxes, n, &tmp));
    k_point_boxes<<<grid_for(n, 256), 256, 0, st>>>(px, py, pz, n, boxes);
    ctx->launches++;
    const int e = build_lbvh_sorted(mkey, boxes, n, 0.0f, kPointLeaf, nodes, &out->dev.root_ref, st, &ctx->launches);
    if (e != 0) return done(ctx_fail(ctx, std::string("point map LBVH: ") + cudaGetErrorString((cudaError_t)e)));
    trace.mark("  shard: boxes + LBVH");
#undef CKT
#undef RCT
    return done(0);
}

// BVH::build over item boxes (bvh 0.6.0): LBVH with exact f32 unions.  perm[k] = item in leaf slot k.
int build_box_bvh(rptx_ctx* ctx, const Aabb32* d_boxes, uint32_t n, BoxBvhHost* out, uint32_t** d_perm_out) {
    out->dev = BoxBvhDev{};
    out->dev.n = n;
    *d_perm_out = nullptr;
    BvhNode* nodes = nullptr;
    CK(dev_alloc(ctx, &nodes, std::max<uint32_t>(n, 1), &out->owned));
    out->dev.nodes = nodes;
    if (n == 0) return 0;
    const int blocks = std::min(grid_for(n, 256), 1184);
    double* d_b = nullptr;
    CK(dev_alloc(ctx, &d_b, (size_t)blocks * 6, nullptr));
    k_box_bounds<<<blocks, 256, 0, ctx->stream>>>(d_boxes, n, d_b);
    ctx->launches++;
    double mn[3], mx[3];
    const int rc0 = reduce_bounds(ctx, d_b, blocks, mn, mx);
    rptx_free_async(d_b, ctx->stream);
    if (rc0) return rc0;
    for (int k = 0; k < 3; k++)
        if (!(mn[k] <= mx[k])) { mn[k] = 0.0; mx[k] = 1.0; }
    uint32_t* perm = nullptr;
    CK(dev_alloc(ctx, &perm, n, nullptr));
    static_assert(sizeof(PrimBox) == sizeof(Aabb32), "box layouts match");
    const int rc = build_lbvh(reinterpret_cast<const PrimBox*>(d_boxes), n, 0.0f, kLeafMax, mn, mx, nodes, perm, &out->dev.root_ref, ctx->stream,
                              &ctx->launches);
    if (rc != 0) { rptx_free_async(perm, ctx->stream); return ctx_fail(ctx, std::string("photon BVH build: ") + cudaGetErrorString((cudaError_t)rc)); }
    *d_perm_out = perm;
    return 0;
}

// every rank's rows of one list, concatenated in rank order (PhotonList::merge across ranks for ONE list)
int allgather_rows(rptx_ctx* ctx, const PhotonRec* rows, uint32_t n_local, PhotonRec** d_full, uint64_t* n_full) {
    using namespace rptx_host;
    cudaStream_t st = ctx->stream;
    const int world = comm_world(ctx), rank = comm_rank(ctx);
    *d_full = nullptr;
    unsigned long long* d_cnt = nullptr;
    CK(dev_alloc(ctx, &d_cnt, (size_t)world, nullptr));
    const unsigned long long mine = n_local;
    std::vector<unsigned long long> cnt((size_t)world), off((size_t)world);
    cudaError_t e = cudaMemcpyAsync(d_cnt + rank, &mine, sizeof(mine), cudaMemcpyHostToDevice, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    int rc = e == cudaSuccess ? comm_allgather_inplace(ctx, d_cnt, sizeof(unsigned long long)) : ctx_fail(ctx, cudaGetErrorString(e));
    if (rc == 0) {
        e = cudaMemcpyAsync(cnt.data(), d_cnt, sizeof(unsigned long long) * world, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) rc = ctx_fail(ctx, cudaGetErrorString(e));
    }
    rptx_free_async(d_cnt, st);
    if (rc) return rc;
    unsigned long long total = 0;
    for (int r = 0; r < world; r++) { off[r] = total; total += cnt[r]; }
    if (total >= (1ull << 32)) return ctx_fail(ctx, "photon list exceeds 2^32 records");
    PhotonRec* full = nullptr;
    CK(dev_alloc(ctx, &full, (size_t)total, nullptr));
    if (n_local) CK(cudaMemcpyAsync(full + off[rank], rows, (size_t)n_local * sizeof(PhotonRec), cudaMemcpyDeviceToDevice, st));
    void* bufs[1] = {full};
    const size_t elem[1] = {sizeof(PhotonRec)};
    rc = comm_allgatherv_inplace(ctx, bufs, elem, 1, off.data(), cnt.data());
    if (rc == 0 && cudaStreamSynchronize(st) != cudaSuccess) rc = ctx_fail(ctx, "all-gather of photon rows failed");
    if (rc) { rptx_free_async(full, st); return rc; }
    *d_full = full;
    *n_full = total;
    return 0;
}

int find_photon_light(const rptx_scene* s, uint32_t* index) {
    for (size_t i = 0; i < s->lights.size(); i++)
        if (s->lights[i].kind == RPTX_LIGHT_OBJECT) { *index = (uint32_t)i; return 0; }
    return 1;
}

}  // namespace

struct rptx_photon_map {
    rptx_ctx* ctx = nullptr;
    rptx_scene* scene = nullptr;
    int kind = 0;
    PointMapHost surface, volume;
    BoxBvhHost bvh;
    BoxBvhHost seg_bvh;  // over the beam segments (BeamSegDev)
    SphereMapDev spheres = {};
    BeamMapDev beams = {};
    std::vector<void*> owned;
    double build_ms = 0;
};

namespace {

// run the emission kernel once over [first, first + count); returns the deposit counts
int emit_once(rptx_ctx* ctx, rptx_scene* scene, const EmitParams& ep, PhotonRec* d_surf, uint64_t cap_s, PhotonRec* d_vol, uint64_t cap_v,
              uint64_t counts[3]) {
    cudaStream_t st = ctx->stream;
    unsigned long long* d_ctr = nullptr;  // [0] surface count, [1] volume count, [2] work head, [3] rays
    CK(rptx_alloc_async((void**)&d_ctr, sizeof(unsigned long long) * 4, st));
    CK(cudaMemsetAsync(d_ctr, 0, sizeof(unsigned long long) * 4, st));
    PhotonSink surf{d_surf, d_ctr, cap_s}, vol{d_vol, d_ctr + 1, cap_v};
    const SceneDev sc = rptx_host::scene_dev(scene);
    const MediumDev md = rptx_host::medium_dev(scene);
    const int blocks = (int)std::min<uint64_t>((uint64_t)ctx->sm_count * 8, (ep.count + 127) / 128);
    if (sc.small_scene) k_photon_trace<true><<<std::max(blocks, 1), 128, 0, st>>>(sc, md, ep, surf, vol, d_ctr + 2, d_ctr + 3);
    else k_photon_trace<false><<<std::max(blocks, 1), 128, 0, st>>>(sc, md, ep, surf, vol, d_ctr + 2, d_ctr + 3);
    ctx->launches++;
    unsigned long long h[4] = {0, 0, 0, 0};
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d_ctr, sizeof(h), cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    rptx_free_async(d_ctr, ctx->stream);
    if (e != cudaSuccess) return ctx_fail(ctx, std::string("photon emission: ") + cudaGetErrorString(e));
    counts[0] = h[0]; counts[1] = h[1]; counts[2] = h[3];
    return 0;
}

template <int KMAX>
void launch_gather(const SceneDev& sc, const MediumDev& md, const PhotonMapDev& pm, const GatherArgs& ga, unsigned long long* stats,
                   cudaStream_t st) {
    const size_t threads = (size_t)ga.n_pixels * ga.n_lanes;
    k_photon_gather<KMAX><<<grid_for(threads, 128), 128, 0, st>>>(sc, md, pm, ga, stats);
}

}  // namespace

extern "C" {

int rptx_photons_create(rptx_ctx* ctx, rptx_photons** out) {
    if (!ctx || !out) return 1;
    rptx_photons* p = new rptx_photons();
    p->ctx = ctx;
    *out = p;
    return 0;
}

void rptx_photons_destroy(rptx_photons* p) {
    if (!p) return;
    cudaSetDevice(p->ctx->device);
    rptx_free_async(p->d_rows[0], p->ctx->stream); rptx_free_async(p->d_rows[1], p->ctx->stream);
    delete p;
}

int rptx_photons_emit(rptx_ctx* ctx, rptx_scene* scene, const rptx_params* params, uint64_t photon_count_total, uint64_t first,
                      uint64_t count, int32_t kind, rptx_photons** out) {
    if (!ctx || !out) return 1;
    *out = nullptr;
    if (!scene || !params) return ctx_fail(ctx, "rptx_photons_emit: null argument");
    if (scene->ctx != ctx) return ctx_fail(ctx, "scene belongs to another context");
    if (!scene->committed) return ctx_fail(ctx, "scene not committed");
    if (kind < RPTX_PHOTON_MAP || kind > RPTX_PHOTON_BEAM_BEAM) return ctx_fail(ctx, "unknown photon render kind");
    if (photon_count_total == 0 || first + count > photon_count_total) return ctx_fail(ctx, "bad photon range");
    if (photon_count_total >= (1ull << 47)) return ctx_fail(ctx, "too many photons");
    EmitParams ep;
    ep.seed = params->seed; ep.first = first; ep.count = count; ep.kind = kind;
    ep.power = params->watts / (double)photon_count_total;
    if (find_photon_light(scene, &ep.light_index)) return ctx_fail(ctx, "Only found non-object lights while photon mapping");  // src/photon.rs:798
    CK(cudaSetDevice(ctx->device));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, ctx->stream));
    rptx_photons* p = new rptx_photons();
    p->ctx = ctx;
    uint64_t counts[3] = {0, 0, 0};
    uint64_t cap_s, cap_v;
    int rc = 0;
    PhaseTrace trace(ctx);
    if (count > (1ull << 17)) {
        // counting-only pilot over the first 2^16 photons sizes the append buffers
        EmitParams pilot = ep;
        pilot.count = 1ull << 16;
        rc = emit_once(ctx, scene, pilot, nullptr, 0, nullptr, 0, counts);
        const double rs = (double)counts[0] / (double)pilot.count, rv = (double)counts[1] / (double)pilot.count;
        cap_s = (uint64_t)((double)count * rs * 1.03) + 65536;
        cap_v = (uint64_t)((double)count * rv * 1.03) + 65536;
    } else {
        cap_s = cap_v = count * 16 + 1024;
    }
    while (rc == 0) {
        rptx_free_async(p->d_rows[0], ctx->stream); rptx_free_async(p->d_rows[1], ctx->stream);
        p->d_rows[0] = p->d_rows[1] = nullptr;
        cudaError_t e = rptx_alloc_async((void**)&p->d_rows[0], std::max<uint64_t>(cap_s, 1) * sizeof(PhotonRec), ctx->stream);
        if (e == cudaSuccess) e = rptx_alloc_async((void**)&p->d_rows[1], std::max<uint64_t>(cap_v, 1) * sizeof(PhotonRec), ctx->stream);
        if (e != cudaSuccess) { rc = ctx_fail(ctx, std::string("photon buffers: ") + cudaGetErrorString(e)); break; }
        p->cap[0] = cap_s; p->cap[1] = cap_v;
        trace.mark("emit: pilot + buffers");
        rc = emit_once(ctx, scene, ep, p->d_rows[0], cap_s, p->d_rows[1], cap_v, counts);
        trace.mark("emit: k_photon_trace");
        if (rc) break;
        if (counts[0] <= cap_s && counts[1] <= cap_v) break;
        cap_s = std::max(cap_s, counts[0]); cap_v = std::max(cap_v, counts[1]);  // deterministic walk: exact sizes now known
    }
    if (rc == 0) {
        cudaEventRecord(e1, ctx->stream);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        p->emit_ms = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc) { rptx_photons_destroy(p); return rc; }
    if (counts[0] >= (1ull << 32) || counts[1] >= (1ull << 32)) { rptx_photons_destroy(p); return ctx_fail(ctx, "photon list exceeds 2^32 records"); }
    p->n[0] = counts[0]; p->n[1] = counts[1]; p->emitted = count; p->rays = counts[2];
    *out = p;
    return 0;
}

static int photons_reserve(rptx_photons* p, int which, uint64_t need) {
    rptx_ctx* ctx = p->ctx;
    if (need <= p->cap[which]) return 0;
    const uint64_t cap = std::max<uint64_t>(need, p->cap[which] * 2);
    PhotonRec* nr = nullptr;
    CK(rptx_alloc_async((void**)&nr, cap * sizeof(PhotonRec), ctx->stream));
    if (p->n[which]) CK(cudaMemcpyAsync(nr, p->d_rows[which], p->n[which] * sizeof(PhotonRec), cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    rptx_free_async(p->d_rows[which], ctx->stream);
    p->d_rows[which] = nr;
    p->cap[which] = cap;
    return 0;
}

}  // extern "C"

namespace rptx_host {
int photons_resize(rptx_photons* p, int which, uint64_t n) {
    if (int rc = photons_reserve(p, which, n)) return rc;
    p->n[which] = n;
    return 0;
}
}  // namespace rptx_host

extern "C" {

int rptx_photons_append_host(rptx_photons* p, int32_t which, const void* rows, size_t n) {
    if (!p) return 1;
    rptx_ctx* ctx = p->ctx;
    if (which < 0 || which > 1) return ctx_fail(ctx, "which must be 0 (surface) or 1 (volume)");
    if (n == 0) return 0;
    if (!rows) return ctx_fail(ctx, "null rows");
    CK(cudaSetDevice(ctx->device));
    if (p->n[which] + n >= (1ull << 32)) return ctx_fail(ctx, "photon list exceeds 2^32 records");
    if (int rc = photons_reserve(p, which, p->n[which] + n)) return rc;
    CK(cudaMemcpyAsync(p->d_rows[which] + p->n[which], rows, n * sizeof(PhotonRec), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    p->n[which] += n;
    return 0;
}

int rptx_photons_append_device(rptx_photons* p, int32_t which, const void* d_rows, size_t n) {
    if (!p) return 1;
    rptx_ctx* ctx = p->ctx;
    if (which < 0 || which > 1) return ctx_fail(ctx, "which must be 0 (surface) or 1 (volume)");
    if (n == 0) return 0;
    if (!d_rows) return ctx_fail(ctx, "null rows");
    CK(cudaSetDevice(ctx->device));
    if (p->n[which] + n >= (1ull << 32)) return ctx_fail(ctx, "photon list exceeds 2^32 records");
    if (int rc = photons_reserve(p, which, p->n[which] + n)) return rc;
    CK(cudaMemcpyAsync(p->d_rows[which] + p->n[which], d_rows, n * sizeof(PhotonRec), cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    p->n[which] += n;
    return 0;
}

int rptx_photons_count(const rptx_photons* p, uint64_t counts[2]) {
    if (!p || !counts) return 1;
    counts[0] = p->n[0]; counts[1] = p->n[1];
    return 0;
}

void* rptx_photons_device_ptr(rptx_photons* p, int32_t which) { return (p && which >= 0 && which <= 1) ? (void*)p->d_rows[which] : nullptr; }

int rptx_photons_download(rptx_photons* p, int32_t which, void* rows) {
    if (!p) return 1;
    rptx_ctx* ctx = p->ctx;
    if (which < 0 || which > 1 || !rows) return ctx_fail(ctx, "rptx_photons_download: bad argument");
    CK(cudaSetDevice(ctx->device));
    if (p->n[which]) CK(cudaMemcpyAsync(rows, p->d_rows[which], p->n[which] * sizeof(PhotonRec), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rptx_photons_stats(const rptx_photons* p, uint64_t* emitted, uint64_t* rays, double* emit_ms) {
    if (!p) return 1;
    if (emitted) *emitted = p->emitted;
    if (rays) *rays = p->rays;
    if (emit_ms) *emit_ms = p->emit_ms;
    return 0;
}

void rptx_photon_map_destroy(rptx_photon_map* m) {
    if (!m) return;
    cudaSetDevice(m->ctx->device);
    free_owned(m->ctx, m->surface.owned); free_owned(m->ctx, m->volume.owned); free_owned(m->ctx, m->bvh.owned); free_owned(m->ctx, m->seg_bvh.owned);
    free_owned(m->ctx, m->owned);
    delete m;
}

// sharded: `photons` is this rank's part of lists spread over the context's communicator; every rank ends up with the same,
// complete map (build_point_map_sharded), the one rptx_photon_map_build makes of the merged lists
static int photon_map_build_impl(rptx_ctx* ctx, rptx_scene* scene, rptx_photons* photons, int32_t kind, bool sharded, rptx_photon_map** out) {
    if (!ctx || !out) return 1;
    *out = nullptr;
    if (!scene || !photons) return ctx_fail(ctx, "rptx_photon_map_build: null argument");
    // (RPTX_FORCE_SHARDED_BUILD: tests run the sharded code on one rank, where the exchanges are local copies)
    if (sharded && rptx_host::comm_world(ctx) == 1 && !std::getenv("RPTX_FORCE_SHARDED_BUILD")) sharded = false;
    if (photons->ctx != ctx || scene->ctx != ctx) return ctx_fail(ctx, "objects belong to another context");
    if (kind < RPTX_PHOTON_MAP || kind > RPTX_PHOTON_BEAM_BEAM) return ctx_fail(ctx, "unknown photon render kind");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    CK(cudaEventRecord(e0, st));
    rptx_photon_map* m = new rptx_photon_map();
    m->ctx = ctx; m->scene = scene; m->kind = kind;
    int rc = 0;
    uint32_t *canon_s = nullptr, *canon_v = nullptr, *perm = nullptr;
    uint32_t ns = (uint32_t)photons->n[0];
    uint32_t nv = (uint32_t)photons->n[1];
    const PhotonRec* srows = photons->d_rows[0];
    const PhotonRec* vrows = photons->d_rows[1];
    PhotonRec *srows_full = nullptr, *vrows_full = nullptr;  // merged rows, where the ranks need them whole
    // Point maps are split by Morton range from 4 ranks on.  Measured at 1e8 photons (profiles/r2_summary.md): the exchange,
    // the segment all-gather and the replicated tree cost about what the split sorts save at 2 ranks (photon-photon step
    // 272 ms merged-and-replicated, 305 split) and win from 4 on (8 ranks: 239 -> 194 ms); below that the rows are merged
    // and every rank builds the maps itself.  The radius pass of the beam-photon map is split at any rank count.
    int min_world = 4;
    if (const char* e = std::getenv("RPTX_SHARDED_MAPS_MIN_WORLD")) min_world = std::max(1, std::atoi(e));
    const bool shard_maps = sharded && (rptx_host::comm_world(ctx) >= min_world || std::getenv("RPTX_FORCE_SHARDED_BUILD"));
    PhaseTrace trace(ctx);
    auto body = [&]() -> int {
        if (sharded && !shard_maps) {
            uint64_t ts = 0, tv = 0;
            if (int r = allgather_rows(ctx, srows, ns, &srows_full, &ts)) return r;
            if (int r = allgather_rows(ctx, vrows, nv, &vrows_full, &tv)) return r;
            srows = srows_full; vrows = vrows_full;
            ns = (uint32_t)ts; nv = (uint32_t)tv;
            trace.mark("build: rows all-gather");
        }
        if (shard_maps) {
            if (int r = build_point_map_sharded(ctx, photons->d_rows[0], ns, &m->surface)) return r;
            trace.mark("build: point map surface (sharded)");
            if (kind == RPTX_PHOTON_MAP || kind == RPTX_PHOTON_POINT_BEAM) {
                if (int r = build_point_map_sharded(ctx, photons->d_rows[1], nv, &m->volume)) return r;
                nv = m->volume.dev.n;
                trace.mark("build: point map volume (sharded)");
            } else {
                // the beams are made of the rows themselves, in canonical order: every rank needs every volume row
                uint64_t total = 0;
                if (int r = allgather_rows(ctx, photons->d_rows[1], nv, &vrows_full, &total)) return r;
                vrows = vrows_full;
                nv = (uint32_t)total;
                if (int r = canonical_order(ctx, vrows, nv, &canon_v)) return r;
                trace.mark("build: volume rows all-gather + canonical");
            }
        } else {
        if (int r = canonical_order(ctx, srows, ns, &canon_s)) return r;
        trace.mark("build: canonical surface");
        if (int r = build_point_map(ctx, srows, canon_s, ns, &m->surface)) return r;
        trace.mark("build: point map surface");
        rptx_free_async(srows_full, ctx->stream); srows_full = nullptr;
        if (int r = canonical_order(ctx, vrows, nv, &canon_v)) return r;
        trace.mark("build: canonical volume");
        if (kind == RPTX_PHOTON_MAP || kind == RPTX_PHOTON_POINT_BEAM) {
            if (int r = build_point_map(ctx, vrows, canon_v, nv, &m->volume)) return r;
            trace.mark("build: point map volume");
        }
        }
        if (kind == RPTX_PHOTON_POINT_BEAM) {
            double *radius = nullptr, *center = nullptr, *rad = nullptr, *dir = nullptr, *power = nullptr;
            Aabb32 *boxes = nullptr, *obox = nullptr;
            CK(dev_alloc(ctx, &radius, nv, &m->owned));
            CK(dev_alloc(ctx, &boxes, nv, &m->owned));
            if (nv && sharded) {
                // the 10-NN radius pass (src/photon.rs:212-236) split by photon index, then all-gathered
                const int world = rptx_host::comm_world(ctx);
                std::vector<unsigned long long> off((size_t)world), cnt((size_t)world);
                for (int r = 0; r < world; r++) {
                    off[r] = (unsigned long long)nv * r / world;
                    cnt[r] = (unsigned long long)nv * (r + 1) / world - off[r];
                }
                const int me = rptx_host::comm_rank(ctx);
                if (cnt[me]) {
                    k_sphere_radius_range<<<grid_for(cnt[me], 128), 128, 0, st>>>(m->volume.dev, (uint32_t)off[me], (uint32_t)cnt[me], radius, boxes);
                    ctx->launches++;
                }
                void* bufs[2] = {radius, boxes};
                const size_t elem[2] = {sizeof(double), sizeof(Aabb32)};
                if (int r = rptx_host::comm_allgatherv_inplace(ctx, bufs, elem, 2, off.data(), cnt.data())) return r;
                CK(cudaStreamSynchronize(st));
                trace.mark("build: sphere radii (sharded)");
            } else if (nv) {
                k_sphere_radius<<<grid_for(nv, 128), 128, 0, st>>>(m->volume.dev, radius, boxes);
                ctx->launches++;
            }
            if (int r = build_box_bvh(ctx, boxes, nv, &m->bvh, &perm)) return r;
            CK(dev_alloc(ctx, &center, 3 * (size_t)nv, &m->owned)); CK(dev_alloc(ctx, &rad, nv, &m->owned));
            CK(dev_alloc(ctx, &dir, 3 * (size_t)nv, &m->owned)); CK(dev_alloc(ctx, &power, 3 * (size_t)nv, &m->owned));
            CK(dev_alloc(ctx, &obox, nv, &m->owned));
            if (nv) {
                k_sphere_gather<<<grid_for(nv, 256), 256, 0, st>>>(m->volume.dev, radius, boxes, perm, nv, center, rad, dir, power, obox);
                ctx->launches++;
            }
            m->bvh.dev.boxes = obox;
            m->spheres.bvh = m->bvh.dev;
            m->spheres.center = center; m->spheres.radius = rad; m->spheres.dir = dir; m->spheres.power = power;
        } else if (kind == RPTX_PHOTON_BEAM_BEAM) {
            double *start = nullptr, *dirn = nullptr, *len = nullptr, *power = nullptr, *ostart = nullptr, *odirn = nullptr, *olen = nullptr,
                   *opower = nullptr;
            Aabb32 *boxes = nullptr, *obox = nullptr;
            CK(dev_alloc(ctx, &start, 3 * (size_t)nv, &m->owned)); CK(dev_alloc(ctx, &dirn, 3 * (size_t)nv, &m->owned));
            CK(dev_alloc(ctx, &len, nv, &m->owned)); CK(dev_alloc(ctx, &power, 3 * (size_t)nv, &m->owned));
            CK(dev_alloc(ctx, &boxes, nv, &m->owned));
            if (nv) {
                k_beam_make<<<grid_for(nv, 256), 256, 0, st>>>(vrows, canon_v, nv, start, dirn, len, power, boxes);
                ctx->launches++;
            }
            if (int r = build_box_bvh(ctx, boxes, nv, &m->bvh, &perm)) return r;
            CK(dev_alloc(ctx, &ostart, 3 * (size_t)nv, &m->owned)); CK(dev_alloc(ctx, &odirn, 3 * (size_t)nv, &m->owned));
            CK(dev_alloc(ctx, &olen, nv, &m->owned)); CK(dev_alloc(ctx, &opower, 3 * (size_t)nv, &m->owned));
            CK(dev_alloc(ctx, &obox, nv, &m->owned));
            if (nv) {
                k_beam_gather<<<grid_for(nv, 256), 256, 0, st>>>(start, dirn, len, power, boxes, perm, nv, ostart, odirn, olen, opower, obox);
                ctx->launches++;
            }
            m->bvh.dev.boxes = obox;
            m->beams.bvh = m->bvh.dev;
            m->beams.start = ostart; m->beams.dirn = odirn; m->beams.len = olen; m->beams.power = opower;
            m->beams.seg = BeamSegDev{};
            if (nv && !std::getenv("RPTX_NO_BEAM_SEGMENTS")) {
                // the segments of the warp camera pass (BeamSegDev): count, offsets, boxes, their own BVH, leaf order
                uint32_t *cnt = nullptr, *off = nullptr, *sbeam = nullptr, *obeam = nullptr, *perm2 = nullptr;
                uint16_t *sk = nullptr, *sK = nullptr, *ok = nullptr, *oK = nullptr;
                Aabb32 *sbox = nullptr, *osbox = nullptr;
                std::vector<void*> tmp;
                auto fin = [&](int rc2) { free_owned(ctx, tmp); rptx_free_async(perm2, st); return rc2; };
                CK(dev_alloc(ctx, &cnt, nv, &tmp)); CK(dev_alloc(ctx, &off, nv, &tmp));
                double seg_len = kBeamSegLen;
                if (const char* e = std::getenv("RPTX_BEAM_SEG_LEN")) seg_len = std::max(0.5, std::atof(e));  // developer knob
                k_seg_count<<<grid_for(nv, 256), 256, 0, st>>>(olen, nv, seg_len, cnt);
                size_t scan_bytes = 0;
                CK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, cnt, off, (int)nv, st));
                void* scan_tmp = nullptr;
                CK(dev_alloc(ctx, (char**)&scan_tmp, scan_bytes, &tmp));
                CK(cub::DeviceScan::ExclusiveSum(scan_tmp, scan_bytes, cnt, off, (int)nv, st));
                uint32_t last[2] = {0, 0};
                CK(cudaMemcpyAsync(&last[0], off + (nv - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                CK(cudaMemcpyAsync(&last[1], cnt + (nv - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
                CK(cudaStreamSynchronize(st));
                const uint32_t ns = last[0] + last[1];
                CK(dev_alloc(ctx, &sbox, ns, &tmp)); CK(dev_alloc(ctx, &sbeam, ns, &tmp)); CK(dev_alloc(ctx, &sk, ns, &tmp)); CK(dev_alloc(ctx, &sK, ns, &tmp));
                k_seg_make<<<grid_for(nv, 128), 128, 0, st>>>(ostart, odirn, olen, off, cnt, nv, sbox, sbeam, sk, sK);
                ctx->launches += 4;
                if (int r = build_box_bvh(ctx, sbox, ns, &m->seg_bvh, &perm2)) return fin(r);
                CK(dev_alloc(ctx, &osbox, ns, &m->owned)); CK(dev_alloc(ctx, &obeam, ns, &m->owned));
                CK(dev_alloc(ctx, &ok, ns, &m->owned)); CK(dev_alloc(ctx, &oK, ns, &m->owned));
                k_seg_gather<<<grid_for(ns, 256), 256, 0, st>>>(sbox, sbeam, sk, sK, perm2, ns, osbox, obeam, ok, oK);
                ctx->launches++;
                CK(cudaStreamSynchronize(st));
                m->seg_bvh.dev.boxes = osbox;
                m->beams.seg.bvh = m->seg_bvh.dev;
                m->beams.seg.beam = obeam; m->beams.seg.k = ok; m->beams.seg.K = oK;
                fin(0);
            }
        }
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(st));
        trace.mark("build: spheres / beams + BVH");
        return 0;
    };
    rc = body();
    rptx_free_async(canon_s, ctx->stream); rptx_free_async(canon_v, ctx->stream); rptx_free_async(perm, ctx->stream);
    rptx_free_async(vrows_full, ctx->stream); rptx_free_async(srows_full, ctx->stream);
    if (rc == 0) {
        cudaEventRecord(e1, st);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        m->build_ms = ms;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (rc) { rptx_photon_map_destroy(m); return rc; }
    *out = m;
    return 0;
}

int rptx_photon_map_build(rptx_ctx* ctx, rptx_scene* scene, rptx_photons* photons, int32_t kind, rptx_photon_map** out) {
    return photon_map_build_impl(ctx, scene, photons, kind, false, out);
}

int rptx_photon_map_build_sharded(rptx_ctx* ctx, rptx_scene* scene, rptx_photons* local_photons, int32_t kind, rptx_photon_map** out) {
    return photon_map_build_impl(ctx, scene, local_photons, kind, true, out);
}

int rptx_photon_map_stats(const rptx_photon_map* m, uint64_t counts[3], double* build_ms) {
    if (!m) return 1;
    if (counts) { counts[0] = m->surface.dev.n; counts[1] = m->volume.dev.n; counts[2] = m->bvh.dev.n; }
    if (build_ms) *build_ms = m->build_ms;
    return 0;
}

int rptx_photon_gather_trace(rptx_film* f, rptx_photon_map* m, uint32_t iterations) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (!m) return ctx_fail(ctx, "null photon map");
    if (m->ctx != ctx || m->scene != f->scene) return ctx_fail(ctx, "photon map belongs to another context or scene");
    if (iterations == 0) return ctx_fail(ctx, "iterations must be > 0");
    const uint64_t kmax = std::max(f->params.gather_size, f->params.gather_size_volume);
    if (kmax > 1024) return ctx_fail(ctx, "gather sizes above 1024 are not supported");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    GatherArgs ga;
    ga.cam = f->cam;
    BatchParams& bp = ga.bp;
    bp.width = f->params.width; bp.height = f->params.height; bp.max_bounces = f->params.max_bounces;
    bp.seed = f->params.seed; bp.batch_base = f->next_sample; bp.iterations = iterations;
    bp.shard_count = f->params.shard_count; bp.shard_index = f->params.shard_index;
    bp.n_local = iterations > bp.shard_index ? (iterations - bp.shard_index + bp.shard_count - 1) / bp.shard_count : 0;
    ga.n_pixels = f->pool.n_pixels;
    ga.n_lanes = std::max<uint32_t>(1, std::min(f->pool.n_lanes, bp.n_local));
    ga.ax = f->pool.ax; ga.ay = f->pool.ay; ga.az = f->pool.az;
    PhotonMapDev pm;
    std::memset(&pm, 0, sizeof(pm));
    pm.kind = m->kind;
    pm.surface = m->surface.dev; pm.volume = m->volume.dev; pm.spheres = m->spheres; pm.beams = m->beams;
    pm.gather_size = (uint32_t)f->params.gather_size; pm.gather_size_volume = (uint32_t)f->params.gather_size_volume;
    const SceneDev sc = rptx_host::scene_dev(f->scene);
    const MediumDev md = rptx_host::medium_dev(f->scene);
    CK(cudaEventRecord(f->ev0, st));
    const bool warp_mode = kmax <= kWarpKnnMaxK && !std::getenv("RPTX_GATHER_PER_THREAD");
    if (warp_mode) {
        // one warp per slot: ~64 K slots keep every SM busy for several rounds
        ga.n_lanes = std::max<uint32_t>(1, std::min<uint32_t>(ga.n_lanes, std::max<uint32_t>(1, 65536u / ga.n_pixels)));
        const size_t warps = (size_t)ga.n_pixels * ga.n_lanes;
        k_photon_gather_warp<<<grid_for(warps, 4), 128, 0, st>>>(sc, md, pm, ga, f->q.stats);
    } else if (kmax <= 32) launch_gather<32>(sc, md, pm, ga, f->q.stats, st);
    else if (kmax <= 128) launch_gather<128>(sc, md, pm, ga, f->q.stats, st);
    else launch_gather<1024>(sc, md, pm, ga, f->q.stats, st);
    k_lane_reduce<<<grid_for(ga.n_pixels, 128), 128, 0, st>>>(ga.ax, ga.ay, ga.az, ga.n_pixels, ga.n_lanes, f->d_batch_sum);
    ctx->launches += 2;
    CK(cudaGetLastError());
    CK(cudaEventRecord(f->ev1, st));
    CK(cudaStreamSynchronize(st));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, f->ev0, f->ev1));
    f->stats.device_ms = ms;
    return 0;
}

int rptx_photon_map_knn(rptx_photon_map* m, int32_t which, const double* q, size_t n, uint32_t k, uint64_t* keys, double* d2,
                        uint32_t* counts) {
    if (!m) return 1;
    rptx_ctx* ctx = m->ctx;
    if (which < 0 || which > 1) return ctx_fail(ctx, "which must be 0 (surface) or 1 (volume)");
    if (k == 0 || k > 1024) return ctx_fail(ctx, "k must be in 1..1024");
    if (n == 0) return 0;
    if (!q || !keys || !d2 || !counts) return ctx_fail(ctx, "rptx_photon_map_knn: null argument");
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    double *d_q = nullptr, *d_d2 = nullptr;
    uint64_t* d_keys = nullptr;
    uint32_t* d_counts = nullptr;
    std::vector<void*> tmp;
    auto done = [&](int rc) { free_owned(ctx, tmp); return rc; };
#define CKT(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return done(ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_))); } while (0)
    CKT(dev_alloc(ctx, &d_q, 3 * n, &tmp)); CKT(dev_alloc(ctx, &d_d2, n * k, &tmp)); CKT(dev_alloc(ctx, &d_keys, n * k, &tmp)); CKT(dev_alloc(ctx, &d_counts, n, &tmp));
    CKT(cudaMemcpyAsync(d_q, q, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    const PointMapDev pmap = which == 0 ? m->surface.dev : m->volume.dev;
    if (k <= 32) k_knn_batch<32><<<grid_for(n, 128), 128, 0, st>>>(pmap, d_q, n, k, d_keys, d_d2, d_counts);
    else if (k <= 128) k_knn_batch<128><<<grid_for(n, 128), 128, 0, st>>>(pmap, d_q, n, k, d_keys, d_d2, d_counts);
    else k_knn_batch<1024><<<grid_for(n, 128), 128, 0, st>>>(pmap, d_q, n, k, d_keys, d_d2, d_counts);
    ctx->launches++;
    CKT(cudaGetLastError());
    CKT(cudaMemcpyAsync(keys, d_keys, sizeof(uint64_t) * n * k, cudaMemcpyDeviceToHost, st));
    CKT(cudaMemcpyAsync(d2, d_d2, sizeof(double) * n * k, cudaMemcpyDeviceToHost, st));
    CKT(cudaMemcpyAsync(counts, d_counts, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, st));
    CKT(cudaStreamSynchronize(st));
#undef CKT
    return done(0);
}

// sphere radii (PhotonPointBeam maps) keyed by photon, for parity checks: radius of the sphere in BVH slot i and its centre
int rptx_photon_map_spheres(rptx_photon_map* m, double* center, double* radius) {
    if (!m) return 1;
    rptx_ctx* ctx = m->ctx;
    if (m->kind != RPTX_PHOTON_POINT_BEAM) return ctx_fail(ctx, "not a point-beam map");
    CK(cudaSetDevice(ctx->device));
    const size_t n = m->bvh.dev.n;
    if (n && center) CK(cudaMemcpyAsync(center, m->spheres.center, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, ctx->stream));
    if (n && radius) CK(cudaMemcpyAsync(radius, m->spheres.radius, sizeof(double) * n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rptx_photon_render(rptx_ctx* ctx, rptx_scene* scene, const rptx_camera* camera, const rptx_params* params, uint64_t photon_count,
                       int32_t kind, uint8_t* out_rgb8, double* out_mean) {
    if (!ctx) return 1;
    if (!scene || !camera || !params) return ctx_fail(ctx, "rptx_photon_render: null argument");
    rptx_photons* ph = nullptr;
    rptx_photon_map* map = nullptr;
    rptx_film* film = nullptr;
    // with a communicator (rptx_comm_init) the three passes shard as photon_render's rayon loops do (src/photon.rs:663-716):
    // this rank emits its photon-index range, the lists are all-gathered, every rank gathers its sample shard of the camera
    // pass and the batch sums are all-reduced
    int world = 1, rank = 0;
    rptx_comm_info(ctx, &world, &rank);
    rptx_params prm = *params;
    if (world > 1) { prm.shard_count = (uint32_t)world; prm.shard_index = (uint32_t)rank; }
    else if (prm.shard_count > 1) return ctx_fail(ctx, "sharded photon_render needs a communicator (rptx_comm_init), or the three passes with the caller's own exchange");
    const uint64_t lo = photon_count * (uint64_t)rank / (uint64_t)world, hi = photon_count * (uint64_t)(rank + 1) / (uint64_t)world;
    int rc = rptx_photons_emit(ctx, scene, &prm, photon_count, lo, hi - lo, kind, &ph);
    if (rc == 0 && world > 1 && std::getenv("RPTX_REPLICATED_MAP_BUILD")) {  // round-1 scheme, kept for comparison runs
        rptx_photons* full = nullptr;
        rc = rptx_photons_allgather(ph, &full);
        rptx_photons_destroy(ph);
        ph = full;
        if (rc == 0) rc = rptx_photon_map_build(ctx, scene, ph, kind, &map);
    } else if (rc == 0) {
        rc = rptx_photon_map_build_sharded(ctx, scene, ph, kind, &map);  // == rptx_photon_map_build without a communicator
    }
    if (ph) { rptx_photons_destroy(ph); ph = nullptr; }
    if (rc == 0) rc = rptx_photon_gather_begin(ctx, scene, camera, &prm, &film);
    if (rc == 0) rc = rptx_photon_gather_trace(film, map, params->num_samples);
    if (rc == 0 && world > 1) rc = rptx_render_sample_allreduce(film);
    if (rc == 0) rc = rptx_render_sample_commit(film, params->num_samples);
    if (rc == 0 && out_rgb8) rc = rptx_film_image_rgb8(film, out_rgb8);
    if (rc == 0 && out_mean) rc = rptx_film_mean(film, out_mean);
    if (film) rptx_film_destroy(film);
    if (map) rptx_photon_map_destroy(map);
    return rc;
}

}  // extern "C"

// rptx_intersect_batch_static: the closest-hit batch of rptx_intersect_batch for callers that may not allocate
// device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
//
// rptx_intersect_batch stages its rays and results in seven cudaMalloc'ed arrays per call.  Here that scratch is
// ONE file-scope `static __device__` array of fixed size (kStaticBytes, ~5.8 MiB for RPTX_INTERSECT_STATIC_MAX_RAYS
// rays), part of the library's device image: every GPU that loads librptx.so holds it until the process exits.
//   * The array's NAME in host code is a host-side stand-in, not a device address.  Its device address is taken
//     once per device, with that device current, by cudaGetSymbolAddress (static_scratch() below), and that
//     pointer is what the copies and the kernel receive.
//   * Every call shares the array and finds the previous call's contents in it: the hit counter at its head is
//     cleared at the start of each call, the ray / result regions are fully overwritten for [0, n) before they
//     are read.  Two calls that overlap on the GPU (different contexts or host threads on one device) race on
//     it; calls through one context are ordered by the context's stream and the final synchronise.
//   * The entry point calls no cudaMalloc / cudaMallocAsync / cudaFree and no allocating library routine.
//
// No counterpart in the reference.
#include <mutex>

#include "host_types.hpp"
#include "trace.cuh"

using namespace rptx;
using rptx_host::ctx_fail;
using rptx_host::scene_dev;

namespace {

constexpr size_t kStaticRays = RPTX_INTERSECT_STATIC_MAX_RAYS;

// layout of the scratch: a 256-byte head (hit counter), then the ray and result arrays, f64 before u32
struct StaticBatch {
    unsigned long long hits;
    unsigned char pad[256 - sizeof(unsigned long long)];
    double o[3 * kStaticRays];
    double d[3 * kStaticRays];
    double t[kStaticRays];
    double normal[3 * kStaticRays];
    uint32_t object_id[kStaticRays];
    uint32_t prim_id[kStaticRays];
    uint32_t inst_id[kStaticRays];
};
constexpr size_t kStaticBytes = sizeof(StaticBatch);
static_assert(kStaticBytes <= (8u << 20), "the static scratch is meant to stay modest (<= 8 MiB per GPU)");

static __device__ __align__(256) unsigned char g_static_batch[kStaticBytes];

// k_intersect_batch (path_kernels.cuh) reading its rays from and writing its results to the static scratch, plus
// a count of the rays that hit something (one atomic per warp)
__global__ void __launch_bounds__(kBlock) k_intersect_batch_static(const __grid_constant__ SceneDev sc, StaticBatch* __restrict__ s, uint32_t n,
                                                                    double t_min) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RayD ray;
    ray.o = mk3(s->o[3 * i], s->o[3 * i + 1], s->o[3 * i + 2]);
    ray.d = mk3(s->d[3 * i], s->d[3 * i + 1], s->d[3 * i + 2]);
    Hit h;
    h.t = CUDART_INF;
    closest_hit<kClosest>(sc, ray, t_min, 0.0, h);
    uint32_t prim = 0, inst = 0;
    d3 n3 = mk3(0, 0, 0);
    const bool hit = h.inst != kNoHit;
    if (hit) n3 = hit_normal(sc, ray, h, &prim, &inst);
    s->t[i] = hit ? h.t : CUDART_INF;
    s->object_id[i] = h.obj;
    s->prim_id[i] = prim;
    s->inst_id[i] = inst;
    s->normal[3 * i] = n3.x; s->normal[3 * i + 1] = n3.y; s->normal[3 * i + 2] = n3.z;
    stat_add(&s->hits, hit ? 1ull : 0ull);
}

// device address of g_static_batch on `device` (which must be current): resolved once per device
constexpr int kMaxDevices = 64;
cudaError_t static_scratch(int device, StaticBatch** out) {
    static std::mutex mu;
    static StaticBatch* ptr[kMaxDevices] = {};
    *out = nullptr;
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_batch);
        if (e != cudaSuccess) return e;
        ptr[device] = static_cast<StaticBatch*>(p);
    }
    *out = ptr[device];
    return cudaSuccess;
}

}  // namespace

#define CK(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) return ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_));   \
    } while (0)

extern "C" int rptx_intersect_batch_static(rptx_scene* scene, const double* o, const double* d, size_t n, double t_min, double* t,
                                           uint32_t* object_id, uint32_t* prim_id, uint32_t* inst_id, double* normal, uint64_t* n_hits) {
    if (!scene) return 1;
    rptx_ctx* ctx = scene->ctx;
    if (!scene->committed) return ctx_fail(ctx, "scene not committed");
    if (n > kStaticRays)
        return ctx_fail(ctx, "rptx_intersect_batch_static: n = " + std::to_string(n) + " rays exceeds RPTX_INTERSECT_STATIC_MAX_RAYS = " +
                                 std::to_string(kStaticRays) + " (the static scratch is fixed-size; split the batch or use rptx_intersect_batch)");
    if (n_hits) *n_hits = 0;
    if (n == 0) return 0;
    if (!o || !d) return ctx_fail(ctx, "null ray arrays");
    CK(cudaSetDevice(ctx->device));
    StaticBatch* s = nullptr;
    CK(static_scratch(ctx->device, &s));
    cudaStream_t st = ctx->stream;
    CK(cudaMemsetAsync(&s->hits, 0, sizeof(s->hits), st));  // the array keeps the last call's count
    CK(cudaMemcpyAsync(s->o, o, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->d, d, sizeof(double) * 3 * n, cudaMemcpyHostToDevice, st));
    const SceneDev sc = scene_dev(scene);
    k_intersect_batch_static<<<(unsigned)((n + kBlock - 1) / kBlock), kBlock, 0, st>>>(sc, s, (uint32_t)n, t_min);
    ctx->launches++;
    CK(cudaGetLastError());
    if (t) CK(cudaMemcpyAsync(t, s->t, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
    if (object_id) CK(cudaMemcpyAsync(object_id, s->object_id, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, st));
    if (prim_id) CK(cudaMemcpyAsync(prim_id, s->prim_id, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, st));
    if (inst_id) CK(cudaMemcpyAsync(inst_id, s->inst_id, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost, st));
    if (normal) CK(cudaMemcpyAsync(normal, s->normal, sizeof(double) * 3 * n, cudaMemcpyDeviceToHost, st));
    unsigned long long hits = 0;
    CK(cudaMemcpyAsync(&hits, &s->hits, sizeof(hits), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    if (n_hits) *n_hits = hits;
    return 0;
}

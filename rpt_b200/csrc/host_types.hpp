// Host-side objects behind the opaque handles of include/rptx.h.
#pragma once
#include <algorithm>
#include <array>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rptx.h"
#include "common.cuh"
#include "path.cuh"
#include "shading.cuh"

struct rptx_ctx {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;
    void* h_pinned = nullptr;
    uint64_t launches = 0;
    std::string err;
    // Path-state arena: one slab per film (tens of GB for a wavefront pool), handed back here when the film is
    // destroyed and reused by the next film of this context, so a render loop maps its HBM once
    // (cudaMalloc of 23 GB costs ~50 ms; iterative_render / repeated render() calls would pay it every time).
    void* spare_arena = nullptr;
    size_t spare_arena_bytes = 0;
    // multi-GPU: NCCL communicator of this rank (comm.cu), null for a single-device context
    struct rptx_comm* comm = nullptr;
};

constexpr uint32_t kMaxBounces = 4096;  // rptx_params::max_bounces above this is rejected (the clamp stack is 64 B per level and slot)

struct HostMesh {
    std::vector<rptx::RawTri> tris;  // original order
    bool has_normals = true;
    double bmin[3], bmax[3];
    // filled by commit
    uint32_t raw_base = 0, tri_base = 0, node_base = 0;
    int root_ref = 0;
    float pad = 0.f;
};

struct HostObject {
    rptx_shape_desc desc;
    std::vector<rptx_shape_desc> children;
    rptx_material material;
};

struct rptx_scene {
    rptx_ctx* ctx = nullptr;
    std::vector<HostMesh> meshes;
    std::vector<HostObject> objects;
    std::vector<rptx_light> lights;
    int medium_kind = RPTX_MEDIUM_NONE;
    double sigma_a = 0, sigma_s = 0;
    bool env_is_hdri = false;
    double env_color[3] = {0, 0, 0};
    uint32_t env_w = 0, env_h = 0;
    std::vector<double> env_pixels;
    // device state
    bool committed = false;
    const rptx::BvhNode* d_nodes = nullptr;
    const rptx::Instance* d_instances = nullptr;
    const rptx::TriRec* d_tris = nullptr;
    const rptx::TriF* d_trif = nullptr;
    const rptx::TriNormals* d_tri_normals = nullptr;
    const uint32_t* d_tri_orig = nullptr;
    const rptx::Material* d_materials = nullptr;
    const rptx::Light* d_lights = nullptr;
    const rptx::RawTri* d_raw = nullptr;
    const double* d_env = nullptr;
    const uint32_t* d_cube_idx = nullptr;
    uint32_t n_cubes = 0;
    int small_scene = 0;
    uint8_t* d_small_slots = nullptr;
    float4* d_small_boxes = nullptr;
    uint32_t small_kind_mask[3] = {0, 0, 0};
    uint32_t small_unbounded_mask = 0;
    rptx::SmallTables small_tables = {};
    int tlas_root_ref = 0;
    uint32_t n_bounded = 0, n_unbounded = 0, n_shadow_lights = 0, shadow_probe_mask = 0;
    double wmin[3] = {0, 0, 0}, wmax[3] = {0, 0, 0};
};

enum FilmMode : int { kFilmWavefront = 0, kFilmFused = 1, kFilmGather = 2, kFilmPersist = 3 };

struct rptx_film {
    rptx_ctx* ctx = nullptr;
    void* arena = nullptr;
    size_t arena_bytes = 0;
    char* d_select_tmp = nullptr;  // cub::DeviceSelect scratch of the queue compaction
    size_t select_tmp_bytes = 0;
    int mode = kFilmWavefront;
    rptx_scene* scene = nullptr;
    rptx_params params;
    rptx::Camera cam;
    rptx::PathPool pool;
    double* d_pstack = nullptr;    // persistent kernel, surface mode: clamp records [level][thread] (persist.cuh)
    int persist_blocks = 0;        // grid of the persistent kernel (resident blocks of the device)
    rptx::Queues q;
    std::vector<void*> owned;
    double* d_batch_sum = nullptr;
    double* d_film_sum = nullptr;
    double* d_scratch = nullptr;
    std::vector<double*> batches;  // one mean image per Renderer::sample call, as Buffer keeps them
    std::vector<double*> pooled_means;  // those of `batches` that came from cudaMallocAsync
    double* d_mean_slots = nullptr;     // the first n_mean_slots batch means live in the arena
    uint32_t n_mean_slots = 0;
    double** d_batch_ptrs = nullptr;
    uint64_t next_sample = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t evw[4][5] = {};   // per-wave kernel boundaries, one set per wave in flight (run_waves)
    cudaEvent_t ev_done[4] = {};  // the wave's live count has landed in pinned memory
    bool no_one_vertex = false;   // RPTX_NO_ONE_VERTEX
    uint32_t shade_rounds = 0;    // RPTX_SHADE_TILE_ROUNDS (0: chosen per wave)
    rptx_stats stats = {};
    bool timing_pending = false;  // ev0 / ev1 of the last traced batch not read yet
};

// Stream-ordered allocation for everything that is created and destroyed per call (scene buffers, BVH build
// scratch): cudaMalloc / cudaFree take a device-wide lock and, with a 20 GB film slab mapped, stalled for
// 20-460 ms at random in the end-to-end loop; the pool (release threshold set in rptx_ctx_create) does neither.
// Large requests (>= 32 MB: photon rows, maps, sort and tree scratch - GBs each at 1e8 photons) do not go to the pool at all.
// The pool sub-allocates: a smaller request may split the cached block a larger one needs a moment later, the larger one then
// maps fresh memory inside a timed step (0.2-0.8 s for 5-20 GB), and which step pays is random (one in three at 1e8 photons, in
// every size-rounding scheme tried: profiles/r2_summary.md section 6).  They are served by a caching allocator over cudaMalloc
// instead (rptx.cu): sizes are sticky (a request reuses the smallest size handed out before that holds it with at most 30 % to
// spare, else gets 10 % headroom rounded to 256 MB), freed blocks wait in a per-stream list of whole blocks that are never split,
// and a block is only ever reused on the stream it was used on (stream order makes that safe without a synchronisation).
// cudaMalloc runs in the first calls only; rptx_ctx_trim / rptx_ctx_destroy give the cached blocks back.
cudaError_t rptx_alloc_async(void** p, size_t bytes, cudaStream_t st);
void rptx_free_async(const void* p, cudaStream_t st);
void rptx_block_cache_release(cudaStream_t st);  // cudaFree every cached block of the stream (which must be idle)

// host helpers shared by rptx.cu and photon.cu
namespace rptx_host {
int ctx_fail(rptx_ctx* ctx, const std::string& msg);  // records the message, returns 1
rptx::SceneDev scene_dev(const rptx_scene* s);
rptx::MediumDev medium_dev(const rptx_scene* s);
// persist.cu
int persist_grid(int sm_count, bool medium, bool probe, bool small);
// comm.cu: collectives of the library's own multi-GPU passes, on the context's stream (no-ops / local copies without a communicator)
int comm_world(const rptx_ctx* ctx);
int comm_rank(const rptx_ctx* ctx);
int comm_allreduce_f64(rptx_ctx* ctx, double* d_buf, size_t n, int op /* 0 sum, 1 min, 2 max */);
int comm_allreduce_u64_sum(rptx_ctx* ctx, unsigned long long* d_buf, size_t n);
int comm_allgather_inplace(rptx_ctx* ctx, void* d_recv, size_t bytes_per_rank);
int comm_allgatherv_inplace(rptx_ctx* ctx, void* const* d_bufs, const size_t* elem, int arrays, const unsigned long long* row_off,
                            const unsigned long long* row_cnt);
int comm_alltoallv(rptx_ctx* ctx, const void* d_send, const size_t* send_off, const size_t* send_cnt, void* d_recv, const size_t* recv_off,
                   const size_t* recv_cnt);
}  // namespace rptx_host

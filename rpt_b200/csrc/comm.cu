// Multi-GPU exchange steps of the path, inside the library (C ABI of include/rptx.h, "multi-GPU" section).
//
// The reference has one address space; its two exchange points are `add_samples` after the rows' collect()
// (src/renderer.rs:169-170) and PhotonList::merge (src/photon.rs:153-160,675).  One process per GPU: every
// rank owns an rptx_ctx; rptx_comm_init joins the ranks' contexts into an NCCL communicator (NVLink 5 /
// NVSwitch on the box) and the two exchanges become
//   * rptx_render_sample_allreduce: ncclAllReduce(sum, f64) of the batch-sum film, 3*w*h doubles;
//   * rptx_photons_allgather: counts first (ncclAllGather of 2 x u64 per rank), then ONE grouped set of
//     ncclBroadcast calls per list straight into the merged list at each rank's row offset — no padding to
//     the longest list and no staging copies.
// Both are enqueued on the context's stream behind the kernels that produce their input.
//
// libnccl is loaded at run time (dlopen "libnccl.so.2"): a single-GPU host needs no NCCL at all, and a Python
// host that already mapped torch's bundled NCCL shares that copy.
#include <dlfcn.h>
#include <nccl.h>

#include <cstring>
#include <string>
#include <vector>

#include "../../include/rptx.h"
#include "host_types.hpp"
#include "photon_host.hpp"

using rptx_host::ctx_fail;

namespace {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};

NcclApi* nccl_api() {
    static NcclApi api;
    if (api.handle || !api.error.empty()) return &api;
    const char* names[] = {std::getenv("RPTX_NCCL_LIBRARY"), "libnccl.so.2", "libnccl.so"};
    for (const char* n : names) {
        if (!n) continue;
        api.handle = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (api.handle) break;
    }
    if (!api.handle) { api.error = std::string("cannot load libnccl.so.2: ") + dlerror(); return &api; }
#define SYM(field, name)                                                               \
    api.field = reinterpret_cast<decltype(api.field)>(dlsym(api.handle, name));        \
    if (!api.field) { api.error = std::string("libnccl lacks ") + name; return &api; }
    SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
    SYM(AllReduce, "ncclAllReduce") SYM(AllGather, "ncclAllGather") SYM(Broadcast, "ncclBroadcast")
    SYM(Send, "ncclSend") SYM(Recv, "ncclRecv")
    SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
    return &api;
}

}  // namespace

struct rptx_comm {
    ncclComm_t comm = nullptr;
    int world = 1, rank = 0;
    unsigned long long* d_counts = nullptr;  // [world][2] row counts of an all-gather
};

#define CKN(call)                                                                                          \
    do {                                                                                                   \
        ncclResult_t r_ = (call);                                                                          \
        if (r_ != ncclSuccess) return ctx_fail(ctx, std::string(#call) + ": " + nccl_api()->GetErrorString(r_)); \
    } while (0)
#define CK(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) return ctx_fail(ctx, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

// ---- collectives for the library's own multi-GPU passes (photon.cu: sharded photon-map build) ----
// All are enqueued on the context's stream; none synchronises.  Without a communicator (world 1) they are no-ops.
namespace rptx_host {

int comm_world(const rptx_ctx* ctx) { return ctx->comm ? ctx->comm->world : 1; }
int comm_rank(const rptx_ctx* ctx) { return ctx->comm ? ctx->comm->rank : 0; }

// op: 0 sum, 1 min, 2 max
int comm_allreduce_f64(rptx_ctx* ctx, double* d_buf, size_t n, int op) {
    if (!ctx->comm || n == 0) return 0;
    const ncclRedOp_t ops[3] = {ncclSum, ncclMin, ncclMax};
    CKN(nccl_api()->AllReduce(d_buf, d_buf, n, ncclFloat64, ops[op], ctx->comm->comm, ctx->stream));
    return 0;
}

int comm_allreduce_u64_sum(rptx_ctx* ctx, unsigned long long* d_buf, size_t n) {
    if (!ctx->comm || n == 0) return 0;
    CKN(nccl_api()->AllReduce(d_buf, d_buf, n, ncclUint64, ncclSum, ctx->comm->comm, ctx->stream));
    return 0;
}

// every rank contributes `bytes` at d_recv + rank * bytes (in place) and receives the others' blocks
int comm_allgather_inplace(rptx_ctx* ctx, void* d_recv, size_t bytes) {
    if (!ctx->comm || bytes == 0) return 0;
    char* base = static_cast<char*>(d_recv);
    CKN(nccl_api()->AllGather(base + (size_t)ctx->comm->rank * bytes, base, bytes, ncclUint8, ctx->comm->comm, ctx->stream));
    return 0;
}

// ragged all-gather in place: segment r = [off[r], off[r] + cnt[r]) bytes of d_buf is valid on rank r and is broadcast from
// there; `arrays` buffers with the same row partition (row sizes elem[a]) go out as ONE group
int comm_allgatherv_inplace(rptx_ctx* ctx, void* const* d_bufs, const size_t* elem, int arrays, const unsigned long long* row_off,
                            const unsigned long long* row_cnt) {
    if (!ctx->comm) return 0;
    NcclApi* a = nccl_api();
    rptx_comm* c = ctx->comm;
    CKN(a->GroupStart());
    for (int k = 0; k < arrays; k++)
        for (int r = 0; r < c->world; r++) {
            const size_t bytes = (size_t)row_cnt[r] * elem[k];
            if (!bytes) continue;
            char* seg = static_cast<char*>(d_bufs[k]) + (size_t)row_off[r] * elem[k];
            const ncclResult_t rr = a->Broadcast(seg, seg, bytes, ncclUint8, r, c->comm, ctx->stream);
            if (rr != ncclSuccess) { a->GroupEnd(); return ctx_fail(ctx, std::string("ncclBroadcast: ") + a->GetErrorString(rr)); }
        }
    CKN(a->GroupEnd());
    return 0;
}

// personalised exchange: send_cnt[r] bytes from d_send + send_off[r] go to rank r and land at d_recv + recv_off[s] on the
// receiver, s the sender (grouped ncclSend / ncclRecv; the block a rank keeps is copied on the stream)
int comm_alltoallv(rptx_ctx* ctx, const void* d_send, const size_t* send_off, const size_t* send_cnt, void* d_recv, const size_t* recv_off,
                   const size_t* recv_cnt) {
    const char* src = static_cast<const char*>(d_send);
    char* dst = static_cast<char*>(d_recv);
    if (!ctx->comm) {
        if (send_cnt[0]) CK(cudaMemcpyAsync(dst + recv_off[0], src + send_off[0], send_cnt[0], cudaMemcpyDeviceToDevice, ctx->stream));
        return 0;
    }
    NcclApi* a = nccl_api();
    rptx_comm* c = ctx->comm;
    if (send_cnt[c->rank]) CK(cudaMemcpyAsync(dst + recv_off[c->rank], src + send_off[c->rank], send_cnt[c->rank], cudaMemcpyDeviceToDevice, ctx->stream));
    CKN(a->GroupStart());
    for (int r = 0; r < c->world; r++) {
        if (r == c->rank) continue;
        ncclResult_t rr = ncclSuccess;
        if (send_cnt[r]) rr = a->Send(src + send_off[r], send_cnt[r], ncclUint8, r, c->comm, ctx->stream);
        if (rr == ncclSuccess && recv_cnt[r]) rr = a->Recv(dst + recv_off[r], recv_cnt[r], ncclUint8, r, c->comm, ctx->stream);
        if (rr != ncclSuccess) { a->GroupEnd(); return ctx_fail(ctx, std::string("ncclSend/Recv: ") + a->GetErrorString(rr)); }
    }
    CKN(a->GroupEnd());
    return 0;
}

}  // namespace rptx_host

extern "C" {

int rptx_comm_unique_id(void* id_out) {
    if (!id_out) return 1;
    NcclApi* a = nccl_api();
    if (!a->error.empty()) return 1;
    ncclUniqueId id;
    if (a->GetUniqueId(&id) != ncclSuccess) return 1;
    static_assert(sizeof(id) == RPTX_COMM_ID_BYTES, "ncclUniqueId is 128 bytes");
    std::memcpy(id_out, &id, sizeof(id));
    return 0;
}

int rptx_comm_init(rptx_ctx* ctx, const void* id, int world, int rank) {
    if (!ctx) return 1;
    if (!id || world < 1 || rank < 0 || rank >= world) return ctx_fail(ctx, "rptx_comm_init: bad argument");
    if (ctx->comm) return ctx_fail(ctx, "context already has a communicator");
    NcclApi* a = nccl_api();
    if (!a->error.empty()) return ctx_fail(ctx, a->error);
    CK(cudaSetDevice(ctx->device));
    ncclUniqueId uid;
    std::memcpy(&uid, id, sizeof(uid));
    rptx_comm* c = new rptx_comm();
    c->world = world; c->rank = rank;
    ncclResult_t r = a->CommInitRank(&c->comm, world, uid, rank);
    if (r != ncclSuccess) { delete c; return ctx_fail(ctx, std::string("ncclCommInitRank: ") + a->GetErrorString(r)); }
    if (cudaMalloc((void**)&c->d_counts, sizeof(unsigned long long) * 2 * (size_t)world) != cudaSuccess) {
        a->CommDestroy(c->comm);
        delete c;
        return ctx_fail(ctx, "rptx_comm_init: out of memory");
    }
    ctx->comm = c;
    return 0;
}

int rptx_comm_info(const rptx_ctx* ctx, int* world, int* rank) {
    if (!ctx) return 1;
    if (world) *world = ctx->comm ? ctx->comm->world : 1;
    if (rank) *rank = ctx->comm ? ctx->comm->rank : 0;
    return 0;
}

void rptx_comm_destroy(rptx_ctx* ctx) {
    if (!ctx || !ctx->comm) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    nccl_api()->CommDestroy(ctx->comm->comm);
    cudaFree(ctx->comm->d_counts);
    delete ctx->comm;
    ctx->comm = nullptr;
}

// `add_samples` across ranks (src/renderer.rs:169-170): every rank ends up with the batch sum of all shards
int rptx_render_sample_allreduce(rptx_film* f) {
    if (!f) return 1;
    rptx_ctx* ctx = f->ctx;
    if (!ctx->comm) return ctx->comm == nullptr && f->params.shard_count == 1 ? 0 : ctx_fail(ctx, "sharded film on a context without a communicator (rptx_comm_init)");
    if ((int)f->params.shard_count != ctx->comm->world || (int)f->params.shard_index != ctx->comm->rank)
        return ctx_fail(ctx, "film shard (shard_count, shard_index) does not match the communicator (world, rank)");
    CK(cudaSetDevice(ctx->device));
    const size_t n3 = (size_t)f->pool.n_pixels * 3;
    CKN(nccl_api()->AllReduce(f->d_batch_sum, f->d_batch_sum, n3, ncclFloat64, ncclSum, ctx->comm->comm, ctx->stream));
    return 0;
}

// PhotonList::merge across ranks (src/photon.rs:153-160,675): rows of rank r land at offset sum(n_0..n_{r-1})
int rptx_photons_allgather(rptx_photons* local, rptx_photons** out) {
    if (!local || !out) return 1;
    rptx_ctx* ctx = local->ctx;
    *out = nullptr;
    if (!ctx->comm) return ctx_fail(ctx, "rptx_photons_allgather: the context has no communicator (rptx_comm_init)");
    rptx_comm* c = ctx->comm;
    NcclApi* a = nccl_api();
    CK(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const unsigned long long mine[2] = {local->n[0], local->n[1]};
    CK(cudaMemcpyAsync(c->d_counts + 2 * c->rank, mine, sizeof(mine), cudaMemcpyHostToDevice, st));
    CKN(a->AllGather(c->d_counts + 2 * c->rank, c->d_counts, 2, ncclUint64, c->comm, st));
    std::vector<unsigned long long> counts(2 * (size_t)c->world);
    CK(cudaMemcpyAsync(counts.data(), c->d_counts, sizeof(unsigned long long) * counts.size(), cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    rptx_photons* full = nullptr;
    if (int rc = rptx_photons_create(ctx, &full)) return rc;
    for (int which = 0; which < 2; which++) {
        uint64_t total = 0;
        for (int r = 0; r < c->world; r++) total += counts[2 * r + which];
        if (total >= (1ull << 32)) { rptx_photons_destroy(full); return ctx_fail(ctx, "photon list exceeds 2^32 records"); }
        if (int rc = rptx_host::photons_resize(full, which, total)) { rptx_photons_destroy(full); return rc; }
        char* dst = static_cast<char*>(rptx_photons_device_ptr(full, which));
        const char* src = static_cast<const char*>(rptx_photons_device_ptr(local, which));
        CKN(a->GroupStart());
        uint64_t off = 0;
        for (int r = 0; r < c->world; r++) {
            const size_t bytes = (size_t)counts[2 * r + which] * RPTX_PHOTON_ROW_BYTES;
            if (bytes) {
                ncclResult_t rr = a->Broadcast(r == c->rank ? src : dst + off, dst + off, bytes, ncclUint8, r, c->comm, st);
                if (rr != ncclSuccess) { a->GroupEnd(); rptx_photons_destroy(full); return ctx_fail(ctx, std::string("ncclBroadcast: ") + a->GetErrorString(rr)); }
            }
            off += bytes;
        }
        CKN(a->GroupEnd());
    }
    full->emitted = local->emitted; full->rays = local->rays; full->emit_ms = local->emit_ms;
    CK(cudaStreamSynchronize(st));
    *out = full;
    return 0;
}

}  // extern "C"

// Wavefront path tracer state: structure-of-arrays path pool in HBM + ballot-compacted queues.
// One slot owns pixel (slot % n_pixels) and sample lane (slot / n_pixels); it walks its
// samples one after another (path regeneration), summing their radiance into its own f64
// accumulator, so the film needs no atomics and sums are deterministic.
#pragma once
#include "common.cuh"
#include "rng.cuh"

namespace rptx {

constexpr int kBlock = 128;

RPTX_D uint32_t lane_id() { return threadIdx.x & 31u; }
RPTX_D void prefetch_l2(const void* ptr) { asm volatile("prefetch.global.L2 [%0];" ::"l"(ptr)); }

// one atomic per warp for the run counters (a per-thread atomic on one address serialises)
RPTX_D void stat_add(unsigned long long* __restrict__ dst, unsigned long long v) {
    const unsigned active = __activemask();
    const unsigned total = __reduce_add_sync(active, (unsigned)v);  // per-thread counts are small
    if (lane_id() == (uint32_t)(__ffs(active) - 1) && total) atomicAdd(dst, (unsigned long long)total);
}

// same for counts that may exceed 32 bits per warp
RPTX_D void stat_add64(unsigned long long* __restrict__ dst, unsigned long long v) {
    const unsigned active = __activemask();
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(active, v, o);
    if (lane_id() == 0 && v) atomicAdd(dst, v);
}

struct PathPool {
    uint32_t n_slots, n_pixels, n_lanes, n_shadow;  // n_shadow = non-ambient, non-directional lights
    uint32_t stack_depth;                           // max_bounces + 1 (surface mode only)
    uint32_t rec_stride;                            // doubles per level record: 8, or 4 when max_bounces == 0 (only `direct` is kept)
    uint32_t record_mode;                           // surface mode: the vertex is assembled in its stack record (k_shade/k_shadow)
    uint32_t width, tiled;                          // tiled: slots walk the image in 8x4-pixel tiles (pool_pixel)
    // current segment
    double *ox, *oy, *oz, *dx, *dy, *dz;
    // closest hit of the segment
    double* ht;
    uint32_t* hobj;
    double *hnx, *hny, *hnz;
    // per-path bookkeeping
    uint64_t* rng;
    uint32_t* sample_j;  // index into this rank's sample list of the path in flight
    uint32_t* depth;
    uint32_t* flags;
    // vertex being assembled
    // medium mode, 64-byte records per slot (whole sectors instead of twelve 8-byte streams):
    double* mvtx;   // vertex being assembled: {direct.xyz, - | continuation weight f.xyz, -}
    double* mpath;  // forward accumulation: {throughput beta.xyz, - | radiance L.xyz, -}
    // NEE candidates, one 64-byte record per [light][slot]: {c.xyz, dist | d.xyz, -}.  Whole 32-byte sectors only: the
    // contribution (always written / read) in the first, the shadow ray (jobs that are traced) in the second.
    double* sjob;
    // per-level records for the clamped back-fold (surface mode): [n_slots][stack_depth][8]
    double* stack;
    // lane film: sum of finished samples of this slot in the current batch
    double *ax, *ay, *az;
};

// Pixel owned by a slot.  Wavefront films whose size allows it map 32 consecutive slots to an 8x4-pixel
// tile rather than to 32 pixels of one row: a warp's primary rays then span a squarer footprint (fewer
// warps straddle a silhouette, closer traversal lengths).  The RNG is keyed by the PIXEL, so the image
// does not depend on the mapping.
RPTX_HD uint32_t pool_pixel(const PathPool& p, uint32_t slot) {
    const uint32_t r = slot % p.n_pixels;
    if (!p.tiled) return r;
    const uint32_t tile = r >> 5, in = r & 31u, tiles_x = p.width >> 3;
    const uint32_t tx = tile % tiles_x, ty = tile / tiles_x;
    return (ty * 4u + (in >> 3)) * p.width + tx * 8u + (in & 7u);
}

enum PathFlags : uint32_t { kFlagTerminate = 1u, kFlagHasShadow = 2u };

RPTX_D void job_store_c(const PathPool& p, size_t o, d3 c, double dist) {
    double2* r = reinterpret_cast<double2*>(p.sjob + o * 8);
    r[0] = make_double2(c.x, c.y);
    r[1] = make_double2(c.z, dist);
}
RPTX_D void job_store_d(const PathPool& p, size_t o, d3 d) {
    double2* r = reinterpret_cast<double2*>(p.sjob + o * 8);
    r[2] = make_double2(d.x, d.y);
    r[3] = make_double2(d.z, 0.0);
}
RPTX_D d3 job_load_c(const PathPool& p, size_t o, double* dist) {
    const double2* r = reinterpret_cast<const double2*>(p.sjob + o * 8);
    const double2 a = r[0], b = r[1];
    *dist = b.y;
    return mk3(a.x, a.y, b.x);
}
RPTX_D d3 job_load_d(const PathPool& p, size_t o) {
    const double2* r = reinterpret_cast<const double2*>(p.sjob + o * 8);
    const double2 a = r[2], b = r[3];
    return mk3(a.x, a.y, b.x);
}

RPTX_D void rec_store(double* base, uint32_t slot, int half, d3 v) {
    double2* r = reinterpret_cast<double2*>(base + (size_t)slot * 8 + half * 4);
    r[0] = make_double2(v.x, v.y);
    r[1] = make_double2(v.z, 0.0);
}
RPTX_D d3 rec_load(const double* base, uint32_t slot, int half) {
    const double2* r = reinterpret_cast<const double2*>(base + (size_t)slot * 8 + half * 4);
    const double2 a = r[0], b = r[1];
    return mk3(a.x, a.y, b.x);
}

struct Queues {
    uint32_t* cur;        // slot ids to process this wave
    uint32_t* next;       // slot ids alive after this wave
    uint8_t* keep;        // [i] whether cur[i] stays in flight (k_connect); next = the kept entries of cur, order preserved
    uint32_t* counters;   // [0] cur count, [1] next count, [2] extend work head, [3] connect work head
    unsigned long long* stats;  // [0] paths started, [1] extend rays, [2] shadow rays
};

struct BatchParams {
    uint32_t width, height;
    uint32_t max_bounces;
    uint64_t seed;
    uint64_t batch_base;    // global index of the batch's first sample
    uint32_t iterations;    // samples per pixel in this batch (all ranks together)
    uint32_t shard_count, shard_index;
    uint32_t n_local;       // samples of this rank in this batch
};

}  // namespace rptx

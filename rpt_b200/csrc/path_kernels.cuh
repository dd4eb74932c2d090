// Wavefront kernels of the path tracer: generate -> extend -> shade (surface / medium scatter)
// -> connect (shadow/NEE resolve, vertex finalise, accumulate, regenerate).  Together they
// evaluate Renderer::trace_ray (src/renderer.rs:187-322) iteratively:
//  * surface branch (:286-321): per-level records (direct, f, pdf, |cos|) are kept in HBM and
//    folded back to front when the path ends, because each level's whole indirect term is
//    clamped per channel to FIREFLY_CLAMP (:311-313);
//  * medium branch (:188-285): linear, so radiance is accumulated forward with a throughput.
#pragma once
#include "path.cuh"
#include "shading.cuh"
#include "trace.cuh"

// resident blocks per SM asked of ptxas for the three heavy wavefront kernels (register cap = 65536 / (128 * n))
#ifndef RPTX_OCC_EXTEND
#define RPTX_OCC_EXTEND 4
#endif
#ifndef RPTX_OCC_SHADE
#define RPTX_OCC_SHADE 5
#endif
#ifndef RPTX_OCC_FINAL
#define RPTX_OCC_FINAL 6
#endif
#ifndef RPTX_SHADE_BLOCK
#define RPTX_SHADE_BLOCK 128
#endif
#ifndef RPTX_OCC_SHADOW
#define RPTX_OCC_SHADOW 6
#endif
// the small-scene variants (closest_hit_small: no traversal stack) are bound by the latency of dependent L1 hits at 16
// warps per SM: 7 / 8 blocks (72 / 64 registers, 90-190 B of spills) measured best (cornell extend 39.2 -> 33.9 ms, shadow 37.3 -> 28.0)
#ifndef RPTX_OCC_EXTEND_SMALL
#define RPTX_OCC_EXTEND_SMALL 7
#endif
#ifndef RPTX_OCC_SHADOW_SMALL
#define RPTX_OCC_SHADOW_SMALL 8
#endif

namespace rptx {

// warp-ballot compaction: append `slot` of every lane with pred to q, one atomic per warp
RPTX_D void queue_push(uint32_t* __restrict__ q, uint32_t* __restrict__ count, bool pred, uint32_t slot) {
    const unsigned active = __activemask();
    const unsigned mask = __ballot_sync(active, pred);
    if (mask == 0) return;
    const int leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if ((int)lane_id() == leader) base = atomicAdd(count, (uint32_t)__popc(mask));
    base = __shfl_sync(active, base, leader);
    if (pred) q[base + __popc(mask & ((1u << lane_id()) - 1u))] = slot;  // (queue_push is unused by the ordered-compaction wave loop)
}

template <bool MEDIUM>
RPTX_D void start_path(const PathPool& p, const BatchParams& bp, const Camera& cam, uint32_t slot, uint32_t j) {
    const uint32_t pixel = pool_pixel(p, slot);
    const uint32_t x = pixel % bp.width, y = pixel / bp.width;
    const uint64_t s = bp.batch_base + bp.shard_index + (uint64_t)j * bp.shard_count;
    Rng rng = Rng::make(bp.seed, kDomainCamera, pixel, s);
    d3 o, d;
    primary_ray(cam, bp.width, bp.height, x, y, rng, &o, &d);
    p.ox[slot] = o.x; p.oy[slot] = o.y; p.oz[slot] = o.z;
    p.dx[slot] = d.x; p.dy[slot] = d.y; p.dz[slot] = d.z;
    p.rng[slot] = rng.state;
    p.sample_j[slot] = j;
    p.depth[slot] = 0;
    p.flags[slot] = 0;
    if (MEDIUM) {
        rec_store(p.mpath, slot, 0, mk3(1.0, 1.0, 1.0));
        rec_store(p.mpath, slot, 1, mk3(0.0, 0.0, 0.0));
    }
}

template <bool MEDIUM>
__global__ void __launch_bounds__(kBlock) k_generate(PathPool p, Queues q, BatchParams bp, Camera cam) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    if (slot >= p.n_slots) return;
    p.ax[slot] = 0.0; p.ay[slot] = 0.0; p.az[slot] = 0.0;
    const uint32_t lane = slot / p.n_pixels;
    const bool active = lane < p.n_lanes && lane < bp.n_local;
    if (active) {  // the active slots are a prefix: the first queue is the identity, i.e. in slot order
        start_path<MEDIUM>(p, bp, cam, slot, lane);
        q.cur[slot] = slot;
    }
    if (slot == 0) q.counters[0] = p.n_pixels * min(p.n_lanes, bp.n_local);
    stat_add(&q.stats[0], active ? 1ull : 0ull);
}

// extend: persistent warps pull 32 rays at a time from the queue
template <bool SMALL>
__global__ void __launch_bounds__(kBlock, SMALL ? RPTX_OCC_EXTEND_SMALL : RPTX_OCC_EXTEND) k_extend(const __grid_constant__ SceneDev sc, PathPool p, Queues q) {
    const uint32_t count = q.counters[0];
    for (;;) {
        uint32_t base = 0;
        if (lane_id() == 0) base = atomicAdd(&q.counters[2], 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= count) break;
        const uint32_t i = base + lane_id();
        if (i < count) {
            const uint32_t slot = q.cur[i];
            RayD ray;
            ray.o = mk3(p.ox[slot], p.oy[slot], p.oz[slot]);
            ray.d = mk3(p.dx[slot], p.dy[slot], p.dz[slot]);
            Hit h;
            h.t = CUDART_INF;
            closest_hit_sel<kClosest, SMALL>(sc, ray, kEpsilon, 0.0, h);
            if (h.inst != kNoHit) {
                uint32_t prim, inst;
                const d3 n = hit_normal(sc, ray, h, &prim, &inst);
                p.ht[slot] = h.t;
                p.hobj[slot] = h.obj;
                p.hnx[slot] = n.x; p.hny[slot] = n.y; p.hnz[slot] = n.z;
            } else {
                p.ht[slot] = CUDART_INF;
                p.hobj[slot] = kNoHit;
            }
        }
    }
    if (threadIdx.x == 0 && blockIdx.x == 0) atomicAdd(&q.stats[1], (unsigned long long)count);
}

// NEE candidates of one vertex (Renderer::sample_lights / sample_lights_for_media).  Ambient
// lights are added to `direct` here; every other light hands a shadow job (k-th shadow-casting
// light, direction, distance, radiance if unoccluded) to `job`.  Directional lights have
// dist = +inf, which can never pass |t - dist| < 1e-12 (src/renderer.rs:396), so no job is
// emitted for them (their RNG draws, if any, still happen).
template <bool SURFACE, class Job>
RPTX_D void sample_lights_each(const SceneDev& sc, const Material* mat, const MediumDev* md, d3 pos, d3 n, d3 wo, d3 mc, Rng& rng,
                               d3* direct, Job&& job) {
    uint32_t k = 0;
    for (uint32_t li = 0; li < sc.n_lights; li++) {
        const Light& l = sc.lights[li];
        if (l.kind == 1) {
            const d3 amb = mk3(l.color[0], l.color[1], l.color[2]);
            *direct = *direct + cmul(amb, SURFACE ? mat_color(*mat) : mc);
            continue;
        }
        d3 intensity, wi;
        double dist;
        illuminate(sc, l, pos, rng, &intensity, &wi, &dist);
        if (l.kind == 2) continue;
        d3 c;
        if (SURFACE) {
            c = cmul(mat_bsdf(*mat, n, wo, wi), intensity) * dot(wi, n);
        } else {
            c = (md->sigma_s / md->extinction()) * cmul(intensity, mc) * md->phase;
        }
        job(k, wi, dist, c);
        k++;
    }
}

template <bool SURFACE>
RPTX_D void sample_lights_dev(const SceneDev& sc, const PathPool& p, uint32_t slot, const Material* mat, const MediumDev* md,
                              d3 pos, d3 n, d3 wo, d3 mc, Rng& rng, d3* direct) {
    sample_lights_each<SURFACE>(sc, mat, md, pos, n, wo, mc, rng, direct, [&](uint32_t k, d3 wi, double dist, d3 c) {
        const size_t o = (size_t)k * p.n_slots + slot;
        job_store_c(p, o, c, dist);
        // a candidate that contributes exactly zero is never traced (k_shadow): its ray need not be stored
        if (!(c.x == 0.0 && c.y == 0.0 && c.z == 0.0)) job_store_d(p, o, wi);
    });
}

// shade: the three kinds of vertex (medium collision, surface hit, miss) run different code, and a warp of 32
// consecutive queue entries holds all three (lampshade: 12 of 32 lanes live per instruction unsorted).  So a
// block takes a TILE of `rounds` x kShadeBlock consecutive queue entries, classifies them (the class needs t,
// the hit object and, with a medium, the free-flight draw), sorts the tile by class in shared memory (stable
// counting sort: ballot ranks, a 32-wide scan per class) and walks the sorted tile kShadeBlock entries at a
// time: only the two warp-rounds that straddle a class boundary are mixed.  Round 1 sorted kShadeBlock entries
// only (two or three mixed warps of four: 19.4 live lanes on lampshade).  The entries stay within one tile of
// their queue position, so the state accesses stay local; `rounds` is chosen by the host from the wave's size
// (a small wave keeps one round, so that it still fills the SMs).
constexpr int kShadeBlock = RPTX_SHADE_BLOCK;
#ifndef RPTX_SHADE_ROUNDS
#define RPTX_SHADE_ROUNDS 8
#endif
constexpr int kShadeRoundsMax = RPTX_SHADE_ROUNDS;
constexpr int kShadeTile = kShadeBlock * kShadeRoundsMax;
constexpr int kShadeWarps = kShadeBlock / 32;
static_assert(kShadeWarps * kShadeRoundsMax <= 32 && kShadeBlock >= 96, "one 32-wide scan per class over the (round, warp) counts");

template <bool MEDIUM>
__global__ void __launch_bounds__(kShadeBlock, RPTX_OCC_SHADE) k_shade(const __grid_constant__ SceneDev sc, PathPool p, Queues q, BatchParams bp,
                                                          MediumDev md, uint32_t rounds) {
    __shared__ uint32_t s_slot[kShadeTile];
    __shared__ uint16_t s_perm[kShadeTile];                   // sorted position -> tile position
    __shared__ uint8_t s_cls[kShadeTile];
    __shared__ uint32_t s_cnt[3][32];                         // per class: entries of (round, warp), then their exclusive scan
    __shared__ uint32_t s_tot[3];
    __shared__ double s_dsamp[MEDIUM ? kShadeTile : 1];       // the free-flight draw travels with the entry
    __shared__ uint64_t s_rng[MEDIUM ? kShadeTile : 1];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    const uint32_t lt = (1u << lane) - 1u;
    {
        const uint32_t count = q.counters[0];
        const uint32_t tile_base = blockIdx.x * (rounds * kShadeBlock);
        if (threadIdx.x < 96) s_cnt[warp][lane] = 0;
        __syncthreads();
        for (uint32_t r = 0; r < rounds; r++) {
            const uint32_t u = r * kShadeBlock + threadIdx.x;
            const uint32_t i = tile_base + u;
            uint32_t cls = 3;  // 0 medium collision, 1 surface, 2 miss, 3 no entry
            if (i < count) {
                const uint32_t slot0 = q.cur[i];
#ifndef RPTX_NO_SHADE_PREFETCH
                // the entry is dealt to another thread of this block below: start its state on the way to L2 now, so
                // that the loads after the re-deal do not add a third DRAM round trip (queue -> hit object -> state)
                prefetch_l2(p.ox + slot0); prefetch_l2(p.oy + slot0); prefetch_l2(p.oz + slot0);
                prefetch_l2(p.dx + slot0); prefetch_l2(p.dy + slot0); prefetch_l2(p.dz + slot0);
                prefetch_l2(p.hnx + slot0); prefetch_l2(p.hny + slot0); prefetch_l2(p.hnz + slot0);
                prefetch_l2(p.depth + slot0);
                if (!MEDIUM) { prefetch_l2(p.rng + slot0); prefetch_l2(p.ht + slot0); }
#endif
                const bool hit0 = p.hobj[slot0] != kNoHit;
                cls = hit0 ? 1u : 2u;
                if (MEDIUM) {
                    Rng r0;
                    r0.state = p.rng[slot0];
                    const double ds = medium_sample_d(md, r0);
                    if (ds < (hit0 ? p.ht[slot0] : kBackgroundDist)) cls = 0;
                    s_dsamp[u] = ds;
                    s_rng[u] = r0.state;
                }
                s_slot[u] = slot0;
            }
            s_cls[u] = (uint8_t)cls;
#pragma unroll
            for (uint32_t c = 0; c < 3; c++) {
                const unsigned m = __ballot_sync(0xffffffffu, cls == c);
                if (lane == 0) s_cnt[c][r * kShadeWarps + warp] = (uint32_t)__popc(m);
            }
        }
        __syncthreads();
        if (warp < 3) {  // warp c: exclusive scan of class c's counts in (round, warp) order
            const uint32_t v = s_cnt[warp][lane];
            uint32_t incl = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t up = __shfl_up_sync(0xffffffffu, incl, o);
                if ((int)lane >= o) incl += up;
            }
            s_cnt[warp][lane] = incl - v;
            if (lane == 31) s_tot[warp] = incl;
        }
        __syncthreads();
        for (uint32_t r = 0; r < rounds; r++) {
            const uint32_t u = r * kShadeBlock + threadIdx.x;
            const uint32_t cls = s_cls[u];
            uint32_t pos = 0;
#pragma unroll
            for (uint32_t c = 0; c < 3; c++) {
                const unsigned m = __ballot_sync(0xffffffffu, cls == c);
                if (cls == c) pos = s_cnt[c][r * kShadeWarps + warp] + (uint32_t)__popc(m & lt) + (c > 0 ? s_tot[0] : 0u) + (c > 1 ? s_tot[1] : 0u);
            }
            RPTX_ASSERT(cls >= 3 || (pos < rounds * kShadeBlock && u < (uint32_t)kShadeTile));
            if (cls < 3) s_perm[pos] = (uint16_t)u;
        }
        __syncthreads();
    }
    const uint32_t total = s_tot[0] + s_tot[1] + s_tot[2];
    for (uint32_t j = threadIdx.x; j < total; j += kShadeBlock) {
    const uint32_t u = s_perm[j];
    const uint32_t slot = s_slot[u];
    double dsamp = 0.0;
    Rng rng;
    if (MEDIUM) { dsamp = s_dsamp[u]; rng.state = s_rng[u]; }
    const d3 o = mk3(p.ox[slot], p.oy[slot], p.oz[slot]);
    const d3 d = mk3(p.dx[slot], p.dy[slot], p.dz[slot]);
    const double t = p.ht[slot];
    const uint32_t obj = p.hobj[slot];
    const uint32_t depth = p.depth[slot];
    if (!MEDIUM) rng.state = p.rng[slot];
    uint32_t flags = 0;
    d3 direct = mk3(0, 0, 0);
    d3 rec_f = mk3(0, 0, 0);
    double rec_pdf = 0.0, rec_ac = 0.0;

    const bool hit = obj != kNoHit;
    const double max_dist = hit ? t : kBackgroundDist;

    if (MEDIUM && dsamp < max_dist) {
        // medium vertex, src/renderer.rs:243-283
        const d3 wo = -normalize(d);
        const d3 collision = o + dsamp * d;
        const d3 mc = md.color(collision);
        if (depth == 0) direct = md.emission * mc;
        sample_lights_dev<false>(sc, p, slot, nullptr, &md, collision, mk3(0, 0, 0), wo, mc, rng, &direct);
        if (p.n_shadow) flags |= kFlagHasShadow;
        p.ox[slot] = collision.x; p.oy[slot] = collision.y; p.oz[slot] = collision.z;
        if (rng.gen() < kRrMedium) {
            d3 wi;
            medium_sample_ph(rng, &wi);
            const double ph_p = md.phase;
            const double sa = md.sigma_s / md.extinction();
            // indirect = ((sa * L) / ph_p) (.) mc * phase / rr_p
            rec_f = mk3(((sa / ph_p) * mc.x * md.phase) / kRrMedium, ((sa / ph_p) * mc.y * md.phase) / kRrMedium,
                        ((sa / ph_p) * mc.z * md.phase) / kRrMedium);
            p.dx[slot] = wi.x; p.dy[slot] = wi.y; p.dz[slot] = wi.z;
        } else {
            flags |= kFlagTerminate;
        }
    } else if (!hit) {
        // miss: environment (medium mode: only reached when d >= background_dist)
        direct = env_color(sc, d);
        flags |= kFlagTerminate;
    } else {
        // surface vertex, src/renderer.rs:207-236 (medium mode) / :289-319
        const d3 n = mk3(p.hnx[slot], p.hny[slot], p.hnz[slot]);
        const d3 pos = o + t * d;
        const Material mat = sc.materials[obj];
        const d3 wo = -normalize(d);
        if (depth == 0) direct = mat_emittance(mat) * mat_color(mat);
        sample_lights_dev<true>(sc, p, slot, &mat, nullptr, pos, n, wo, mk3(0, 0, 0), rng, &direct);
        if (p.n_shadow) flags |= kFlagHasShadow;
        p.ox[slot] = pos.x; p.oy[slot] = pos.y; p.oz[slot] = pos.z;
        bool go = MEDIUM ? (rng.gen() < kRrMedium) : (depth < bp.max_bounces);
        if (go) {
            d3 wi;
            double pdf;
            if (mat_sample_f(mat, n, wo, rng, &wi, &pdf)) {
                const d3 f = mat_bsdf(mat, n, wo, wi);
                const double ac = fabs(dot(wi, n));
                if (MEDIUM) {
                    const double s = 1.0 / (pdf * kRrMedium);
                    rec_f = mk3(s * f.x * ac, s * f.y * ac, s * f.z * ac);
                } else {
                    rec_f = f; rec_pdf = pdf; rec_ac = ac;
                }
                p.dx[slot] = wi.x; p.dy[slot] = wi.y; p.dz[slot] = wi.z;
            } else {
                go = false;
            }
        }
        if (!go) flags |= kFlagTerminate;
    }
    if (MEDIUM) {
        rec_store(p.mvtx, slot, 0, direct);
        if (!(flags & kFlagTerminate)) rec_store(p.mvtx, slot, 1, rec_f);
    } else {
        // surface mode assembles the vertex directly in its level record of the clamp stack
        // ([slot][depth]{direct, f, pdf, |cos|}, 64 B): k_shadow adds the visible light samples to `direct`
        // there and k_connect has nothing to copy.  A vertex that ends the path only needs `direct`.
        double2* r = reinterpret_cast<double2*>(p.stack + ((size_t)slot * p.stack_depth + depth) * p.rec_stride);
        r[0] = make_double2(direct.x, direct.y);
        r[1] = make_double2(direct.z, rec_f.x);
        if (!(flags & kFlagTerminate)) {
            r[2] = make_double2(rec_f.y, rec_f.z);
            r[3] = make_double2(rec_pdf, rec_ac);
        }
    }
    p.flags[slot] = flags;
    p.rng[slot] = rng.state;
    }
}

// shadow: NEE visibility of the vertex's light samples (nee_visible, trace.cuh).  Passing
// candidates are added to the vertex's direct light in light order.  A candidate whose
// contribution is exactly zero (light sample facing away, bsdf 0) is not traced: adding it
// cannot change the sum.  About half of the Cornell / lampshade vertices have no job, and a warp of
// 32 consecutive queue entries ran its traversal with 16 live lanes.  So a block pulls kShadowChunk
// entries at a time, keeps the ones with a job (ballot ranks + per-warp counts, queue order kept) and
// deals them to its threads densely: every warp but the last runs with 32 live lanes.
// PROBE: the scene has point lights (probed at their sample first); all lights take that path then.
constexpr int kShadowPer = 2;                        // queue entries classified per thread and pull
constexpr int kShadowChunk = kShadowPer * kBlock;

template <bool PROBE, bool SMALL>
__global__ void __launch_bounds__(kBlock, SMALL ? RPTX_OCC_SHADOW_SMALL : RPTX_OCC_SHADOW) k_shadow_block(const __grid_constant__ SceneDev sc, PathPool p, Queues q) {
    __shared__ uint32_t s_base;
    __shared__ uint32_t s_cnt[kShadowPer * (kBlock / 32)];
    __shared__ uint32_t s_slot[kShadowChunk];
    __shared__ double s_c[3][kShadowChunk];
    const uint32_t count = q.counters[0];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    unsigned long long my_queries = 0, my_traced = 0, my_occl = 0;
    for (;;) {
        __syncthreads();  // the previous chunk's readers of s_base / s_slot are done
        if (threadIdx.x == 0) s_base = atomicAdd(&q.counters[3], (uint32_t)kShadowChunk);
        __syncthreads();
        const uint32_t base = s_base;
        if (base >= count) break;
        uint32_t slot[kShadowPer];
        bool has[kShadowPer];
#pragma unroll
        for (int r = 0; r < kShadowPer; r++) {
            const uint32_t i = base + (uint32_t)r * kBlock + threadIdx.x;
            slot[r] = i < count ? q.cur[i] : 0u;
            has[r] = i < count && (p.flags[slot[r]] & kFlagHasShadow);
        }
        for (uint32_t k = 0; k < p.n_shadow; k++) {  // light order; the adds of light k are visible before light k + 1 starts
            if (k) __syncthreads();
            uint32_t rank[kShadowPer];
            bool job[kShadowPer];
            d3 c[kShadowPer];
#pragma unroll
            for (int r = 0; r < kShadowPer; r++) {
                c[r] = mk3(0, 0, 0);
                if (has[r]) {
                    const size_t off = (size_t)k * p.n_slots + slot[r];
                    double dist_unused;
                    c[r] = job_load_c(p, off, &dist_unused);
                    my_queries++;
                }
                job[r] = has[r] && !(c[r].x == 0.0 && c[r].y == 0.0 && c[r].z == 0.0);
                const unsigned m = __ballot_sync(0xffffffffu, job[r]);
                rank[r] = (uint32_t)__popc(m & ((1u << lane) - 1u));
                if (lane == 0) s_cnt[r * (kBlock / 32) + warp] = (uint32_t)__popc(m);
            }
            __syncthreads();
            uint32_t total = 0, before[kShadowPer];
#pragma unroll
            for (int r = 0; r < kShadowPer; r++) before[r] = 0;
#pragma unroll
            for (uint32_t w = 0; w < kShadowPer * (kBlock / 32); w++) {
                const uint32_t n = s_cnt[w];
#pragma unroll
                for (int r = 0; r < kShadowPer; r++)
                    if (w < r * (kBlock / 32) + warp) before[r] += n;
                total += n;
            }
#pragma unroll
            for (int r = 0; r < kShadowPer; r++)
                if (job[r]) {
                    const uint32_t at = before[r] + rank[r];
                    RPTX_ASSERT(at < (uint32_t)kShadowChunk && slot[r] < p.n_slots);
                    s_slot[at] = slot[r];
                    s_c[0][at] = c[r].x; s_c[1][at] = c[r].y; s_c[2][at] = c[r].z;
                }
            __syncthreads();
            for (uint32_t j = threadIdx.x; j < total; j += kBlock) {
                const uint32_t js = s_slot[j];
                const d3 cj = mk3(s_c[0][j], s_c[1][j], s_c[2][j]);
                const size_t off = (size_t)k * p.n_slots + js;
                RayD ray;
                ray.o = mk3(p.ox[js], p.oy[js], p.oz[js]);
                ray.d = job_load_d(p, off);
                double dist;
                job_load_c(p, off, &dist);
                my_traced++;
                if (nee_visible<PROBE, SMALL>(sc, ray, dist, &my_occl)) {
                    // direct += contribution, in light order like the reference's loop
                    if (p.record_mode) {
                        double2* rec = reinterpret_cast<double2*>(p.stack + ((size_t)js * p.stack_depth + p.depth[js]) * p.rec_stride);
                        const double2 a = rec[0], b = rec[1];
                        rec[0] = make_double2(a.x + cj.x, a.y + cj.y);
                        rec[1] = make_double2(b.x + cj.z, b.y);
                    } else {
                        rec_store(p.mvtx, js, 0, rec_load(p.mvtx, js, 0) + cj);
                    }
                }
            }
        }
    }
    stat_add64(&q.stats[2], my_queries);
    stat_add64(&q.stats[3], my_traced);
    stat_add64(&q.stats[4], my_occl);
}

// Warp-level form of the same idea, without block barriers (they were 4 of 15 stall cycles per issue in the block
// form on the small scenes; used there: cornell shadow 28.0 -> 23.7 ms, lampshade 50.4 -> 45.5; the BVH scenes, whose
// point-light probes end at the root, keep the block form: 2.6 against 4.8 ms): a warp keeps a FIFO of pending jobs in shared memory, appends the jobs of each 32 entries it pulls (ballot
// ranks, queue order kept) and traces 32 at a time as soon as it holds that many; the remainder (< 32) is carried to
// the next pull and flushed at the end.  With several shadow-casting lights the FIFO is drained after every light,
// so that a vertex's contributions are still added in light order by one lane at a time.
template <bool PROBE, bool SMALL>
__global__ void __launch_bounds__(kBlock, SMALL ? RPTX_OCC_SHADOW_SMALL : RPTX_OCC_SHADOW) k_shadow_warp(const __grid_constant__ SceneDev sc, PathPool p, Queues q) {
    constexpr int kW = kBlock / 32;
    __shared__ uint32_t s_slot[kW][64];
    __shared__ double s_c[kW][3][64];
    const uint32_t count = q.counters[0];
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31u;
    unsigned long long my_queries = 0, my_traced = 0, my_occl = 0;
    uint32_t fill = 0;  // warp-uniform
    // trace the first n (<= 32) pending jobs of light k, then move the rest of the FIFO to its front
    auto run = [&](uint32_t n, uint32_t k) {
        if (lane < n) {
            const uint32_t js = s_slot[warp][lane];
            const d3 cj = mk3(s_c[warp][0][lane], s_c[warp][1][lane], s_c[warp][2][lane]);
            const size_t off = (size_t)k * p.n_slots + js;
            RayD ray;
            ray.o = mk3(p.ox[js], p.oy[js], p.oz[js]);
            ray.d = job_load_d(p, off);
            double dist;
            job_load_c(p, off, &dist);
            my_traced++;
            if (nee_visible<PROBE, SMALL>(sc, ray, dist, &my_occl)) {
                // direct += contribution, in light order like the reference's loop
                if (p.record_mode) {
                    double2* rec = reinterpret_cast<double2*>(p.stack + ((size_t)js * p.stack_depth + p.depth[js]) * p.rec_stride);
                    const double2 a = rec[0], b = rec[1];
                    rec[0] = make_double2(a.x + cj.x, a.y + cj.y);
                    rec[1] = make_double2(b.x + cj.z, b.y);
                } else {
                    rec_store(p.mvtx, js, 0, rec_load(p.mvtx, js, 0) + cj);
                }
            }
        }
        __syncwarp();
        const uint32_t rest = fill - n;  // < 32
        uint32_t ts = 0;
        double t0 = 0.0, t1 = 0.0, t2 = 0.0;
        if (lane < rest) { ts = s_slot[warp][32 + lane]; t0 = s_c[warp][0][32 + lane]; t1 = s_c[warp][1][32 + lane]; t2 = s_c[warp][2][32 + lane]; }
        __syncwarp();
        if (lane < rest) { s_slot[warp][lane] = ts; s_c[warp][0][lane] = t0; s_c[warp][1][lane] = t1; s_c[warp][2][lane] = t2; }
        __syncwarp();
        fill = rest;
    };
    // one loop with ONE call site of run() (nee_visible is inlined there; three copies cost the big-scene variant 70 %)
    bool exhausted = false, drain = false;
    uint32_t k = p.n_shadow, kcur = 0, slot = 0;
    bool has = false;
    for (;;) {
        if (!drain && !exhausted && fill < 32u) {
            if (k == p.n_shadow) {  // next 32 queue entries
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(&q.counters[3], 32u);
                base = __shfl_sync(0xffffffffu, base, 0);
                if (base >= count) {
                    exhausted = true;
                } else {
                    const uint32_t i = base + lane;
                    slot = i < count ? q.cur[i] : 0u;
                    has = i < count && (p.flags[slot] & kFlagHasShadow);
                    k = 0;
                }
            }
            if (!exhausted) {  // append the entries' jobs of light k
                d3 c = mk3(0, 0, 0);
                if (has) {
                    const size_t off = (size_t)k * p.n_slots + slot;
                    double dist_unused;
                    c = job_load_c(p, off, &dist_unused);
                    my_queries++;
                }
                const bool job = has && !(c.x == 0.0 && c.y == 0.0 && c.z == 0.0);
                const unsigned m = __ballot_sync(0xffffffffu, job);
                if (job) {
                    const uint32_t at = fill + (uint32_t)__popc(m & ((1u << lane) - 1u));
                    RPTX_ASSERT(at < 64u && slot < p.n_slots);
                    s_slot[warp][at] = slot;
                    s_c[warp][0][at] = c.x; s_c[warp][1][at] = c.y; s_c[warp][2][at] = c.z;
                }
                fill += (uint32_t)__popc(m);
                __syncwarp();
                kcur = k;
                k++;
                drain = p.n_shadow > 1;  // several lights: finish this light's jobs before the next light's are queued
            }
        }
        if (exhausted) drain = true;
        if (fill >= 32u || (drain && fill > 0)) run(fill < 32u ? fill : 32u, kcur);
        if (fill == 0) {
            if (exhausted) break;
            drain = false;
        }
    }
    stat_add64(&q.stats[2], my_queries);
    stat_add64(&q.stats[3], my_traced);
    stat_add64(&q.stats[4], my_occl);
}

// connect: finalise the vertex; finished paths are folded (surface mode), added to the slot's
// lane film and the slot regenerates its next sample.  Returns whether the slot stays in flight.
// The queue is then compacted IN ORDER (cub::DeviceSelect::Flagged on q.keep, rptx.cu): it starts
// as the identity and a regenerated path keeps its slot, so every wave's queue is sorted by slot.
// A warp's 32 entries are then as close in memory as the survivors allow (coalesced SoA access)
// and belong to neighbouring pixels.  An atomic append per warp (the previous scheme) shuffled
// the warps' chunks every wave: k_shade was 1.6x slower after a few regenerations.
template <bool MEDIUM>
RPTX_D bool connect_vertex(const PathPool& p, const BatchParams& bp, const Camera& cam, uint32_t slot, d3 direct, unsigned long long* n_started) {
    const uint32_t flags = p.flags[slot];
    const uint32_t depth = p.depth[slot];
    const bool finished = (flags & kFlagTerminate) != 0;
    d3 L;
    if (MEDIUM) {
        const d3 beta = rec_load(p.mpath, slot, 0);
        L = rec_load(p.mpath, slot, 1) + cmul(beta, direct);
        if (!finished) {
            const d3 f = rec_load(p.mvtx, slot, 1);
            rec_store(p.mpath, slot, 1, L);
            rec_store(p.mpath, slot, 0, mk3(beta.x * f.x, beta.y * f.y, beta.z * f.z));
        }
    } else {
        // per-slot stack, one 64-byte record per level: [slot][level]{direct, f, pdf, |cos|}.  Only the
        // few lanes whose path ends fold, so a slot's levels are contiguous (4 x double2 per level).
        // (k_shade assembled this level's record, k_shadow added the visible light samples to it)
        const double2* st = reinterpret_cast<const double2*>(p.stack + ((size_t)slot * p.stack_depth) * p.rec_stride);
        if (finished) {
            const size_t lvl = p.rec_stride >> 1;  // double2 per level
            const double2 d0 = st[(size_t)depth * lvl], d1 = st[(size_t)depth * lvl + 1];
            L = mk3(d0.x, d0.y, d1.x);
            for (int j = (int)depth - 1; j >= 0; j--) {
                const double2* r = st + (size_t)j * lvl;
                const double2 a = r[0], b = r[1], c = r[2], e = r[3];
                const d3 f = mk3(b.y, c.x, c.y);
                const d3 ind = 1.0 / e.x * cmul(f, L) * e.y;
                L = mk3(a.x + fmin(ind.x, kFireflyClamp), a.y + fmin(ind.y, kFireflyClamp), b.x + fmin(ind.z, kFireflyClamp));
            }
        }
    }
    if (finished) {
        p.ax[slot] += L.x; p.ay[slot] += L.y; p.az[slot] += L.z;
        const uint32_t jn = p.sample_j[slot] + p.n_lanes;
        if (jn < bp.n_local) {
            start_path<MEDIUM>(p, bp, cam, slot, jn);
            (*n_started)++;
            return true;
        }
        return false;
    }
    p.depth[slot] = depth + 1;
    return true;
}

template <bool MEDIUM>
__global__ void __launch_bounds__(kBlock) k_connect(PathPool p, Queues q, BatchParams bp, Camera cam) {
    const uint32_t count = q.counters[0];
    unsigned long long my_started = 0;
    {
        const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
        bool alive = false;
        uint32_t slot = 0;
        if (i < count) {
            slot = q.cur[i];
            const d3 direct = MEDIUM ? rec_load(p.mvtx, slot, 0) : mk3(0, 0, 0);
            alive = connect_vertex<MEDIUM>(p, bp, cam, slot, direct, &my_started);
            q.keep[i] = alive ? 1 : 0;
        } else if (i < p.n_slots) {
            q.keep[i] = 0;  // between the live count and the host's (lagged) bound: nothing to keep
        }
    }
    stat_add(&q.stats[0], my_started);
}

// end of wave: next becomes cur (pointers are swapped on the host), heads reset
__global__ void k_wave_end(Queues q) {
    q.counters[0] = q.counters[1];
    q.counters[1] = 0;
    q.counters[2] = 0;
    q.counters[3] = 0;
}

// ---- fused path kernel -------------------------------------------------------------------
// For scenes whose whole accelerator sits in L1 (a dozen analytic objects: Cornell, lampshade)
// the wavefront's per-vertex state round trip through HBM (~650 B) and its five launches per
// wave cost more than the divergence they remove.  Here one thread owns a (pixel, lane) slot
// and walks its samples' whole paths in registers: extend -> shade -> NEE -> next vertex,
// regenerating the next sample in place.  Same arithmetic and RNG draw order as the wavefront
// kernels; NEE contributions are added in the reference's light order.
constexpr int kFusedMaxDepth = 16;

struct FusedArgs {
    Camera cam;
    BatchParams bp;
    uint32_t n_pixels, n_lanes;
    double *ax, *ay, *az;  // [n_lanes][n_pixels] lane sums of the batch
};

struct PathCounters {
    unsigned long long extend, queries, traced, occl;
};

template <bool SURFACE>
RPTX_D void nee_inline(const SceneDev& sc, const Material* mat, const MediumDev* md, d3 pos, d3 n, d3 wo, d3 mc, Rng& rng, d3* direct,
                       PathCounters& ctr) {
    sample_lights_each<SURFACE>(sc, mat, md, pos, n, wo, mc, rng, direct, [&](uint32_t k, d3 wi, double dist, d3 c) {
        ctr.queries++;
        if (c.x == 0.0 && c.y == 0.0 && c.z == 0.0) return;
        RayD sr;
        sr.o = pos;
        sr.d = wi;
        ctr.traced++;
        if (sc.shadow_probe_mask ? nee_visible<true>(sc, sr, dist, &ctr.occl) : nee_visible<false>(sc, sr, dist, &ctr.occl)) *direct = *direct + c;
    });
}

// One-vertex paths (surface mode with max_bounces == 0 - the reference's builder default, examples/teapot.rs and
// fractal_teapots.rs): the vertex k_extend found ends the path.  Shading, NEE visibility, the film add and the
// regeneration of the slot's next sample then run in ONE kernel and nothing of the vertex goes through HBM
// (k_shade + k_shadow + k_connect write and re-read ~400 B per vertex for it).  Same arithmetic, same RNG draws and
// the same light order as the three kernels.
template <bool PROBE, bool SMALL>
__global__ void __launch_bounds__(kBlock, RPTX_OCC_FINAL) k_shade_final(const __grid_constant__ SceneDev sc, PathPool p, Queues q, BatchParams bp, Camera cam) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long my_started = 0, my_queries = 0, my_traced = 0, my_occl = 0;
    if (i < q.counters[0]) {
        const uint32_t slot = q.cur[i];
        const d3 o = mk3(p.ox[slot], p.oy[slot], p.oz[slot]);
        const d3 d = mk3(p.dx[slot], p.dy[slot], p.dz[slot]);
        const uint32_t obj = p.hobj[slot];
        Rng rng;
        rng.state = p.rng[slot];
        d3 direct;
        if (obj == kNoHit) {
            direct = env_color(sc, d);
        } else {
            const double t = p.ht[slot];
            const d3 n = mk3(p.hnx[slot], p.hny[slot], p.hnz[slot]);
            const d3 pos = o + t * d;
            const Material mat = sc.materials[obj];
            const d3 wo = -normalize(d);
            direct = mat_emittance(mat) * mat_color(mat);  // depth 0, src/renderer.rs:295-299
            sample_lights_each<true>(sc, &mat, nullptr, pos, n, wo, mk3(0, 0, 0), rng, &direct, [&](uint32_t, d3 wi, double dist, d3 c) {
                my_queries++;
                if (c.x == 0.0 && c.y == 0.0 && c.z == 0.0) return;
                RayD sr;
                sr.o = pos;
                sr.d = wi;
                my_traced++;
                if (nee_visible<PROBE, SMALL>(sc, sr, dist, &my_occl)) direct = direct + c;
            });
        }
        p.ax[slot] += direct.x; p.ay[slot] += direct.y; p.az[slot] += direct.z;
        const uint32_t jn = p.sample_j[slot] + p.n_lanes;
        bool alive = false;
        if (jn < bp.n_local) {
            start_path<false>(p, bp, cam, slot, jn);
            my_started++;
            alive = true;
        }
        q.keep[i] = alive ? 1 : 0;
    } else if (i < p.n_slots) {
        q.keep[i] = 0;  // between the live count and the host's (lagged) bound: nothing to keep
    }
    stat_add(&q.stats[0], my_started);
    stat_add64(&q.stats[2], my_queries);
    stat_add64(&q.stats[3], my_traced);
    stat_add64(&q.stats[4], my_occl);
}

// trace_ray, surface branch (src/renderer.rs:286-321)
RPTX_D d3 path_radiance_surface(const SceneDev& sc, RayD ray, Rng& rng, uint32_t max_bounces, PathCounters& ctr) {
    double st_dir[kFusedMaxDepth][3], st_f[kFusedMaxDepth][3], st_pdf[kFusedMaxDepth], st_ac[kFusedMaxDepth];
    d3 direct = mk3(0, 0, 0);
    uint32_t depth = 0;
    for (;; depth++) {
        Hit h;
        h.t = CUDART_INF;
        closest_hit<kClosest>(sc, ray, kEpsilon, 0.0, h);
        ctr.extend++;
        if (h.inst == kNoHit) { direct = env_color(sc, ray.d); break; }
        uint32_t prim, inst;
        const d3 n = hit_normal(sc, ray, h, &prim, &inst);
        const d3 pos = ray.o + h.t * ray.d;
        const Material mat = sc.materials[h.obj];
        const d3 wo = -normalize(ray.d);
        direct = depth == 0 ? mat_emittance(mat) * mat_color(mat) : mk3(0, 0, 0);
        nee_inline<true>(sc, &mat, nullptr, pos, n, wo, mk3(0, 0, 0), rng, &direct, ctr);
        if (!(depth < max_bounces)) break;
        d3 wi;
        double pdf;
        if (!mat_sample_f(mat, n, wo, rng, &wi, &pdf)) break;
        const d3 f = mat_bsdf(mat, n, wo, wi);
        st_dir[depth][0] = direct.x; st_dir[depth][1] = direct.y; st_dir[depth][2] = direct.z;
        st_f[depth][0] = f.x; st_f[depth][1] = f.y; st_f[depth][2] = f.z;
        st_pdf[depth] = pdf;
        st_ac[depth] = fabs(dot(wi, n));
        ray.o = pos;
        ray.d = wi;
    }
    d3 L = direct;
    for (int j = (int)depth - 1; j >= 0; j--) {
        const d3 f = mk3(st_f[j][0], st_f[j][1], st_f[j][2]);
        const d3 ind = 1.0 / st_pdf[j] * cmul(f, L) * st_ac[j];
        L = mk3(st_dir[j][0] + fmin(ind.x, kFireflyClamp), st_dir[j][1] + fmin(ind.y, kFireflyClamp), st_dir[j][2] + fmin(ind.z, kFireflyClamp));
    }
    return L;
}

// trace_ray, medium branch (src/renderer.rs:188-285)
RPTX_D d3 path_radiance_medium(const SceneDev& sc, const MediumDev& md, RayD ray, Rng& rng, PathCounters& ctr) {
    d3 beta = mk3(1, 1, 1), L = mk3(0, 0, 0);
    for (uint32_t depth = 0;; depth++) {
        const double dsamp = medium_sample_d(md, rng);
        Hit h;
        h.t = CUDART_INF;
        closest_hit<kClosest>(sc, ray, kEpsilon, 0.0, h);
        ctr.extend++;
        const bool hit = h.inst != kNoHit;
        const double max_dist = hit ? h.t : kBackgroundDist;
        d3 direct = mk3(0, 0, 0), fc = mk3(0, 0, 0);
        bool go = false;
        if (dsamp < max_dist) {
            const d3 wo = -normalize(ray.d);
            const d3 collision = ray.o + dsamp * ray.d;
            const d3 mc = md.color(collision);
            if (depth == 0) direct = md.emission * mc;
            nee_inline<false>(sc, nullptr, &md, collision, mk3(0, 0, 0), wo, mc, rng, &direct, ctr);
            ray.o = collision;
            if (rng.gen() < kRrMedium) {
                d3 wi;
                medium_sample_ph(rng, &wi);
                const double ph_p = md.phase;
                const double sa = md.sigma_s / md.extinction();
                fc = mk3(((sa / ph_p) * mc.x * md.phase) / kRrMedium, ((sa / ph_p) * mc.y * md.phase) / kRrMedium,
                         ((sa / ph_p) * mc.z * md.phase) / kRrMedium);
                ray.d = wi;
                go = true;
            }
        } else if (!hit) {
            direct = env_color(sc, ray.d);
        } else {
            uint32_t prim, inst;
            const d3 n = hit_normal(sc, ray, h, &prim, &inst);
            const d3 pos = ray.o + h.t * ray.d;
            const Material mat = sc.materials[h.obj];
            const d3 wo = -normalize(ray.d);
            if (depth == 0) direct = mat_emittance(mat) * mat_color(mat);
            nee_inline<true>(sc, &mat, nullptr, pos, n, wo, mk3(0, 0, 0), rng, &direct, ctr);
            ray.o = pos;
            if (rng.gen() < kRrMedium) {
                d3 wi;
                double pdf;
                if (mat_sample_f(mat, n, wo, rng, &wi, &pdf)) {
                    const d3 f = mat_bsdf(mat, n, wo, wi);
                    const double ac = fabs(dot(wi, n));
                    const double s = 1.0 / (pdf * kRrMedium);
                    fc = mk3(s * f.x * ac, s * f.y * ac, s * f.z * ac);
                    ray.d = wi;
                    go = true;
                }
            }
        }
        L = L + cmul(beta, direct);
        if (!go) return L;
        beta = mk3(beta.x * fc.x, beta.y * fc.y, beta.z * fc.z);
    }
}

template <bool MEDIUM>
__global__ void __launch_bounds__(kBlock) k_path_fused(const __grid_constant__ SceneDev sc, const MediumDev md,
                                                        const __grid_constant__ FusedArgs fa, unsigned long long* __restrict__ stats) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    PathCounters ctr = {0, 0, 0, 0};
    unsigned long long my_paths = 0;
    if (slot < fa.n_pixels * fa.n_lanes) {
        const uint32_t pixel = slot % fa.n_pixels, lane = slot / fa.n_pixels;
        const uint32_t x = pixel % fa.bp.width, y = pixel / fa.bp.width;
        d3 sum = mk3(0, 0, 0);
        for (uint32_t j = lane; j < fa.bp.n_local; j += fa.n_lanes) {
            const uint64_t s = fa.bp.batch_base + fa.bp.shard_index + (uint64_t)j * fa.bp.shard_count;
            Rng rng = Rng::make(fa.bp.seed, kDomainCamera, pixel, s);
            RayD ray;
            primary_ray(fa.cam, fa.bp.width, fa.bp.height, x, y, rng, &ray.o, &ray.d);
            const d3 L = MEDIUM ? path_radiance_medium(sc, md, ray, rng, ctr) : path_radiance_surface(sc, ray, rng, fa.bp.max_bounces, ctr);
            sum = sum + L;
            my_paths++;
        }
        fa.ax[slot] = sum.x; fa.ay[slot] = sum.y; fa.az[slot] = sum.z;
    }
    stat_add64(&stats[0], my_paths);
    stat_add64(&stats[1], ctr.extend);
    stat_add64(&stats[2], ctr.queries);
    stat_add64(&stats[3], ctr.traced);
    stat_add64(&stats[4], ctr.occl);
}

// raw closest-hit batch (parity hook + kernel-only benchmark)
__global__ void __launch_bounds__(kBlock) k_intersect_batch(const __grid_constant__ SceneDev sc, const double* __restrict__ o,
                                                             const double* __restrict__ d, size_t n, double t_min,
                                                             double* __restrict__ t, uint32_t* __restrict__ object_id,
                                                             uint32_t* __restrict__ prim_id, uint32_t* __restrict__ inst_id,
                                                             double* __restrict__ normal) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RayD ray;
    ray.o = mk3(o[3 * i], o[3 * i + 1], o[3 * i + 2]);
    ray.d = mk3(d[3 * i], d[3 * i + 1], d[3 * i + 2]);
    Hit h;
    h.t = CUDART_INF;
    closest_hit<kClosest>(sc, ray, t_min, 0.0, h);
    uint32_t prim = 0, inst = 0;
    d3 n3 = mk3(0, 0, 0);
    if (h.inst != kNoHit) n3 = hit_normal(sc, ray, h, &prim, &inst);
    if (t) t[i] = h.inst != kNoHit ? h.t : CUDART_INF;
    if (object_id) object_id[i] = h.obj;
    if (prim_id) prim_id[i] = prim;
    if (inst_id) inst_id[i] = inst;
    if (normal) { normal[3 * i] = n3.x; normal[3 * i + 1] = n3.y; normal[3 * i + 2] = n3.z; }
}

// ---- film ----

// batch_sum[pixel] = sum over lanes of the lane films (fixed order)
__global__ void k_batch_reduce(PathPool p, double* __restrict__ batch_sum) {
    const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= p.n_pixels) return;
    const uint32_t pixel = pool_pixel(p, r);
    double x = 0, y = 0, z = 0;
    for (uint32_t l = 0; l < p.n_lanes; l++) {
        const size_t s = (size_t)l * p.n_pixels + r;
        x += p.ax[s]; y += p.ay[s]; z += p.az[s];
    }
    batch_sum[3 * (size_t)pixel] = x; batch_sum[3 * (size_t)pixel + 1] = y; batch_sum[3 * (size_t)pixel + 2] = z;
}

// Renderer::get_color's mean and exposure (src/renderer.rs:183) + Buffer::add_samples
__global__ void k_batch_commit(const double* __restrict__ batch_sum, size_t n3, double iterations, double exposure,
                               double* __restrict__ batch_mean, double* __restrict__ film_sum) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n3) return;
    const double m = batch_sum[i] / iterations * exposure;
    batch_mean[i] = m;
    film_sum[i] += m;
}

// Buffer::image (src/buffer.rs:43-56,75-94) + color_bytes (src/color.rs:18-24)
__global__ void k_image(const double* __restrict__ film_sum, uint32_t w, uint32_t h, uint32_t r, uint32_t nbatches,
                        uint8_t* __restrict__ out) {
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= w || y >= h) return;
    double cx = 0, cy = 0, cz = 0;
    uint64_t count = 0;
    const uint32_t i0 = x >= r ? x - r : 0, j0 = y >= r ? y - r : 0;
    for (uint32_t i = i0; i <= x + r; i++)
        for (uint32_t j = j0; j <= y + r; j++)
            if (i < w && j < h) {
                const size_t idx = 3 * ((size_t)j * w + i);
                cx += film_sum[idx]; cy += film_sum[idx + 1]; cz += film_sum[idx + 2];
                count += nbatches;
            }
    const double c[3] = {cx / (double)count, cy / (double)count, cz / (double)count};
    for (int k = 0; k < 3; k++) {
        const double v = pow(fmin(fmax(c[k], 0.0), 1.0), 1.0 / 2.2) * 255.0;
        out[3 * ((size_t)y * w + x) + k] = (uint8_t)(v != v ? 0 : (v >= 255.0 ? 255 : (v <= 0.0 ? 0 : (int)v)));
    }
}

// Buffer::variance (src/buffer.rs:59-73): per-pixel unbiased variance of the batch means
__global__ void k_variance(const double* const* __restrict__ batches, uint32_t nbatches, const double* __restrict__ film_sum,
                           size_t n_pixels, double* __restrict__ out) {
    double acc = 0.0;
    for (size_t pidx = (size_t)blockIdx.x * blockDim.x + threadIdx.x; pidx < n_pixels; pidx += (size_t)gridDim.x * blockDim.x) {
        const double mx = film_sum[3 * pidx] / (double)nbatches, my = film_sum[3 * pidx + 1] / (double)nbatches,
                     mz = film_sum[3 * pidx + 2] / (double)nbatches;
        double ss = 0.0;
        for (uint32_t b = 0; b < nbatches; b++) {
            const double* s = batches[b] + 3 * pidx;
            const double ex = s[0] - mx, ey = s[1] - my, ez = s[2] - mz;
            ss += ex * ex + ey * ey + ez * ez;
        }
        acc += ss / ((double)nbatches - 1.0);
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

__global__ void k_film_mean(const double* __restrict__ film_sum, size_t n3, double nbatches, double* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n3) out[i] = film_sum[i] / nbatches;
}

// ---- mesh preparation ----
__global__ void k_tri_boxes(const RawTri* __restrict__ raw, uint32_t n, PrimBox* __restrict__ boxes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const RawTri t = raw[i];
    PrimBox b;
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const double lo = fmin(fmin(t.v[k], t.v[3 + k]), t.v[6 + k]);
        const double hi = fmax(fmax(t.v[k], t.v[3 + k]), t.v[6 + k]);
        b.mn[k] = __double2float_rd(lo);
        b.mx[k] = __double2float_ru(hi);
    }
    boxes[i] = b;
}

// per-triangle constants of Triangle::intersect, same f64 operations as src/shape/mesh.rs:51-66
__global__ void k_tri_prepare(const RawTri* __restrict__ raw, const uint32_t* __restrict__ perm, uint32_t n, TriRec* __restrict__ out,
                              TriF* __restrict__ outf, TriNormals* __restrict__ nrm, uint32_t* __restrict__ orig) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t src = perm[i];
    const RawTri t = raw[src];
    const d3 v1 = mk3(t.v[0], t.v[1], t.v[2]), v2 = mk3(t.v[3], t.v[4], t.v[5]), v3 = mk3(t.v[6], t.v[7], t.v[8]);
    const d3 d0 = v2 - v1, d1 = v3 - v1;
    const d3 pn = normalize(cross(d0, d1));
    TriRec r;
    r.v1[0] = v1.x; r.v1[1] = v1.y; r.v1[2] = v1.z;
    r.d0[0] = d0.x; r.d0[1] = d0.y; r.d0[2] = d0.z;
    r.d1[0] = d1.x; r.d1[1] = d1.y; r.d1[2] = d1.z;
    r.pn[0] = pn.x; r.pn[1] = pn.y; r.pn[2] = pn.z;
    r.d00 = dot(d0, d0); r.d01 = dot(d0, d1); r.d11 = dot(d1, d1);
    r.denom = r.d00 * r.d11 - r.d01 * r.d01;
    out[i] = r;
    TriF tf;
    tf.v0[0] = (float)v1.x; tf.v0[1] = (float)v1.y; tf.v0[2] = (float)v1.z; tf.pad0 = 0.f;
    tf.e1[0] = (float)d0.x; tf.e1[1] = (float)d0.y; tf.e1[2] = (float)d0.z; tf.pad1 = 0.f;
    tf.e2[0] = (float)d1.x; tf.e2[1] = (float)d1.y; tf.e2[2] = (float)d1.z; tf.pad2 = 0.f;
    outf[i] = tf;
    TriNormals tn;
    for (int k = 0; k < 3; k++) { tn.n1[k] = t.n[k]; tn.n2[k] = t.n[3 + k]; tn.n3[k] = t.n[6 + k]; }
    nrm[i] = tn;
    orig[i] = src;
}

// Triangle::from_vertices (src/shape/mesh.rs:27-38) for meshes uploaded without normals
__global__ void k_face_normals(RawTri* __restrict__ raw, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RawTri t = raw[i];
    const d3 v1 = mk3(t.v[0], t.v[1], t.v[2]), v2 = mk3(t.v[3], t.v[4], t.v[5]), v3 = mk3(t.v[6], t.v[7], t.v[8]);
    const d3 nn = normalize(cross(v2 - v1, v3 - v1));
    for (int k = 0; k < 3; k++) { raw[i].n[3 * k] = nn.x; raw[i].n[3 * k + 1] = nn.y; raw[i].n[3 * k + 2] = nn.z; }
}

}  // namespace rptx

// Persistent-thread path kernel: Renderer::sample (src/renderer.rs:158-184) with the whole of trace_ray
// (:187-322) in ONE launch per batch and no path state in HBM.
//
// A thread owns a slot (pixel, sample lane) at a time and walks the slot's samples one vertex per trip of a
// single flat loop: extend -> vertex (medium collision | surface | miss) -> NEE visibility -> continuation.
// A lane whose path ended starts the slot's next sample at the top of the very next trip, and a lane whose
// slot ran out of samples claims the next free slot (one atomic per warp and claim), so no lane ever waits
// for the longest path of its warp: that wait is what made the per-thread kernel of round 1 (k_path_fused,
// one path per thread per outer loop trip) 7-11x slower than the wavefront with Russian roulette.  The three
// vertex kinds share the light sampling, the NEE trace and the accumulation code; they diverge only where
// their arithmetic differs (hit normal + material vs medium colour; BSDF sample vs phase sample).
//
// Block-synchronous phases.  The loop body is ~60 KB of hot SASS (f64 primitive tests, light sampling, BSDF
// sampling) against 32 KB of L1.5 instruction cache per SM; with every warp at its own place in the loop the
// kernel was bound by instruction fetch (ncu: 19 % of the issue slots, "no instruction" the top stall at every
// line; profiles/r2_summary.md).  So ONE block owns an SM and its warps walk the loop in step: a
// __syncthreads() separates the phases (refill | extend | vertex + light sample | NEE visibility |
// continuation), each phase's code fits the instruction cache and is fetched once per SM and trip instead of
// once per warp.  The NEE phase uses the barrier it already pays for: the block's shadow jobs (about half of
// the vertices have one: light sample facing away, bsdf 0) are compacted through shared memory and traced by
// the first `total` threads, so the traversal runs with full warps, and the verdict returns to the owner.
//
// Same arithmetic, the same RNG draw order and the same summation order as the wavefront kernels
// (path_kernels.cuh): a slot adds its samples' radiance in sample order, the lanes of a pixel are added in
// lane order by k_batch_reduce, so the batch does not depend on which thread ran which slot.
#pragma once
#include "path.cuh"
#include "shading.cuh"
#include "trace.cuh"

#ifndef RPTX_PERSIST_BLOCK
#define RPTX_PERSIST_BLOCK 512
#endif
#ifndef RPTX_PERSIST_OCC
#define RPTX_PERSIST_OCC 1
#endif

namespace rptx {

constexpr uint32_t kNoSlot = 0xffffffffu;
constexpr int kPersistBlock = RPTX_PERSIST_BLOCK;

struct PersistArgs {
    Camera cam;
    BatchParams bp;
    uint32_t n_pixels, n_lanes, n_slots, width, tiled;
    uint32_t stack_depth;        // surface mode: max_bounces + 1 clamp records per thread
    double *ax, *ay, *az;        // [n_lanes][n_pixels] slot sums of the batch
    double* stack;               // surface mode: [level][thread]{direct.xyz, f.xyz, pdf, |cos|}, 64 B records
    uint32_t* next_slot;         // claim counter
};

RPTX_D uint32_t persist_pixel(const PersistArgs& a, uint32_t slot) {
    const uint32_t r = slot % a.n_pixels;
    if (!a.tiled) return r;
    const uint32_t tile = r >> 5, in = r & 31u, tiles_x = a.width >> 3;
    const uint32_t tx = tile % tiles_x, ty = tile / tiles_x;
    return (ty * 4u + (in >> 3)) * a.width + tx * 8u + (in & 7u);
}

// shared memory of one block: the shadow-job exchange and the per-thread accumulators that are only touched at the
// end of a trip (kept out of the register file while the traversals run); every array is [field][thread]: conflict-free
struct PersistShared {
    double job[7][kPersistBlock];   // shadow ray origin, direction, distance to the light sample
    double acc[9][kPersistBlock];   // slot sum, path throughput, path radiance
    uint16_t owner[kPersistBlock];  // thread that owns job j
    uint8_t visible[kPersistBlock]; // verdict for the job of thread t
    uint32_t count[kPersistBlock / 32];
};

template <bool MEDIUM, bool PROBE, bool SMALL>
__global__ void __launch_bounds__(kPersistBlock, RPTX_PERSIST_OCC)
k_path_persist(const __grid_constant__ SceneDev sc, const __grid_constant__ MediumDev md, const __grid_constant__ PersistArgs a,
               unsigned long long* __restrict__ stats) {
    extern __shared__ __align__(16) unsigned char persist_smem[];
    PersistShared& sh = *reinterpret_cast<PersistShared*>(persist_smem);
    const uint32_t tx = threadIdx.x;
    const uint32_t tid = blockIdx.x * blockDim.x + tx, n_threads = gridDim.x * blockDim.x;
    const uint32_t lane = tx & 31u, warp = tx >> 5;
    unsigned long long n_paths = 0, n_extend = 0, n_queries = 0, n_traced = 0, n_occl = 0;
    uint32_t slot = kNoSlot, j = 0, pixel = 0, depth = 0;
    bool idle = false, alive = false;
    RayD ray;
    ray.o = mk3(0, 0, 0); ray.d = mk3(0, 0, 1);
    Rng rng;
    rng.state = 0;
    double2* const my_stack = MEDIUM ? nullptr : reinterpret_cast<double2*>(a.stack) + (size_t)tid * 4;
    const size_t level_stride = (size_t)n_threads * 4;  // double2 per level
    auto acc_get = [&](int f) { return mk3(sh.acc[f][tx], sh.acc[f + 1][tx], sh.acc[f + 2][tx]); };
    auto acc_set = [&](int f, d3 v) { sh.acc[f][tx] = v.x; sh.acc[f + 1][tx] = v.y; sh.acc[f + 2][tx] = v.z; };
    constexpr int kSum = 0, kBeta = 3, kRad = 6;

    for (;;) {
        // ================= phase 0: next sample of the slot, or the next slot =================
        if (!alive && !idle && slot != kNoSlot) {
            j += a.n_lanes;
            if (j >= a.bp.n_local) {
                const d3 sum = acc_get(kSum);
                a.ax[slot] = sum.x; a.ay[slot] = sum.y; a.az[slot] = sum.z;
                slot = kNoSlot;
            }
        }
        {
            const bool want = !alive && !idle && slot == kNoSlot;
            const unsigned m = __ballot_sync(0xffffffffu, want);
            if (m) {
                uint32_t base = 0;
                const int leader = __ffs(m) - 1;
                if ((int)lane == leader) base = atomicAdd(a.next_slot, (uint32_t)__popc(m));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (want) {
                    const uint32_t s = base + (uint32_t)__popc(m & ((1u << lane) - 1u));
                    if (s < a.n_slots) {
                        slot = s;
                        j = s / a.n_pixels;  // the slot's lane: its samples are j, j + n_lanes, ...
                        pixel = persist_pixel(a, s);
                        acc_set(kSum, mk3(0, 0, 0));
                    } else {
                        idle = true;
                    }
                }
            }
        }
        if (!alive && !idle) {
            const uint32_t x = pixel % a.bp.width, y = pixel / a.bp.width;
            const uint64_t s = a.bp.batch_base + a.bp.shard_index + (uint64_t)j * a.bp.shard_count;
            rng = Rng::make(a.bp.seed, kDomainCamera, pixel, s);
            primary_ray(a.cam, a.bp.width, a.bp.height, x, y, rng, &ray.o, &ray.d);
            depth = 0;
            if (MEDIUM) { acc_set(kBeta, mk3(1, 1, 1)); acc_set(kRad, mk3(0, 0, 0)); }
            alive = true;
            n_paths++;
        }
        if (!__syncthreads_or(alive ? 1 : 0)) break;  // block-uniform: every thread ran out of slots

        // ================= phase 1: extend =================
        double dsamp = CUDART_INF;
        Hit h;
        h.t = CUDART_INF; h.inst = kNoHit; h.obj = kNoHit; h.prim = 0; h.v = 0; h.w = 0;
        d3 n = mk3(0, 0, 0);
        if (alive) {
            if (MEDIUM) dsamp = medium_sample_d(md, rng);
            closest_hit_sel<kClosest, SMALL>(sc, ray, kEpsilon, 0.0, h);
            n_extend++;
            if (h.inst != kNoHit && !(MEDIUM && dsamp < h.t)) {
                uint32_t prim, inst;
                n = hit_normal(sc, ray, h, &prim, &inst);
            }
        }
        const bool hit = h.inst != kNoHit;
        const bool vol = alive && MEDIUM && dsamp < (hit ? h.t : kBackgroundDist);
        const bool vertex = alive && (vol || hit);  // a medium collision or a surface hit; otherwise the path left the scene
        __syncthreads();

        // ================= phase 2: vertex + light samples, phase 3: NEE visibility (per light) =================
        d3 direct = mk3(0, 0, 0), pos = ray.o, wo = mk3(0, 0, 0), mc = mk3(0, 0, 0);
        Material mat;
        mat.kind = 0; mat.emittance = 0; mat.shininess = 0; mat.ior = 0; mat.albedo[0] = mat.albedo[1] = mat.albedo[2] = 0;
        if (alive && !vertex) direct = env_color(sc, ray.d);  // src/renderer.rs:198-206 / :291
        if (vertex) {
            wo = -normalize(ray.d);
            if (vol) {  // medium vertex, src/renderer.rs:243-283
                pos = ray.o + dsamp * ray.d;
                mc = md.color(pos);
                if (depth == 0) direct = md.emission * mc;
            } else {    // surface vertex, src/renderer.rs:207-236 / :289-319
                pos = ray.o + h.t * ray.d;
                mat = sc.materials[h.obj];
                if (depth == 0) direct = mat_emittance(mat) * mat_color(mat);
            }
        }
        // NEE, Renderer::sample_lights / sample_lights_for_media (one body for both kinds of vertex); the light count is
        // uniform, so the barriers inside the loop are reached by every thread
        for (uint32_t li = 0; li < sc.n_lights; li++) {
            const Light& l = sc.lights[li];
            if (l.kind == 1) {
                if (vertex) {
                    const d3 amb = mk3(l.color[0], l.color[1], l.color[2]);
                    direct = direct + cmul(amb, vol ? mc : mat_color(mat));
                }
                continue;
            }
            d3 c = mk3(0, 0, 0), wi = mk3(0, 0, 0);
            double dist = 0.0;
            bool job = false;
            if (vertex) {
                d3 intensity;
                illuminate(sc, l, pos, rng, &intensity, &wi, &dist);
                if (l.kind != 2) {  // a directional light's dist = inf never passes the 1e-12 test (src/renderer.rs:396)
                    if (vol) c = DDIV(md.sigma_s, md.extinction()) * cmul(intensity, mc) * md.phase;
                    else c = cmul(mat_bsdf(mat, n, wo, wi), intensity) * dot(wi, n);
                    n_queries++;
                    job = !(c.x == 0.0 && c.y == 0.0 && c.z == 0.0);
                }
            }
            if (l.kind == 2) continue;
            // ---- phase 3: the block's jobs, densely packed ----
            const unsigned m = __ballot_sync(0xffffffffu, job);
            if (lane == 0) sh.count[warp] = (uint32_t)__popc(m);
            __syncthreads();
            uint32_t before = 0, total = 0;
#pragma unroll
            for (uint32_t w = 0; w < kPersistBlock / 32; w++) {
                const uint32_t cw = sh.count[w];
                if (w < warp) before += cw;
                total += cw;
            }
            if (job) {
                const uint32_t at = before + (uint32_t)__popc(m & ((1u << lane) - 1u));
                sh.job[0][at] = pos.x; sh.job[1][at] = pos.y; sh.job[2][at] = pos.z;
                sh.job[3][at] = wi.x; sh.job[4][at] = wi.y; sh.job[5][at] = wi.z;
                sh.job[6][at] = dist;
                sh.owner[at] = (uint16_t)tx;
            }
            __syncthreads();
            if (tx < total) {
                RayD sr;
                sr.o = mk3(sh.job[0][tx], sh.job[1][tx], sh.job[2][tx]);
                sr.d = mk3(sh.job[3][tx], sh.job[4][tx], sh.job[5][tx]);
                n_traced++;
                sh.visible[sh.owner[tx]] = nee_visible<PROBE, SMALL>(sc, sr, sh.job[6][tx], &n_occl) ? 1 : 0;
            }
            __syncthreads();
            if (job && sh.visible[tx]) direct = direct + c;
        }

        // ================= phase 4: continuation + accumulation =================
        d3 fc = mk3(0, 0, 0);
        double rec_pdf = 0.0, rec_ac = 0.0;
        bool go = false;
        if (vertex) {
            ray.o = pos;
            // Russian roulette 0.8 with a medium (src/renderer.rs:193,222,262), max_bounces without
            if (MEDIUM ? (rng.gen() < kRrMedium) : (depth < a.bp.max_bounces)) {
                if (vol) {
                    d3 wi;
                    medium_sample_ph(rng, &wi);
                    const double ph_p = md.phase;
                    const double sa = DDIV(md.sigma_s, md.extinction());
                    const double sp = DDIV(sa, ph_p);
                    fc = mk3(DDIV((sp * mc.x) * md.phase, kRrMedium), DDIV((sp * mc.y) * md.phase, kRrMedium), DDIV((sp * mc.z) * md.phase, kRrMedium));
                    ray.d = wi;
                    go = true;
                } else {
                    d3 wi;
                    double pdf;
                    if (mat_sample_f(mat, n, wo, rng, &wi, &pdf)) {
                        const d3 f = mat_bsdf(mat, n, wo, wi);
                        const double ac = fabs(dot(wi, n));
                        if (MEDIUM) {
                            const double s = DDIV(1.0, pdf * kRrMedium);
                            fc = mk3(s * f.x * ac, s * f.y * ac, s * f.z * ac);
                        } else {
                            fc = f; rec_pdf = pdf; rec_ac = ac;
                        }
                        ray.d = wi;
                        go = true;
                    }
                }
            }
        }
        if (alive) {
            d3 L;
            if (MEDIUM) {
                const d3 beta = acc_get(kBeta);
                L = acc_get(kRad) + cmul(beta, direct);
                if (go) { acc_set(kRad, L); acc_set(kBeta, mk3(beta.x * fc.x, beta.y * fc.y, beta.z * fc.z)); }
            } else if (go) {
                // level record of the clamped back-fold (src/renderer.rs:311-313)
                double2* r = my_stack + (size_t)depth * level_stride;
                r[0] = make_double2(direct.x, direct.y);
                r[1] = make_double2(direct.z, fc.x);
                r[2] = make_double2(fc.y, fc.z);
                r[3] = make_double2(rec_pdf, rec_ac);
            } else {
                L = direct;
                for (int k = (int)depth - 1; k >= 0; k--) {
                    const double2* r = my_stack + (size_t)k * level_stride;
                    const double2 r0 = r[0], r1 = r[1], r2 = r[2], r3 = r[3];
                    const d3 f = mk3(r1.y, r2.x, r2.y);
                    const d3 ind = DDIV(1.0, r3.x) * cmul(f, L) * r3.y;
                    L = mk3(r0.x + fmin(ind.x, kFireflyClamp), r0.y + fmin(ind.y, kFireflyClamp), r1.x + fmin(ind.z, kFireflyClamp));
                }
            }
            if (go) {
                depth++;
            } else {
                acc_set(kSum, acc_get(kSum) + L);
                alive = false;
            }
        }
        // phase 0 of the next trip touches only this thread's own shared slots: no barrier needed here (the
        // __syncthreads_or at its end orders the trip)
    }
    stat_add64(&stats[0], n_paths);
    stat_add64(&stats[1], n_extend);
    stat_add64(&stats[2], n_queries);
    stat_add64(&stats[3], n_traced);
    stat_add64(&stats[4], n_occl);
}

}  // namespace rptx

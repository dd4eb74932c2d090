// Persistent-thread path kernel: Renderer::sample (src/renderer.rs:158-184) with the whole of trace_ray
// (:187-322) in ONE launch per batch and no path state in HBM.
//
// A thread owns a slot (pixel, sample lane) at a time and walks the slot's samples one vertex per trip of a
// single flat loop: extend -> vertex (medium collision | surface | miss) -> NEE visibility -> continuation.
// A lane whose path ended starts the slot's next sample at the top of the very next trip, and a lane whose
// slot ran out of samples claims the next free slot (one atomic per warp and claim), so no lane ever waits
// for the longest path of its warp: that wait is what made the per-thread kernel of round 1 (k_path_fused,
// one path per thread per outer loop trip) 2-3x slower than the wavefront with Russian roulette.  The three
// vertex kinds share the light sampling, the NEE trace and the accumulation code; they diverge only where
// their arithmetic differs (hit normal + material vs medium colour; BSDF sample vs phase sample).
//
// Same arithmetic, the same RNG draw order and the same summation order as the wavefront kernels
// (path_kernels.cuh): a slot adds its samples' radiance in sample order, the lanes of a pixel are added in
// lane order by k_batch_reduce, so the batch does not depend on which thread ran which slot.
#pragma once
#include "path.cuh"
#include "shading.cuh"
#include "trace.cuh"

#ifndef RPTX_OCC_PERSIST
#define RPTX_OCC_PERSIST 4
#endif
#ifndef RPTX_OCC_PERSIST_SMALL
#define RPTX_OCC_PERSIST_SMALL 4
#endif

namespace rptx {

constexpr uint32_t kNoSlot = 0xffffffffu;

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

template <bool MEDIUM, bool PROBE, bool SMALL>
__global__ void __launch_bounds__(kBlock, SMALL ? RPTX_OCC_PERSIST_SMALL : RPTX_OCC_PERSIST)
k_path_persist(const __grid_constant__ SceneDev sc, const __grid_constant__ MediumDev md, const __grid_constant__ PersistArgs a,
               unsigned long long* __restrict__ stats) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, n_threads = gridDim.x * blockDim.x;
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long n_paths = 0, n_extend = 0, n_queries = 0, n_traced = 0, n_occl = 0;
    uint32_t slot = kNoSlot, j = 0, pixel = 0, depth = 0;
    bool idle = false, alive = false;
    d3 sum = mk3(0, 0, 0), beta = mk3(1, 1, 1), L = mk3(0, 0, 0);
    RayD ray;
    ray.o = mk3(0, 0, 0); ray.d = mk3(0, 0, 1);
    Rng rng;
    rng.state = 0;
    double2* const my_stack = MEDIUM ? nullptr : reinterpret_cast<double2*>(a.stack) + (size_t)tid * 4;
    const size_t level_stride = (size_t)n_threads * 4;  // double2 per level

    for (;;) {
        // ---- next sample of the slot, or the next slot ----
        if (!alive && !idle && slot != kNoSlot) {
            j += a.n_lanes;
            if (j >= a.bp.n_local) {
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
                        sum = mk3(0, 0, 0);
                    } else {
                        idle = true;
                    }
                }
            }
        }
        if (__all_sync(0xffffffffu, idle)) break;
        if (!alive && !idle) {
            const uint32_t x = pixel % a.bp.width, y = pixel / a.bp.width;
            const uint64_t s = a.bp.batch_base + a.bp.shard_index + (uint64_t)j * a.bp.shard_count;
            rng = Rng::make(a.bp.seed, kDomainCamera, pixel, s);
            primary_ray(a.cam, a.bp.width, a.bp.height, x, y, rng, &ray.o, &ray.d);
            depth = 0;
            if (MEDIUM) { beta = mk3(1, 1, 1); L = mk3(0, 0, 0); }
            alive = true;
            n_paths++;
        }
        if (!alive) continue;  // idle lane of a warp that still has work

        // ---- one vertex ----
        const double dsamp = MEDIUM ? medium_sample_d(md, rng) : CUDART_INF;
        Hit h;
        h.t = CUDART_INF;
        closest_hit_sel<kClosest, SMALL>(sc, ray, kEpsilon, 0.0, h);
        n_extend++;
        const bool hit = h.inst != kNoHit;
        const bool vol = MEDIUM && dsamp < (hit ? h.t : kBackgroundDist);
        d3 direct = mk3(0, 0, 0), fc = mk3(0, 0, 0);
        double rec_pdf = 0.0, rec_ac = 0.0;
        bool go = false;
        if (!vol && !hit) {
            direct = env_color(sc, ray.d);  // src/renderer.rs:198-206 / :291
        } else {
            d3 pos, n = mk3(0, 0, 0), mc = mk3(0, 0, 0);
            Material mat;
            const d3 wo = -normalize(ray.d);
            if (vol) {  // medium vertex, src/renderer.rs:243-283
                pos = ray.o + dsamp * ray.d;
                mc = md.color(pos);
                if (depth == 0) direct = md.emission * mc;
            } else {    // surface vertex, src/renderer.rs:207-236 / :289-319
                uint32_t prim, inst;
                n = hit_normal(sc, ray, h, &prim, &inst);
                pos = ray.o + h.t * ray.d;
                mat = sc.materials[h.obj];
                if (depth == 0) direct = mat_emittance(mat) * mat_color(mat);
            }
            // NEE, Renderer::sample_lights / sample_lights_for_media (one body for both kinds of vertex)
            for (uint32_t li = 0; li < sc.n_lights; li++) {
                const Light& l = sc.lights[li];
                if (l.kind == 1) {
                    const d3 amb = mk3(l.color[0], l.color[1], l.color[2]);
                    direct = direct + cmul(amb, vol ? mc : mat_color(mat));
                    continue;
                }
                d3 intensity, wi;
                double dist;
                illuminate(sc, l, pos, rng, &intensity, &wi, &dist);
                if (l.kind == 2) continue;  // dist = inf never passes the 1e-12 test (src/renderer.rs:396)
                d3 c;
                if (vol) c = (md.sigma_s / md.extinction()) * cmul(intensity, mc) * md.phase;
                else c = cmul(mat_bsdf(mat, n, wo, wi), intensity) * dot(wi, n);
                n_queries++;
                if (c.x == 0.0 && c.y == 0.0 && c.z == 0.0) continue;
                RayD sr;
                sr.o = pos;
                sr.d = wi;
                n_traced++;
                if (nee_visible<PROBE, SMALL>(sc, sr, dist, &n_occl)) direct = direct + c;
            }
            ray.o = pos;
            // continuation: Russian roulette 0.8 with a medium (src/renderer.rs:193,222,262), max_bounces without
            if (MEDIUM ? (rng.gen() < kRrMedium) : (depth < a.bp.max_bounces)) {
                if (vol) {
                    d3 wi;
                    medium_sample_ph(rng, &wi);
                    const double ph_p = md.phase;
                    const double sa = md.sigma_s / md.extinction();
                    fc = mk3(((sa / ph_p) * mc.x * md.phase) / kRrMedium, ((sa / ph_p) * mc.y * md.phase) / kRrMedium,
                             ((sa / ph_p) * mc.z * md.phase) / kRrMedium);
                    ray.d = wi;
                    go = true;
                } else {
                    d3 wi;
                    double pdf;
                    if (mat_sample_f(mat, n, wo, rng, &wi, &pdf)) {
                        const d3 f = mat_bsdf(mat, n, wo, wi);
                        const double ac = fabs(dot(wi, n));
                        if (MEDIUM) {
                            const double s = 1.0 / (pdf * kRrMedium);
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
        if (MEDIUM) {
            L = L + cmul(beta, direct);
            if (go) beta = mk3(beta.x * fc.x, beta.y * fc.y, beta.z * fc.z);
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
                const d3 ind = 1.0 / r3.x * cmul(f, L) * r3.y;
                L = mk3(r0.x + fmin(ind.x, kFireflyClamp), r0.y + fmin(ind.y, kFireflyClamp), r1.x + fmin(ind.z, kFireflyClamp));
            }
        }
        if (go) {
            depth++;
        } else {
            sum = sum + L;
            alive = false;
        }
    }
    stat_add64(&stats[0], n_paths);
    stat_add64(&stats[1], n_extend);
    stat_add64(&stats[2], n_queries);
    stat_add64(&stats[3], n_traced);
    stat_add64(&stats[4], n_occl);
}

}  // namespace rptx

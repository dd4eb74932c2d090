// Photon integrators of the B200 core: device records and kernels replacing src/photon.rs of
// the reference — photon emission and tracing (shoot_photon :724-799, trace_photon :803-946),
// the photon-map structures (PhotonMap::new_* :193-305; the kd-tree 0.4.1 and bvh 0.6.0 crates
// behind them) and the three radiance estimators (estimate_indirect :316-627) with the camera
// pass (get_color_with_photon_map :950-985).
//
// Layout in HBM:
//  * emission appends 104-byte PhotonRec rows (the reference's Photon, :25-34, plus a 64-bit
//    key = photon index << 16 | depth that makes the list order canonical after a sort, so
//    results do not depend on thread scheduling or on how many GPUs emitted the photons);
//  * a point map keeps its photons in 63-bit Morton order as structure-of-arrays (f64 position
//    triples read by the kNN search; direction / power / key touched only for the k winners)
//    under a Karras LBVH built on those codes (64-byte nodes holding both children's f32 boxes,
//    leaves of <= kPointLeaf consecutive photons).  Splitting at Morton-cell boundaries matters:
//    an implicit equal-count tree over the same order visited 8x more nodes per query;
//  * photon spheres / beams sit under the same LBVH the object accelerator uses
//    (bvh_build.cu), with their own f32 boxes computed exactly as the reference does
//    (:76-105, :119-128) so the candidate set of BVH::traverse is reproduced.
#pragma once
#include <cooperative_groups.h>
#include <cooperative_groups/reduce.h>

#include "path.cuh"
#include "shading.cuh"
#include "trace.cuh"

// resident blocks per SM asked of ptxas for the warp-per-sample gather kernel (register cap = 65536 / (128 * n))
#ifndef RPTX_OCC_GATHER
#define RPTX_OCC_GATHER 4
#endif

namespace rptx {

namespace cg = cooperative_groups;

struct PhotonRec {
    double pos[3], dir[3], power[3], start[3];
    uint64_t key;
};
static_assert(sizeof(PhotonRec) == 104, "PhotonRec is 13 x 8 bytes (rows of the C ABI)");

constexpr double kPhotonRrDiffuse = 0.7;  // src/photon.rs:823
constexpr double kBeamThin = 0.001;       // src/photon.rs:780
constexpr double kBeamRadius = 3.0;       // src/photon.rs:277
constexpr uint32_t kSphereKnn = 10;       // src/photon.rs:217
constexpr int kPointLeaf = 8;

RPTX_HD uint64_t photon_key(uint64_t photon_index, uint32_t depth) {
    return (photon_index << 16) | (uint64_t)(depth < 65535u ? depth : 65535u);
}

struct PhotonSink {
    PhotonRec* rows;
    unsigned long long* count;  // keeps counting past cap (the host then grows the buffer and re-runs)
    unsigned long long cap;
};

// one atomic per group of converged lanes
RPTX_D void photon_deposit(const PhotonSink& s, d3 pos, d3 dir, d3 power, d3 start, uint64_t key) {
    const cg::coalesced_group g = cg::coalesced_threads();
    unsigned long long base = 0;
    if (g.thread_rank() == 0) base = atomicAdd(s.count, (unsigned long long)g.size());
    base = g.shfl(base, 0);
    const unsigned long long i = base + g.thread_rank();
    if (i >= s.cap) return;
    PhotonRec* r = s.rows + i;
    r->pos[0] = pos.x; r->pos[1] = pos.y; r->pos[2] = pos.z;
    r->dir[0] = dir.x; r->dir[1] = dir.y; r->dir[2] = dir.z;
    r->power[0] = power.x; r->power[1] = power.y; r->power[2] = power.z;
    r->start[0] = start.x; r->start[1] = start.y; r->start[2] = start.z;
    r->key = key;
}

struct EmitParams {
    uint64_t seed;
    uint64_t first, count;  // this call shoots photon indices [first, first + count)
    double power;           // watts / photon_count (src/photon.rs:671)
    int kind;               // PhotonRenderKind
    uint32_t light_index;   // first Light::Object of the scene (src/photon.rs:727-729)
};

// shoot_photon + trace_photon as one walk per lane.  Persistent warps; every pass of the loop is one segment
// of each live lane's photon, and a lane whose photon ended takes the next photon index at the top of the
// loop (one atomic per warp for all idle lanes, once at least kRefillLanes are idle), so lanes do not
// wait for the longest walk among 32.  Rows leave in arbitrary order; their key (photon index, depth)
// restores the canonical one (photon.cu canonical_order).  The reference's recursion deposits a vertex
// only when the walk continues past it (:842-866), which the loop reproduces by depositing right after a
// successful sample_f.
// idle lanes at which a warp pulls new photons (one atomic for all of them).  Measured at 1e8 photons, emission ms for the
// photon-photon / beam-beam lampshade: 16 lanes 139 / 233, 8 (round 1) 111 / 223, 4: 99 / 175, 2: 96 / 160, 1: 96 / 159 - the
// atomic is cheap next to a lane idling through a neighbour's whole walk
#ifndef RPTX_REFILL_LANES
#define RPTX_REFILL_LANES 1
#endif
constexpr int kRefillLanes = RPTX_REFILL_LANES;

// resident blocks asked of ptxas for the walk (registers = 65536 / (128 * n)).  Measured at 1e8 photons, emission ms for the
// photon-photon / beam-beam lampshade: 3 blocks (142 registers as compiled freely) 115 / 250, 4 (128) 108 / 217, 5 (96) 108 / 216,
// 6 (80, 352 B of spills) 113 / 230
#ifndef RPTX_OCC_EMIT
#define RPTX_OCC_EMIT 4
#endif
template <bool SMALL>
__global__ void __launch_bounds__(128, RPTX_OCC_EMIT) k_photon_trace(const __grid_constant__ SceneDev sc, const MediumDev md, const EmitParams ep,
                                                      PhotonSink surf, PhotonSink vol, unsigned long long* __restrict__ work_head,
                                                      unsigned long long* __restrict__ stats) {
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long my_rays = 0;
    bool alive = false, exhausted = false;
    uint64_t pid = 0;
    uint32_t depth = 0;
    Rng rng = Rng::make(ep.seed, kDomainPhoton, 0, 0);
    RayD ray;
    ray.o = mk3(0, 0, 0); ray.d = mk3(0, 0, 1);
    d3 power = mk3(0, 0, 0);
    for (;;) {
        const unsigned idle = __ballot_sync(0xffffffffu, !alive);
        if (!exhausted && (__popc(idle) >= kRefillLanes || idle == 0xffffffffu)) {
            unsigned long long base = 0;
            if (lane == 0) base = atomicAdd(work_head, (unsigned long long)__popc(idle));
            base = __shfl_sync(0xffffffffu, base, 0);
            exhausted = base + (unsigned long long)__popc(idle) >= ep.count;
            const unsigned long long i = base + (unsigned long long)__popc(idle & ((1u << lane) - 1u));
            if (!alive && i < ep.count) {
                pid = ep.first + i;
                rng = Rng::make(ep.seed, kDomainPhoton, pid, 0);
                const Light& l = sc.lights[ep.light_index];
                d3 pos, n;
                double light_pdf;
                sample_light_shape(sc, l, mk3(0, 0, 0), rng, &pos, &n, &light_pdf);
                const double phi = 2.0 * kPi * rng.gen();
                const double ct = 1.0 - rng.gen(), st = sin_from_cos(ct);  // theta = acos(1 - u), src/photon.rs:737
                double sp, cp;
                sincos(phi, &sp, &cp);
                const d3 hemi = mk3(st * cp, ct, st * sp);
                const Rot3 rot = rotation_from_y(n);
                ray.o = pos;
                ray.d = rot_apply(rot, hemi);
                power = ep.power * mat_color(l.material);
                depth = 0;
                alive = true;
            }
        }
        if (__ballot_sync(0xffffffffu, alive) == 0) break;
        if (alive) {
            const d3 wo = -normalize(ray.d);
            Hit h;
            h.t = CUDART_INF;
            closest_hit_sel<kClosest, SMALL>(sc, ray, kEpsilon, 0.0, h);
            my_rays++;
            const bool hit = h.inst != kNoHit;
            double dsamp = CUDART_INF;
            if (md.kind != 0) dsamp = medium_sample_d(md, rng);
            else if (!hit) alive = false;
            if (!alive) {
                // left the scene (no medium)
            } else if (md.kind != 0 && (!hit || dsamp < h.t)) {
                // trace_in_volume, src/photon.rs:877-927
                const d3 collision = ray.o + dsamp * ray.d;
                const d3 mc = md.color(collision);
                const double scat = md.sigma_s, extinction = md.sigma_a + md.sigma_s;
                const d3 attenuated = cmul(power, mc) * scat / extinction;
                const double rr_prob = scat / extinction;
                const bool go = rng.gen() < rr_prob;
                d3 wi = mk3(0, 0, 0);
                if (go) medium_sample_ph(rng, &wi);
                if (ep.kind == 2) {  // beam-beam thinning, src/photon.rs:773-793
                    Rng thin = Rng::make(ep.seed, kDomainThin, pid, depth);
                    if (thin.gen() < kBeamThin) photon_deposit(vol, collision, wo, power / kBeamThin, ray.o, photon_key(pid, depth));
                } else {
                    photon_deposit(vol, collision, wo, power, ray.o, photon_key(pid, depth));
                }
                if (!go) {
                    alive = false;
                } else {
                    power = attenuated * md.phase / md.phase;  // phase(wo, wi) / ph_p, both the medium's constant
                    ray.o = collision;
                    ray.d = wi;
                }
            } else {
                // trace_on_surface, src/photon.rs:813-874
                uint32_t prim, inst;
                const d3 hn = hit_normal(sc, ray, h, &prim, &inst);
                const d3 world_pos = ray.o + h.t * ray.d;
                const Material mat = sc.materials[h.obj];
                d3 wi;
                double pdf;
                if (!(rng.gen() < kPhotonRrDiffuse)) {
                    alive = false;  // absorbed
                } else if (!mat_sample_f(mat, hn, wo, rng, &wi, &pdf)) {
                    alive = false;  // total internal reflection
                } else {
                    const d3 f = mat_bsdf(mat, hn, wo, wi);
                    const double wn = dot(wi, hn);
                    const double cosine_term = wn > 0.0 ? wn : 1.0;
                    const d3 attenuated = cmul(power, f) * cosine_term / pdf / kPhotonRrDiffuse;
                    if (!mat_is_mirror(mat)) photon_deposit(surf, world_pos, wo, power, world_pos, photon_key(pid, depth));
                    power = attenuated;
                    ray.o = world_pos;
                    ray.d = wi;
                }
            }
            depth++;
        }
        __syncwarp();
    }
    const unsigned long long tot = __reduce_add_sync(0xffffffffu, (unsigned)my_rays) ;
    if (lane == 0 && tot) atomicAdd(&stats[0], tot);
}

// ------------------------------------------------------------------ point map (kd-tree 0.4.1)

struct PointMapDev {
    uint32_t n;           // photons
    int root_ref;         // LBVH root (leaf ref when n <= kPointLeaf)
    const BvhNode* nodes; // Karras LBVH over the 63-bit Morton order: 64 B, both children's f32 boxes (rounded outward)
    const double *px, *py, *pz;  // Morton order
    const double* dir;    // [n][3]
    const double* power;  // [n][3]
    const uint64_t* key;  // [n]
};
constexpr int kKnnStack = 96;  // Karras depth <= 63 key bits + 32 tie-break bits

// k nearest so far: binary max-heap on the squared distance
template <int KMAX>
struct KnnHeap {
    double d2[KMAX];
    uint32_t idx[KMAX];
    uint32_t size, k;
    RPTX_D double worst() const { return size < k ? CUDART_INF : d2[0]; }
    RPTX_D void push(double d, uint32_t id) {
        if (size < k) {
            uint32_t c = size++;
            while (c > 0) {
                const uint32_t p = (c - 1) >> 1;
                if (!(d2[p] < d)) break;
                d2[c] = d2[p]; idx[c] = idx[p];
                c = p;
            }
            d2[c] = d; idx[c] = id;
        } else if (d < d2[0]) {
            sift_down(0, size, d, id);
        }
    }
    RPTX_D void sift_down(uint32_t c, uint32_t n, double d, uint32_t id) {
        for (;;) {
            uint32_t l = 2 * c + 1;
            if (l >= n) break;
            if (l + 1 < n && d2[l + 1] > d2[l]) l++;
            if (!(d2[l] > d)) break;
            d2[c] = d2[l]; idx[c] = idx[l];
            c = l;
        }
        d2[c] = d; idx[c] = id;
    }
    // heap -> ascending order in place (nearests() returns ascending)
    RPTX_D void sort_ascending() {
        for (uint32_t n = size; n > 1; n--) {
            const double d = d2[n - 1];
            const uint32_t id = idx[n - 1];
            d2[n - 1] = d2[0]; idx[n - 1] = idx[0];
            sift_down(0, n - 1, d, id);
        }
    }
};

constexpr double kBoxShrink = 0.99999999999999;  // rounding slack of the box lower bound

RPTX_D double box_dist2(float4 mn, float4 mx, d3 q) {
    const double dx = fmax(fmax((double)mn.x - q.x, q.x - (double)mx.x), 0.0);
    const double dy = fmax(fmax((double)mn.y - q.y, q.y - (double)mx.y), 0.0);
    const double dz = fmax(fmax((double)mn.z - q.z, q.z - (double)mx.z), 0.0);
    return dx * dx + dy * dy + dz * dz;
}

// KdTree::nearests(q, k): exact k nearest by squared Euclidean distance
template <int KMAX>
RPTX_DN void knn_query(const PointMapDev& m, d3 q, KnnHeap<KMAX>& h) {
    h.size = 0;
    if (m.n == 0 || h.k == 0) return;
    int stack[kKnnStack];
    double stack_d[kKnnStack];
    int sp = 0;
    int cur = m.root_ref;
    for (;;) {
        while (cur >= 0) {
            const float4* np = reinterpret_cast<const float4*>(m.nodes + cur);
            const float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2), d = __ldg(np + 3);
            const double d0 = box_dist2(a, b, q), d1 = box_dist2(c, d, q);
            const int left = __float_as_int(a.w), right = __float_as_int(c.w);
            const double w = h.worst();
            const bool v0 = d0 * kBoxShrink <= w, v1 = d1 * kBoxShrink <= w;
            if (v0 && v1) {
                const bool first0 = d0 <= d1;
                RPTX_ASSERT(sp < kKnnStack);
                stack[sp] = first0 ? right : left;
                stack_d[sp] = first0 ? d1 : d0;
                sp++;
                cur = first0 ? left : right;
            } else if (v0) {
                cur = left;
            } else if (v1) {
                cur = right;
            } else {
                cur = 0x7fffffff;  // nothing to visit below this node
                break;
            }
        }
        if (cur < 0) {
            const uint32_t code = (uint32_t)(~cur);
            const uint32_t first = code >> 3, last = first + (code & 7u) + 1u;
            for (uint32_t i = first; i < last; i++) {
                const double dx = __ldg(m.px + i) - q.x, dy = __ldg(m.py + i) - q.y, dz = __ldg(m.pz + i) - q.z;
                h.push(dx * dx + dy * dy + dz * dz, i);
            }
        }
        for (;;) {
            if (sp == 0) return;
            sp--;
            if (stack_d[sp] * kBoxShrink <= h.worst()) { cur = stack[sp]; break; }
        }
    }
}

// ------------------------------------------------------------------ box BVH (bvh 0.6.0)

struct Aabb32 {
    float mn[3], mx[3];
};

// bvh::ray::Ray::new + intersects_aabb (Williams et al. slab test in f32)
struct Ray32 {
    float o[3], inv[3];
    int sign[3];
};
RPTX_D Ray32 make_ray32(d3 origin, d3 dir) {
    Ray32 r;
    float d[3] = {(float)dir.x, (float)dir.y, (float)dir.z};
    const float len = sqrtf(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
    r.o[0] = (float)origin.x; r.o[1] = (float)origin.y; r.o[2] = (float)origin.z;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        d[i] = d[i] / len;
        r.inv[i] = 1.0f / d[i];
        r.sign[i] = d[i] < 0.0f ? 1 : 0;
    }
    return r;
}
// LINE = true: the whole line through the ray, both directions (beam segments: the reference's beam test has no t > 0
// condition, a closest approach behind the camera counts as long as the ray hits the beam's own box in front of it)
template <bool LINE = false>
RPTX_D bool ray32_hits(const Ray32& r, const float* mn, const float* mx) {
    float ray_min = ((r.sign[0] ? mx[0] : mn[0]) - r.o[0]) * r.inv[0];
    float ray_max = ((r.sign[0] ? mn[0] : mx[0]) - r.o[0]) * r.inv[0];
    const float y_min = ((r.sign[1] ? mx[1] : mn[1]) - r.o[1]) * r.inv[1];
    const float y_max = ((r.sign[1] ? mn[1] : mx[1]) - r.o[1]) * r.inv[1];
    if (ray_min > y_max || y_min > ray_max) return false;
    if (y_min > ray_min) ray_min = y_min;
    if (y_max < ray_max) ray_max = y_max;
    const float z_min = ((r.sign[2] ? mx[2] : mn[2]) - r.o[2]) * r.inv[2];
    const float z_max = ((r.sign[2] ? mn[2] : mx[2]) - r.o[2]) * r.inv[2];
    if (ray_min > z_max || z_min > ray_max) return false;
    if (z_max < ray_max) ray_max = z_max;
    return LINE || ray_max > 0.0f;
}

struct BoxBvhDev {
    uint32_t n;
    int root_ref;
    const BvhNode* nodes;  // child boxes are exact f32 unions of the item boxes (pad 0)
    const Aabb32* boxes;   // per item, leaf order
};

// PhotonSphere (src/photon.rs:62-73), leaf order, structure-of-arrays
struct SphereMapDev {
    BoxBvhDev bvh;
    const double* center;  // [n][3]
    const double* radius;  // [n]
    const double* dir;     // [n][3]
    const double* power;   // [n][3]
};
// PhotonBeam (src/photon.rs:46-59), leaf order
// Culling structure of the warp camera pass over photon beams.  A beam's own box (the reference's, src/photon.rs:76-105) is
// the box of a long thin cylinder: at 1e8 photons a camera ray hits the boxes of 10 000 of the 2e5 beams and 97 % of them
// pass nowhere near the beam.  The beams are therefore cut into segments of ~kBeamSegLen (16 radii), each with the box of ITS piece of the
// cylinder, under a second BVH.  A segment stands for the contributions whose closest-approach parameter beam_t falls into its
// piece - floor(beam_t * K / len) == k, a partition of [0, len], so every contribution is counted exactly once - and the ray
// that makes such a contribution passes within the radius of that piece, i.e. through the segment's box.  The beam's own box
// is still tested (the reference never sees a beam whose box the ray misses); the result is the per-beam walk's.
// measured at 1e8 photons (2e5 beams), camera pass ms at segment lengths 4 / 8 / 12 / 24 / 48 / 96 / 192: 207 / 197 / 192 / 179 /
// 170 / 179 / 224 (one box per beam: 700); RPTX_BEAM_SEG_LEN overrides
constexpr double kBeamSegLen = 48.0;
constexpr uint32_t kBeamMaxSeg = 64;
struct BeamSegDev {
    BoxBvhDev bvh;         // over the segments; bvh.boxes = segment boxes, leaf order
    const uint32_t* beam;  // [n] beam (BeamMapDev index) of the segment
    const uint16_t* k;     // [n] piece of the beam ...
    const uint16_t* K;     // [n] ... of so many
};

struct BeamMapDev {
    BeamSegDev seg;        // seg.bvh.n == 0: not built
    BoxBvhDev bvh;
    const double* start;  // [n][3]
    const double* dirn;   // [n][3] normalised start -> end
    const double* len;    // [n] |end - start|
    const double* power;  // [n][3]
};

// BVH::traverse: call f(item) for every item whose own f32 box the ray hits.  Inner boxes are
// exact unions, and the slab test is monotone under enclosure, so culling by them loses nothing.
template <class F>
RPTX_D void box_bvh_traverse(const BoxBvhDev& b, const Ray32& r, F&& f) {
    if (b.n == 0) return;
    if (b.n == 1) { f(0u); return; }  // a single-leaf root is returned untested
    int stack[kStackSize];
    int sp = 0;
    int cur = b.root_ref;
    for (;;) {
        if (cur >= 0) {
            const float4* np = reinterpret_cast<const float4*>(b.nodes + cur);
            const float4 a = __ldg(np), bb = __ldg(np + 1), c = __ldg(np + 2), d = __ldg(np + 3);
            const float lmn[3] = {a.x, a.y, a.z}, lmx[3] = {bb.x, bb.y, bb.z}, rmn[3] = {c.x, c.y, c.z}, rmx[3] = {d.x, d.y, d.z};
            const bool hl = ray32_hits(r, lmn, lmx), hr = ray32_hits(r, rmn, rmx);
            const int left = __float_as_int(a.w), right = __float_as_int(c.w);
            if (hl && hr) { RPTX_ASSERT(sp < kStackSize); stack[sp++] = right; cur = left; continue; }
            if (hl) { cur = left; continue; }
            if (hr) { cur = right; continue; }
        } else {
            const uint32_t code = (uint32_t)(~cur);
            const uint32_t first = code >> 3, count = (code & 7u) + 1u;
            for (uint32_t i = first; i < first + count; i++) {
                const Aabb32 bx = b.boxes[i];
                if (ray32_hits(r, bx.mn, bx.mx)) f(i);
            }
        }
        if (sp == 0) return;
        cur = stack[--sp];
    }
}

// ------------------------------------------------------------------ estimators

struct PhotonMapDev {
    int kind;  // PhotonRenderKind
    PointMapDev surface, volume;
    SphereMapDev spheres;
    BeamMapDev beams;
    uint32_t gather_size, gather_size_volume;
};

RPTX_D double k2_kernel(double sq) {  // src/photon.rs:466-469
    const double tmp = 1.0 - sq;
    return (3.0 / kPi) * tmp * tmp;
}
RPTX_D double transmittance(const MediumDev& md, double t) { return exp(-(md.extinction() * t)); }  // src/medium.rs:126-130

struct GatherCounters {
    unsigned long long knn, visibility, candidates;
};

// ---- candidate tests of the beam estimators (volume_estimate, src/photon.rs:439-593), shared by the per-thread and the
// warp-cooperative camera pass.  *_cull: conservative f32 rejection (true = cannot contribute); *_exact: the reference's f64 test
// and contribution.
struct VolumeQuery {
    RayD ray;
    bool hit;
    double t_hit, extinction;
    d3 mc;
    float fox, foy, foz, fdx, fdy, fdz;  // unit direction (primary_ray normalises)
};

RPTX_D VolumeQuery make_volume_query(const MediumDev& md, const RayD& ray, bool hit, double t_hit) {
    VolumeQuery q;
    q.ray = ray; q.hit = hit; q.t_hit = t_hit;
    q.extinction = md.extinction();
    q.mc = md.color(mk3(0, 0, 0));
    q.fox = (float)ray.o.x; q.foy = (float)ray.o.y; q.foz = (float)ray.o.z;
    q.fdx = (float)ray.d.x; q.fdy = (float)ray.d.y; q.fdz = (float)ray.d.z;
    return q;
}

// The sphere contributes only if its centre lies within `radius` of the ray's line.  perp^2 = |c - o|^2 - ((c - o).d)^2 in
// f32 carries a relative error of a few 2^-23 of |c - o|^2; the margin is 16x that, and the exact f64 test decides everything
// inside it.
RPTX_D bool sphere_cull(const SphereMapDev& sm, const VolumeQuery& q, uint32_t i) {
    const float lx = (float)sm.center[3 * (size_t)i] - q.fox, ly = (float)sm.center[3 * (size_t)i + 1] - q.foy,
                lz = (float)sm.center[3 * (size_t)i + 2] - q.foz;
    const float l2 = lx * lx + ly * ly + lz * lz;
    const float s = lx * q.fdx + ly * q.fdy + lz * q.fdz;
    const float rf = (float)sm.radius[i];
    return l2 - s * s > rf * rf * 1.0001f + l2 * 4e-6f;
}

RPTX_D void sphere_exact(const SphereMapDev& sm, const MediumDev& md, const VolumeQuery& q, uint32_t i, d3* volume_color) {
    const d3 c = mk3(sm.center[3 * (size_t)i], sm.center[3 * (size_t)i + 1], sm.center[3 * (size_t)i + 2]);
    const d3 origin_to_center = c - q.ray.o;
    if (q.hit && length(origin_to_center) > q.t_hit) return;
    const double radius = sm.radius[i];
    const double radius_squared = radius * radius;
    const double disk_distance = dot(origin_to_center, q.ray.d);
    const double dn = length((q.ray.o + disk_distance * q.ray.d) - c);
    const double distance_squared = dn * dn;
    if (disk_distance > 0.0 && distance_squared < radius_squared) {
        const double weight = k2_kernel(distance_squared / radius_squared) / radius_squared;
        const double tr = exp(-q.extinction * disk_distance);
        const d3 ppow = mk3(sm.power[3 * (size_t)i], sm.power[3 * (size_t)i + 1], sm.power[3 * (size_t)i + 2]);
        *volume_color = *volume_color + tr * cmul(ppow, q.mc) * md.phase * weight * 1.0;
    }
}

// piece of [0, len] that beam_t falls into, of K equal pieces (both ends of [0, len] included)
RPTX_HD uint32_t beam_segment_of(double beam_t, double len, uint32_t K) {
    const double x = beam_t * (double)K / len;
    const uint32_t k = x > 0.0 ? (uint32_t)x : 0u;  // NaN (len == 0) -> 0
    return k < K ? k : K - 1u;
}

// `dist` of the exact test is the distance between a point of the camera ray's line and a point of the beam's line, so it is
// at least the distance between the two lines, |l . (d x b)| / |d x b|.  Beams whose line passes farther than the radius (with
// a margin of 16x the f32 rounding of the triple product, relative to |l| |d x b|) contribute nothing; nearly parallel lines
// and everything inside the margin go to the exact f64 test.
RPTX_D bool beam_cull(const BeamMapDev& bm, const VolumeQuery& q, uint32_t i) {
    const size_t i3 = 3 * (size_t)i;
    const float bx = (float)bm.dirn[i3], by = (float)bm.dirn[i3 + 1], bz = (float)bm.dirn[i3 + 2];
    const float nx = q.fdy * bz - q.fdz * by, ny = q.fdz * bx - q.fdx * bz, nz = q.fdx * by - q.fdy * bx;
    const float n2 = nx * nx + ny * ny + nz * nz;
    const float lx = (float)bm.start[i3] - q.fox, ly = (float)bm.start[i3 + 1] - q.foy, lz = (float)bm.start[i3 + 2] - q.foz;
    const float trip = lx * nx + ly * ny + lz * nz;
    const float l2 = lx * lx + ly * ly + lz * lz;
    const float lim = (float)kBeamRadius * 1.0001f;
    return n2 > 1e-6f && trip * trip > (lim * lim + l2 * 4e-6f) * n2;
}

// seg_K > 1: only the contribution whose beam_t lies in piece seg_k of seg_K (BeamSegDev)
RPTX_D void beam_exact(const BeamMapDev& bm, const MediumDev& md, const VolumeQuery& q, uint32_t i, d3* volume_color, uint32_t seg_k = 0,
                       uint32_t seg_K = 1) {
    const size_t i3 = 3 * (size_t)i;
    const d3 bstart = mk3(bm.start[i3], bm.start[i3 + 1], bm.start[i3 + 2]);
    const d3 bdir = mk3(bm.dirn[i3], bm.dirn[i3 + 1], bm.dirn[i3 + 2]);
    const d3 l = bstart - q.ray.o;
    const d3 u = normalize(cross(l, bdir));
    const d3 n = normalize(cross(bdir, u));
    const double t = dot(n, l) / dot(n, q.ray.d);
    const d3 query_collision = q.ray.o + t * q.ray.d;
    if (q.hit && t >= q.t_hit) return;
    const double dd = dot(q.ray.d, bdir);
    const double inv_sin_theta = 1.0 / sqrt(fmax(0.0, 1.0 - dd * dd));
    const double beam_t = dot(bdir, query_collision - bstart);
    const double beam_len = bm.len[i];
    if (beam_t < 0.0 || beam_t > beam_len) return;
    if (seg_K > 1u && beam_segment_of(beam_t, beam_len, seg_K) != seg_k) return;
    const d3 beam_collision = bstart + beam_t * bdir;
    const double dist = length(query_collision - beam_collision);
    if (dist >= kBeamRadius) return;
    const d3 ppow = mk3(bm.power[i3], bm.power[i3 + 1], bm.power[i3 + 2]);
    const d3 color = q.extinction * cmul(ppow, q.mc) * md.phase * inv_sin_theta * transmittance(md, t) * transmittance(md, beam_t) *
                     k2_kernel(dist / kBeamRadius) / (2.0 * kBeamRadius);
    *volume_color = *volume_color + color;
}

// surface_estimate, src/photon.rs:327-375
template <int KMAX>
RPTX_DN d3 surface_estimate(const SceneDev& sc, const PhotonMapDev& pm, const RayD& ray, d3 wo, double t_hit, d3 normal,
                           const Material& mat, GatherCounters& ctr) {
    const d3 world_pos = ray.o + t_hit * ray.d;
    KnnHeap<KMAX> heap;
    heap.k = pm.gather_size;
    knn_query(pm.surface, world_pos, heap);
    heap.sort_ascending();
    ctr.knn++;
    double max_d2 = 0.0;
    for (uint32_t j = 0; j < heap.size; j++) max_d2 = fmax(max_d2, heap.d2[j]);
    d3 color = mat_emittance(mat) * mat_color(mat);
    for (uint32_t j = 0; j < heap.size; j++) {
        const uint32_t i = heap.idx[j];
        const d3 ppos = mk3(pm.surface.px[i], pm.surface.py[i], pm.surface.pz[i]);
        const d3 disp = world_pos - ppos;
        RayD vr;
        vr.o = ppos;
        vr.d = normalize(disp);
        const double len = length(disp);
        // the closest hit must not be nearer than the shading point: search below `len`, stop at the first find
        Hit vh;
        vh.t = len;
        closest_hit<kAnyBelow>(sc, vr, kEpsilon, len, vh);
        ctr.visibility++;
        if (vh.inst != kNoHit && len > vh.t) continue;
        const d3 pdir = mk3(pm.surface.dir[3 * (size_t)i], pm.surface.dir[3 * (size_t)i + 1], pm.surface.dir[3 * (size_t)i + 2]);
        const d3 ppow = mk3(pm.surface.power[3 * (size_t)i], pm.surface.power[3 * (size_t)i + 1], pm.surface.power[3 * (size_t)i + 2]);
        color = color + cmul(mat_bsdf(mat, normal, wo, pdir), ppow) * clampd(dot(pdir, normal), 0.0, 1.0);
    }
    return color * (1.0 / (kPi * max_d2));
}

// volume_estimate, src/photon.rs:379-594.  `hit` false = the camera ray left the scene.
template <int KMAX>
RPTX_DN d3 volume_estimate(const SceneDev& sc, const MediumDev& md, const PhotonMapDev& pm, const RayD& ray, d3 wo, bool hit, double t_hit,
                          d3 normal, const Material& mat, Rng& rng, GatherCounters& ctr) {
    const double extinction = md.extinction();
    if (pm.kind == 0) {
        const double random = rng.gen_range(0.0, 1.0);
        const double d = -log(random) / extinction;
        const double tr = exp(-(extinction * d));
        const double d_pdf = extinction * tr, d_cdf = 1.0 - tr;
        if (!hit || d < t_hit) {
            const d3 collision = ray.o + d * ray.d;
            const d3 mc = md.color(collision);
            KnnHeap<KMAX> heap;
            heap.k = pm.gather_size_volume;
            knn_query(pm.volume, collision, heap);
            heap.sort_ascending();
            ctr.knn++;
            double max_d2 = 0.0;
            for (uint32_t j = 0; j < heap.size; j++) max_d2 = fmax(max_d2, heap.d2[j]);
            d3 color = mk3(0, 0, 0);
            for (uint32_t j = 0; j < heap.size; j++) {
                const size_t i = heap.idx[j];
                const d3 ppow = mk3(pm.volume.power[3 * i], pm.volume.power[3 * i + 1], pm.volume.power[3 * i + 2]);
                color = color + cmul(ppow, mc) * md.phase;
            }
            color = color / ((4.0 / 3.0) * kPi * pow(max_d2, 1.5));
            color = color / extinction;
            color = color * transmittance(md, d);
            color = color / d_pdf;
            return color;
        }
        return surface_estimate<KMAX>(sc, pm, ray, wo, t_hit, normal, mat, ctr) * transmittance(md, t_hit) / (1.0 - d_cdf);
    }
    const Ray32 r32 = make_ray32(ray.o, ray.d);
    const VolumeQuery vq = make_volume_query(md, ray, hit, t_hit);
    d3 volume_color = mk3(0, 0, 0);
    if (pm.kind == 1) {
        const SphereMapDev& sm = pm.spheres;
        box_bvh_traverse(sm.bvh, r32, [&](uint32_t i) {
            ctr.candidates++;
            if (sphere_cull(sm, vq, i)) return;
            sphere_exact(sm, md, vq, i, &volume_color);
        });
        return volume_color;
    }
    const BeamMapDev& bm = pm.beams;
    box_bvh_traverse(bm.bvh, r32, [&](uint32_t i) {
        ctr.candidates++;
        if (beam_cull(bm, vq, i)) return;
        beam_exact(bm, md, vq, i, &volume_color);
    });
    return volume_color;
}

// estimate_indirect, src/photon.rs:316-627 (dispatch at :596-627)
template <int KMAX>
RPTX_D d3 estimate_indirect(const SceneDev& sc, const MediumDev& md, const PhotonMapDev& pm, const RayD& ray, Rng& rng,
                            GatherCounters& ctr) {
    const d3 wo = -normalize(ray.d);
    Hit h;
    h.t = CUDART_INF;
    closest_hit<kClosest>(sc, ray, kEpsilon, 0.0, h);
    const bool medium = md.kind != 0;
    Material mat;
    mat.kind = 0; mat.albedo[0] = mat.albedo[1] = mat.albedo[2] = 0.0; mat.emittance = 0.0; mat.shininess = 0.0; mat.ior = 1.0;
    const bool hit = h.inst != kNoHit;
    d3 n = mk3(0, 0, 0);
    if (hit) {
        uint32_t prim, inst;
        n = hit_normal(sc, ray, h, &prim, &inst);
        mat = sc.materials[h.obj];
    }
    if (!medium) return hit ? surface_estimate<KMAX>(sc, pm, ray, wo, h.t, n, mat, ctr) : env_color(sc, ray.d);
    const d3 volume_color = volume_estimate<KMAX>(sc, md, pm, ray, wo, hit, hit ? h.t : CUDART_INF, n, mat, rng, ctr);
    if (!hit || pm.kind == 0) return volume_color;  // point-point: the surface term is inside (src/photon.rs:612-615)
    const d3 surface_color = surface_estimate<KMAX>(sc, pm, ray, wo, h.t, n, mat, ctr) * transmittance(md, h.t);
    return surface_color + volume_color;
}

// ------------------------------------------------------------------ warp-cooperative gather
// The per-thread kNN above spends its time in heap sift loops over local memory with 2-3 of 32
// lanes active (profiles/r1_v2_summary.md).  For the point-point estimator (k = 30..100) a whole
// warp owns one camera sample instead: all lanes walk the tree together (uniform control flow),
// the photons of a leaf (at most kPointLeaf = 8: the leaf reference keeps the count in 3 bits) are tested one per
// lane - so a leaf visit uses up to 8 of the 32 lanes; what the warp buys is the uniform tree walk, the
// shared candidate list and the visibility rays - accepted candidates are appended to a shared-memory
// list by ballot, and when the list is full a bitonic sort keeps the k best and tightens the radius.
// The k visibility rays of the surface estimate then run 32 at a time.
constexpr int kWarpCand = 256;       // candidate list capacity per warp (>= k + 32)
constexpr uint32_t kWarpKnnMaxK = 128;

struct WarpKnnSmem {
    double d2[kWarpCand];
    uint32_t idx[kWarpCand];
    double stack_d[kKnnStack];
    uint32_t stack[kKnnStack];
};

// Beam estimators by a warp (warp_volume_boxes): the lanes share ONE camera ray and walk the sphere / beam BVH together.
// The per-thread walk ran 6-9 of 32 lanes (ncu, profiles/r2_summary.md): neighbouring samples sit in different parts of
// node walk / leaf loop / f32 cull / f64 test.  Here every stage is its own dense loop over a shared list:
//   nodes:  a shared LIFO of node references; 32 are popped at a time, one per lane; inner nodes push the children the ray
//           hits, leaves append their items to `cand`;
//   cand:   items whose own box and f32 cull are still to be tested, 32 at a time; survivors go to `exact`;
//   exact:  the reference's f64 test and contribution, 32 at a time; per-lane partial sums, added over the warp at the end.
constexpr int kWarpBoxStack = 1536;   // node references (a pop of 32 pushes at most 64; near the top the walk pops one at a time)
constexpr int kWarpBoxCand = 320;     // < 32 left over + 32 leaves x 8 items
constexpr int kWarpBoxExact = 64;     // < 32 left over + 32 survivors
struct WarpBoxSmem {
    int stack[kWarpBoxStack];
    uint32_t cand[kWarpBoxCand];
    uint32_t exact[kWarpBoxExact];
};
union WarpGatherSmem {  // the kNN of the surface estimate and the box walk never overlap in time
    WarpKnnSmem knn;
    WarpBoxSmem box;
};

// ascending bitonic sort of the first n (power of two, <= kWarpCand) candidates by d2
RPTX_D void warp_sort_candidates(WarpKnnSmem& sm, uint32_t n, uint32_t lane) {
    for (uint32_t size = 2; size <= n; size <<= 1) {
        for (uint32_t stride = size >> 1; stride > 0; stride >>= 1) {
            for (uint32_t t = lane; t < (n >> 1); t += 32) {
                const uint32_t lo = 2 * t - (t & (stride - 1));  // index of the lower partner
                const uint32_t hi = lo + stride;
                const bool up = (lo & size) == 0;
                const double a = sm.d2[lo], b = sm.d2[hi];
                if ((a > b) == up) {
                    sm.d2[lo] = b; sm.d2[hi] = a;
                    const uint32_t ia = sm.idx[lo];
                    sm.idx[lo] = sm.idx[hi]; sm.idx[hi] = ia;
                }
            }
            __syncwarp();
        }
    }
}

// keep the k smallest of the `count` candidates (sorted ascending); returns the new count
RPTX_D uint32_t warp_compact_candidates(WarpKnnSmem& sm, uint32_t count, uint32_t k, uint32_t lane) {
    uint32_t n = 2;
    while (n < count) n <<= 1;
    for (uint32_t i = count + lane; i < n; i += 32) { sm.d2[i] = CUDART_INF; sm.idx[i] = 0xffffffffu; }
    __syncwarp();
    warp_sort_candidates(sm, n, lane);
    return count < k ? count : k;
}

// KdTree::nearests(q, k) by a whole warp; every lane passes the same q and k (k <= kWarpKnnMaxK).
// On return sm.d2 / sm.idx hold the result in ascending order; returns its size.
RPTX_D uint32_t warp_knn(const PointMapDev& m, d3 q, uint32_t k, WarpKnnSmem& sm) {
    const uint32_t lane = threadIdx.x & 31u;
    if (m.n == 0 || k == 0) return 0;
    uint32_t count = 0, sp = 0;
    int cur = m.root_ref;
    double worst = CUDART_INF;
    for (;;) {
        while (cur >= 0) {
            const float4* np = reinterpret_cast<const float4*>(m.nodes + cur);
            const float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2), d = __ldg(np + 3);
            const double d0 = box_dist2(a, b, q), d1 = box_dist2(c, d, q);
            const int left = __float_as_int(a.w), right = __float_as_int(c.w);
            const bool v0 = d0 * kBoxShrink <= worst, v1 = d1 * kBoxShrink <= worst;
            if (v0 && v1) {
                const bool first0 = d0 <= d1;
                RPTX_ASSERT(sp < (uint32_t)kKnnStack);
                if (lane == 0) { sm.stack[sp] = (uint32_t)(first0 ? right : left); sm.stack_d[sp] = first0 ? d1 : d0; }
                sp++;
                cur = first0 ? left : right;
            } else if (v0) {
                cur = left;
            } else if (v1) {
                cur = right;
            } else {
                cur = 0x7fffffff;
                break;
            }
        }
        if (cur < 0) {
            if (count + 32 > (uint32_t)kWarpCand) {
                count = warp_compact_candidates(sm, count, k, lane);
                if (count == k) worst = sm.d2[k - 1];
            }
            const uint32_t code = (uint32_t)(~cur);
            const uint32_t first = code >> 3, n_here = (code & 7u) + 1u;
            const uint32_t i = first + lane;
            double dd = CUDART_INF;
            if (lane < n_here) {
                const double dx = __ldg(m.px + i) - q.x, dy = __ldg(m.py + i) - q.y, dz = __ldg(m.pz + i) - q.z;
                dd = dx * dx + dy * dy + dz * dz;
            }
            const bool take = lane < n_here && dd < worst;
            const unsigned mask = __ballot_sync(0xffffffffu, take);
            if (take) {
                const uint32_t at = count + __popc(mask & ((1u << lane) - 1u));
                RPTX_ASSERT(at < (uint32_t)kWarpCand);
                sm.d2[at] = dd; sm.idx[at] = i;
            }
            count += __popc(mask);
        }
        __syncwarp();
        for (;;) {
            if (sp == 0) return warp_compact_candidates(sm, count, k, lane);
            sp--;
            if (sm.stack_d[sp] * kBoxShrink <= worst) { cur = (int)sm.stack[sp]; break; }
        }
        __syncwarp();  // every lane has read the popped entry before lane 0 may overwrite the slot
    }
}

RPTX_D d3 warp_sum(d3 v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
        v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
        v.z += __shfl_xor_sync(0xffffffffu, v.z, o);
    }
    return v;
}

// surface_estimate (src/photon.rs:327-375) by a warp: lanes share the kNN, then take the photons'
// visibility rays 32 at a time.  All lanes return the same colour.
RPTX_D d3 warp_surface_estimate(const SceneDev& sc, const PhotonMapDev& pm, const RayD& ray, d3 wo, double t_hit, d3 normal,
                                const Material& mat, WarpKnnSmem& sm, GatherCounters& ctr) {
    const uint32_t lane = threadIdx.x & 31u;
    const d3 world_pos = ray.o + t_hit * ray.d;
    const uint32_t count = warp_knn(pm.surface, world_pos, pm.gather_size, sm);
    if (lane == 0) ctr.knn++;
    const double max_d2 = count ? fmax(0.0, sm.d2[count - 1]) : 0.0;
    d3 part = mk3(0, 0, 0);
    for (uint32_t j = lane; j < count; j += 32) {
        const uint32_t i = sm.idx[j];
        const d3 ppos = mk3(pm.surface.px[i], pm.surface.py[i], pm.surface.pz[i]);
        const d3 disp = world_pos - ppos;
        RayD vr;
        vr.o = ppos;
        vr.d = normalize(disp);
        const double len = length(disp);
        Hit vh;
        vh.t = len;
        closest_hit<kAnyBelow>(sc, vr, kEpsilon, len, vh);
        ctr.visibility++;
        if (vh.inst != kNoHit && len > vh.t) continue;
        const d3 pdir = mk3(pm.surface.dir[3 * (size_t)i], pm.surface.dir[3 * (size_t)i + 1], pm.surface.dir[3 * (size_t)i + 2]);
        const d3 ppow = mk3(pm.surface.power[3 * (size_t)i], pm.surface.power[3 * (size_t)i + 1], pm.surface.power[3 * (size_t)i + 2]);
        part = part + cmul(mat_bsdf(mat, normal, wo, pdir), ppow) * clampd(dot(pdir, normal), 0.0, 1.0);
    }
    __syncwarp();
    const d3 color = mat_emittance(mat) * mat_color(mat) + warp_sum(part);
    return color * (1.0 / (kPi * max_d2));
}

// estimate_indirect for the point-point map (PhotonRenderKind::PhotonMap) by a warp
RPTX_D d3 warp_estimate_point_point(const SceneDev& sc, const MediumDev& md, const PhotonMapDev& pm, const RayD& ray, Rng& rng,
                                    WarpKnnSmem& sm, GatherCounters& ctr) {
    const uint32_t lane = threadIdx.x & 31u;
    const d3 wo = -normalize(ray.d);
    Hit h;
    h.t = CUDART_INF;
    closest_hit<kClosest>(sc, ray, kEpsilon, 0.0, h);
    const bool medium = md.kind != 0;
    const bool hit = h.inst != kNoHit;
    Material mat;
    mat.kind = 0; mat.albedo[0] = mat.albedo[1] = mat.albedo[2] = 0.0; mat.emittance = 0.0; mat.shininess = 0.0; mat.ior = 1.0;
    d3 n = mk3(0, 0, 0);
    if (hit) {
        uint32_t prim, inst;
        n = hit_normal(sc, ray, h, &prim, &inst);
        mat = sc.materials[h.obj];
    }
    if (!medium) return hit ? warp_surface_estimate(sc, pm, ray, wo, h.t, n, mat, sm, ctr) : env_color(sc, ray.d);
    // volume_estimate, point-point branch (src/photon.rs:386-437)
    const double extinction = md.extinction();
    const double random = rng.gen_range(0.0, 1.0);
    const double d = -log(random) / extinction;
    const double tr = exp(-(extinction * d));
    const double d_pdf = extinction * tr, d_cdf = 1.0 - tr;
    if (!hit || d < h.t) {
        const d3 collision = ray.o + d * ray.d;
        const d3 mc = md.color(collision);
        const uint32_t count = warp_knn(pm.volume, collision, pm.gather_size_volume, sm);
        if (lane == 0) ctr.knn++;
        const double max_d2 = count ? fmax(0.0, sm.d2[count - 1]) : 0.0;
        d3 part = mk3(0, 0, 0);
        for (uint32_t j = lane; j < count; j += 32) {
            const size_t i = sm.idx[j];
            const d3 ppow = mk3(pm.volume.power[3 * i], pm.volume.power[3 * i + 1], pm.volume.power[3 * i + 2]);
            part = part + cmul(ppow, mc) * md.phase;
        }
        __syncwarp();
        d3 color = warp_sum(part);
        color = color / ((4.0 / 3.0) * kPi * pow(max_d2, 1.5));
        color = color / extinction;
        color = color * transmittance(md, d);
        color = color / d_pdf;
        return color;
    }
    return warp_surface_estimate(sc, pm, ray, wo, h.t, n, mat, sm, ctr) * transmittance(md, h.t) / (1.0 - d_cdf);
}

// volume_estimate for the beam-query maps (kind 1: photon spheres, kind 2: photon beams) by a warp.  Same candidates, same
// tests and same contributions as the per-thread walk (BVH::traverse semantics: an item is a candidate when the ray hits its
// own f32 box); only the order of the additions differs.  All lanes return the same colour.
RPTX_D d3 warp_volume_boxes(const MediumDev& md, const PhotonMapDev& pm, const RayD& ray, bool hit, double t_hit, WarpBoxSmem& sm,
                            GatherCounters& ctr) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const bool segs = pm.kind == 2 && pm.beams.seg.bvh.n != 0;  // beams: walk the segments (BeamSegDev)
    const BoxBvhDev& b = pm.kind == 1 ? pm.spheres.bvh : (segs ? pm.beams.seg.bvh : pm.beams.bvh);
    const VolumeQuery vq = make_volume_query(md, ray, hit, t_hit);
    const Ray32 r = make_ray32(ray.o, ray.d);
    d3 part = mk3(0, 0, 0);
    if (b.n == 0) return part;
    uint32_t top = 0, ncand = 0, nexact = 0;  // warp-uniform
    // a one-item BVH is returned untested (box_bvh_traverse); with segments that is the one BEAM's own box
    const bool untested_single = segs ? pm.beams.bvh.n == 1 : b.n == 1;
    if (lane == 0) sm.stack[0] = b.root_ref;
    top = 1;
    __syncwarp();
    auto run_exact = [&](uint32_t n) {  // the last n entries of `exact`, one per lane
        if (lane < n) {
            const uint32_t i = sm.exact[nexact - n + lane];
            if (pm.kind == 1) sphere_exact(pm.spheres, md, vq, i, &part);
            else if (segs) beam_exact(pm.beams, md, vq, pm.beams.seg.beam[i], &part, pm.beams.seg.k[i], pm.beams.seg.K[i]);
            else beam_exact(pm.beams, md, vq, i, &part);
        }
        nexact -= n;
        __syncwarp();
    };
    auto run_cand = [&](uint32_t n) {  // the last n entries of `cand`: own box + f32 cull, survivors -> exact
        bool keep = false;
        uint32_t i = 0;
        if (lane < n) {
            i = sm.cand[ncand - n + lane];
            const Aabb32 bx = b.boxes[i];
            if (segs) {
                // the segment's box prunes; the beam's own box is the reference's candidate test
                if (ray32_hits<true>(r, bx.mn, bx.mx)) {
                    const uint32_t beam = pm.beams.seg.beam[i];
                    const Aabb32 own = pm.beams.bvh.boxes[beam];
                    if (untested_single || ray32_hits(r, own.mn, own.mx)) {
                        ctr.candidates++;
                        keep = !beam_cull(pm.beams, vq, beam);
                    }
                }
            } else if (untested_single || ray32_hits(r, bx.mn, bx.mx)) {
                ctr.candidates++;
                keep = pm.kind == 1 ? !sphere_cull(pm.spheres, vq, i) : !beam_cull(pm.beams, vq, i);
            }
        }
        ncand -= n;
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        RPTX_ASSERT(nexact + 32u <= (uint32_t)kWarpBoxExact);
        if (keep) sm.exact[nexact + __popc(m & lt)] = i;
        nexact += __popc(m);
        __syncwarp();
        if (nexact >= 32u) run_exact(32u);
    };
    while (top > 0 || ncand > 0) {
        if (ncand >= 32u || top == 0) { run_cand(ncand < 32u ? ncand : 32u); continue; }
        // pop: 32 references, or one at a time near the top of the stack (a depth-first walk then adds at most one entry
        // per level of the subtree it is in)
        const uint32_t n = top + 128u > (uint32_t)kWarpBoxStack ? 1u : (top < 32u ? top : 32u);
        top -= n;
        int ref = 0;
        bool have = lane < n;
        if (have) ref = sm.stack[top + lane];
        __syncwarp();
        bool hl = false, hr = false;
        int left = 0, right = 0;
        uint32_t first = 0, count = 0;
        if (have && ref >= 0) {
            const float4* np = reinterpret_cast<const float4*>(b.nodes + ref);
            const float4 a = __ldg(np), bb = __ldg(np + 1), c = __ldg(np + 2), d = __ldg(np + 3);
            const float lmn[3] = {a.x, a.y, a.z}, lmx[3] = {bb.x, bb.y, bb.z}, rmn[3] = {c.x, c.y, c.z}, rmx[3] = {d.x, d.y, d.z};
            hl = segs ? ray32_hits<true>(r, lmn, lmx) : ray32_hits(r, lmn, lmx);
            hr = segs ? ray32_hits<true>(r, rmn, rmx) : ray32_hits(r, rmn, rmx);
            left = __float_as_int(a.w); right = __float_as_int(c.w);
        } else if (have) {
            const uint32_t code = (uint32_t)(~ref);
            first = code >> 3; count = (code & 7u) + 1u;
        }
        // children: the right ones first, so that the left child of the last lane is on top (the per-thread walk's order)
        const unsigned mr = __ballot_sync(0xffffffffu, hr);
        RPTX_ASSERT(top + 64u <= (uint32_t)kWarpBoxStack || n == 1u);
        RPTX_ASSERT(top + 2u <= (uint32_t)kWarpBoxStack);
        if (hr) sm.stack[top + __popc(mr & lt)] = right;
        top += __popc(mr);
        const unsigned ml = __ballot_sync(0xffffffffu, hl);
        if (hl) sm.stack[top + __popc(ml & lt)] = left;
        top += __popc(ml);
        // leaf items
#pragma unroll 1
        for (uint32_t k = 0; k < 8u; k++) {
            const unsigned mk = __ballot_sync(0xffffffffu, count > k);
            if (mk == 0) break;
            RPTX_ASSERT(ncand + 32u <= (uint32_t)kWarpBoxCand);
            if (count > k) sm.cand[ncand + __popc(mk & lt)] = first + k;
            ncand += __popc(mk);
        }
        __syncwarp();
    }
    if (nexact) run_exact(nexact);
    return warp_sum(part);
}

// estimate_indirect (src/photon.rs:596-627) by a warp, all three maps
RPTX_D d3 warp_estimate_indirect(const SceneDev& sc, const MediumDev& md, const PhotonMapDev& pm, const RayD& ray, Rng& rng,
                                 WarpGatherSmem& sm, GatherCounters& ctr) {
    if (pm.kind == 0) return warp_estimate_point_point(sc, md, pm, ray, rng, sm.knn, ctr);
    const d3 wo = -normalize(ray.d);
    Hit h;
    h.t = CUDART_INF;
    closest_hit<kClosest>(sc, ray, kEpsilon, 0.0, h);
    const bool medium = md.kind != 0;
    const bool hit = h.inst != kNoHit;
    Material mat;
    mat.kind = 0; mat.albedo[0] = mat.albedo[1] = mat.albedo[2] = 0.0; mat.emittance = 0.0; mat.shininess = 0.0; mat.ior = 1.0;
    d3 n = mk3(0, 0, 0);
    if (hit) {
        uint32_t prim, inst;
        n = hit_normal(sc, ray, h, &prim, &inst);
        mat = sc.materials[h.obj];
    }
    if (!medium) return hit ? warp_surface_estimate(sc, pm, ray, wo, h.t, n, mat, sm.knn, ctr) : env_color(sc, ray.d);
    const d3 volume_color = warp_volume_boxes(md, pm, ray, hit, hit ? h.t : CUDART_INF, sm.box, ctr);
    if (!hit) return volume_color;
    __syncwarp();
    const d3 surface_color = warp_surface_estimate(sc, pm, ray, wo, h.t, n, mat, sm.knn, ctr) * transmittance(md, h.t);
    return surface_color + volume_color;
}

struct GatherArgs {
    Camera cam;
    BatchParams bp;
    uint32_t n_pixels, n_lanes;
    double *ax, *ay, *az;  // [n_lanes][n_pixels] lane sums of the batch
};

// Camera pass (get_color_with_photon_map, src/photon.rs:950-985): one thread per (pixel, lane)
// walks its share of the batch's samples and keeps their sum; k_batch_reduce then adds the lanes
// in a fixed order.
template <int KMAX>
__global__ void __launch_bounds__(128) k_photon_gather(const __grid_constant__ SceneDev sc, const MediumDev md,
                                                       const __grid_constant__ PhotonMapDev pm, const __grid_constant__ GatherArgs ga,
                                                       unsigned long long* __restrict__ stats) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    GatherCounters ctr = {0, 0, 0};
    unsigned long long my_samples = 0;
    if (slot < ga.n_pixels * ga.n_lanes) {
        const uint32_t pixel = slot % ga.n_pixels, lane = slot / ga.n_pixels;
        const uint32_t x = pixel % ga.bp.width, y = pixel / ga.bp.width;
        d3 sum = mk3(0, 0, 0);
        for (uint32_t j = lane; j < ga.bp.n_local; j += ga.n_lanes) {
            const uint64_t s = ga.bp.batch_base + ga.bp.shard_index + (uint64_t)j * ga.bp.shard_count;
            Rng rng = Rng::make(ga.bp.seed, kDomainGather, pixel, s);
            RayD ray;
            primary_ray(ga.cam, ga.bp.width, ga.bp.height, x, y, rng, &ray.o, &ray.d);
            sum = sum + estimate_indirect<KMAX>(sc, md, pm, ray, rng, ctr);
            my_samples++;
        }
        ga.ax[slot] = sum.x; ga.ay[slot] = sum.y; ga.az[slot] = sum.z;
    }
    const cg::coalesced_group g = cg::coalesced_threads();
    const unsigned long long a = cg::reduce(g, my_samples, cg::plus<unsigned long long>());
    const unsigned long long b = cg::reduce(g, ctr.knn, cg::plus<unsigned long long>());
    const unsigned long long c = cg::reduce(g, ctr.visibility, cg::plus<unsigned long long>());
    const unsigned long long d = cg::reduce(g, ctr.candidates, cg::plus<unsigned long long>());
    if (g.thread_rank() == 0) {
        if (a) atomicAdd(&stats[0], a);
        if (b) atomicAdd(&stats[5], b);
        if (c) atomicAdd(&stats[6], c);
        if (d) atomicAdd(&stats[7], d);
    }
}

// Camera pass by warps (all three maps): one warp per (pixel, lane) slot walks that slot's samples.
__global__ void __launch_bounds__(128, RPTX_OCC_GATHER) k_photon_gather_warp(const __grid_constant__ SceneDev sc, const MediumDev md,
                                                            const __grid_constant__ PhotonMapDev pm, const __grid_constant__ GatherArgs ga,
                                                            unsigned long long* __restrict__ stats) {
    __shared__ WarpGatherSmem smem[4];
    WarpGatherSmem& sm = smem[threadIdx.x >> 5];
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t slot = blockIdx.x * 4 + (threadIdx.x >> 5);
    GatherCounters ctr = {0, 0, 0};
    unsigned long long my_samples = 0;
    if (slot < ga.n_pixels * ga.n_lanes) {
        const uint32_t pixel = slot % ga.n_pixels, slane = slot / ga.n_pixels;
        const uint32_t x = pixel % ga.bp.width, y = pixel / ga.bp.width;
        d3 sum = mk3(0, 0, 0);
        for (uint32_t j = slane; j < ga.bp.n_local; j += ga.n_lanes) {
            const uint64_t s = ga.bp.batch_base + ga.bp.shard_index + (uint64_t)j * ga.bp.shard_count;
            Rng rng = Rng::make(ga.bp.seed, kDomainGather, pixel, s);
            RayD ray;
            primary_ray(ga.cam, ga.bp.width, ga.bp.height, x, y, rng, &ray.o, &ray.d);
            sum = sum + warp_estimate_indirect(sc, md, pm, ray, rng, sm, ctr);
            if (lane == 0) my_samples++;
            __syncwarp();
        }
        if (lane == 0) { ga.ax[slot] = sum.x; ga.ay[slot] = sum.y; ga.az[slot] = sum.z; }
    }
    __syncwarp();
    const unsigned long long a = __reduce_add_sync(0xffffffffu, (unsigned)my_samples);
    const unsigned long long b = __reduce_add_sync(0xffffffffu, (unsigned)ctr.knn);
    const unsigned long long c = __reduce_add_sync(0xffffffffu, (unsigned)ctr.visibility);
    unsigned long long d = ctr.candidates;  // may pass 2^32 per warp at 1e8 photons
    for (int o = 16; o > 0; o >>= 1) d += __shfl_down_sync(0xffffffffu, d, o);
    if (lane == 0) {
        if (a) { atomicAdd(&stats[0], a); atomicAdd(&stats[1], a); }
        if (b) atomicAdd(&stats[5], b);
        if (c) atomicAdd(&stats[6], c);
        if (d) atomicAdd(&stats[7], d);
    }
}

}  // namespace rptx

// Closest-hit query of the B200 core: replaces Renderer::get_closest_hit
// (src/renderer.rs:416-425), every Shape::intersect (src/shape/*.rs, src/shape.rs:129-138) and
// the kd-tree traversal (src/kdtree.rs:132-226).
//
// Two-level BVH (instances -> triangles).  Box tests run in f32 on outward-padded boxes with
// the ray origin first advanced (in f64) to the box entry point, so they are conservative;
// every primitive test runs in f64 with the reference's own operation order, so hit/miss,
// primitive id and t reproduce the reference bit for bit.  Ties on t resolve to the lower
// object index, as the reference's strict `<` over its ordered object loop does.
#pragma once
#include "common.cuh"

namespace rptx {

struct RayD {
    d3 o, d;
};

struct Hit {
    double t;       // +inf (or the caller's t_max) while nothing was accepted
    uint32_t inst;  // index into SceneDev::instances, kNoHit = none
    uint32_t obj;   // object id of `inst`
    uint32_t prim;  // MESH: triangle slot (BLAS order) relative to tri_base; CUBE: normal code
    double v, w;    // MESH: barycentrics of the accepted hit
};

struct RayF {
    float ox, oy, oz, ix, iy, iz;
};

constexpr float kSlabSlack = 1e-6f;
constexpr int kStackSize = 64;

RPTX_D float f32_up(double x) { return __double2float_ru(x); }

RPTX_D RayF make_rayf(d3 o, d3 d) {
    RayF r;
    r.ox = (float)o.x; r.oy = (float)o.y; r.oz = (float)o.z;
    r.ix = 1.0f / (float)d.x; r.iy = 1.0f / (float)d.y; r.iz = 1.0f / (float)d.z;
    return r;
}

// conservative f32 slab test; boxes are padded at build time (see bvh_build.cu)
RPTX_D bool slab_f32(float bx0, float by0, float bz0, float bx1, float by1, float bz1, const RayF& r, float tbest,
                     float* tnear) {
    const float t0x = (bx0 - r.ox) * r.ix, t1x = (bx1 - r.ox) * r.ix;
    const float t0y = (by0 - r.oy) * r.iy, t1y = (by1 - r.oy) * r.iy;
    const float t0z = (bz0 - r.oz) * r.iz, t1z = (bz1 - r.oz) * r.iz;
    float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    tn = tn - fabsf(tn) * kSlabSlack;
    tf = tf + fabsf(tf) * kSlabSlack;
    *tnear = tn;
    return tn <= tf && tf >= 0.0f && tn <= tbest;
}

// f64 entry/exit of a ray against a padded box (fmin/fmax drop NaNs from 0/0)
RPTX_D void clip_f64(const double* bmin, const double* bmax, d3 o, d3 d, double* t0, double* t1) {
    const double ax = (bmin[0] - o.x) / d.x, bx = (bmax[0] - o.x) / d.x;
    const double ay = (bmin[1] - o.y) / d.y, by = (bmax[1] - o.y) / d.y;
    const double az = (bmin[2] - o.z) / d.z, bz = (bmax[2] - o.z) / d.z;
    *t0 = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmin(az, bz));
    *t1 = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmax(az, bz));
}

// Generic BVH2 traversal.  leaf(first, count, tbest) tests primitives [first, first+count),
// may lower tbest, returns true to stop the whole traversal.
template <class LeafFn>
RPTX_D bool bvh_traverse(const BvhNode* __restrict__ nodes, int root_ref, const RayF& r, float& tbest, LeafFn&& leaf) {
    int stack[kStackSize];
    float stack_t[kStackSize];
    int sp = 0;
    int cur = root_ref;
    for (;;) {
        if (cur < 0) {
            const uint32_t code = (uint32_t)(~cur);
            if (leaf(code >> 3, (code & 7u) + 1u, tbest)) return true;
        } else {
            const float4* np = reinterpret_cast<const float4*>(nodes + cur);
            const float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2), d = __ldg(np + 3);
            float tl, tr;
            const bool hl = slab_f32(a.x, a.y, a.z, b.x, b.y, b.z, r, tbest, &tl);
            const bool hr = slab_f32(c.x, c.y, c.z, d.x, d.y, d.z, r, tbest, &tr);
            const int left = __float_as_int(a.w), right = __float_as_int(c.w);
            if (hl && hr) {
                const bool lf = tl <= tr;
                stack[sp] = lf ? right : left;
                stack_t[sp] = lf ? tr : tl;
                sp++;
                cur = lf ? left : right;
                continue;
            } else if (hl) { cur = left; continue; }
            else if (hr) { cur = right; continue; }
        }
        // pop
        for (;;) {
            if (sp == 0) return false;
            sp--;
            if (stack_t[sp] <= tbest) { cur = stack[sp]; break; }
        }
    }
}

RPTX_D bool accept(const Hit& h, double t, uint32_t obj) { return t < h.t || (t == h.t && obj < h.obj); }

// Triangle::intersect, src/shape/mesh.rs:50-83, on the prepared record
RPTX_D bool tri_test(const TriRec& tr, d3 o, d3 d, double t_min, double t_max, bool allow_eq, double* t, double* vv, double* ww) {
    const d3 pn = mk3(tr.pn[0], tr.pn[1], tr.pn[2]);
    const double cosine = dot(pn, d);
    if (fabs(cosine) < 1e-8) return false;
    const d3 v1 = mk3(tr.v1[0], tr.v1[1], tr.v1[2]);
    const double time = dot(pn, v1 - o) / cosine;
    if (time < t_min || !(time < t_max || (allow_eq && time == t_max))) return false;
    const d3 d0 = mk3(tr.d0[0], tr.d0[1], tr.d0[2]), d1 = mk3(tr.d1[0], tr.d1[1], tr.d1[2]);
    const d3 d2 = (o + time * d) - v1;
    const double d20 = dot(d2, d0), d21 = dot(d2, d1);
    const double v = (tr.d11 * d20 - tr.d01 * d21) / tr.denom;
    const double w = (tr.d00 * d21 - tr.d01 * d20) / tr.denom;
    const double u = 1.0 - v - w;
    if (u >= 0.0 && v >= 0.0 && w >= 0.0) { *t = time; *vv = v; *ww = w; return true; }
    return false;
}

RPTX_D void load_tri(const TriRec* __restrict__ p, TriRec* out) {
    const double2* s = reinterpret_cast<const double2*>(p);
    double2* q = reinterpret_cast<double2*>(out);
#pragma unroll
    for (int i = 0; i < 8; i++) q[i] = __ldg(s + i);
}

// Sphere::intersect, src/shape/sphere.rs:14-46
RPTX_D bool sphere_test(d3 o, d3 d, double t_min, double* t) {
    const double a = dot(d, d);
    const double b = dot(d, o);
    const double c = dot(o, o) - 1.0;
    double disc = b * b - a * c;
    if (signbit(disc)) return false;
    disc = sqrt(disc);
    const double t_minus = (-b - disc) / a;
    if (t_minus < t_min) {
        const double t_plus = (-b + disc) / a;
        if (t_plus < t_min) return false;
        *t = t_plus;
    } else {
        *t = t_minus;
    }
    return true;
}

// Cube::intersect, src/shape/cube.rs:22-74; code = axis*2 + (normal component > 0)
RPTX_D bool cube_test(d3 o, d3 d, double t_min, double* t, uint32_t* code) {
    double lo[3], hi[3];
    uint32_t lc[3], hc[3];
#pragma unroll
    for (int dim = 0; dim < 3; dim++) {
        const double oo = comp(o, dim), dd = comp(d, dim);
        double x1 = (-0.5 - oo) / dd;
        double x2 = (0.5 - oo) / dd;
        uint32_t c1 = dim * 2, c2 = dim * 2 + 1;
        if (x1 > x2) { const double s = x1; x1 = x2; x2 = s; const uint32_t cs = c1; c1 = c2; c2 = cs; }
        lo[dim] = x1; hi[dim] = x2; lc[dim] = c1; hc[dim] = c2;
    }
    double start, end;
    uint32_t sc, ec;
    if (lo[0] > lo[1] && lo[0] > lo[2]) { start = lo[0]; sc = lc[0]; }
    else if (lo[1] > lo[2]) { start = lo[1]; sc = lc[1]; }
    else { start = lo[2]; sc = lc[2]; }
    if (hi[0] < hi[1] && hi[0] < hi[2]) { end = hi[0]; ec = hc[0]; }
    else if (hi[1] < hi[2]) { end = hi[1]; ec = hc[1]; }
    else { end = hi[2]; ec = hc[2]; }
    if (start > end || end < t_min) return false;
    if (start < t_min) { *t = end; *code = ec; } else { *t = start; *code = sc; }
    return true;
}

// Plane::intersect, src/shape/plane.rs:17-32
RPTX_D bool plane_test(const Instance& in, d3 o, d3 d, double t_min, double* t) {
    const d3 n = mk3(in.plane_n[0], in.plane_n[1], in.plane_n[2]);
    const double cosine = dot(n, d);
    if (fabs(cosine) < 1e-8) return false;
    const double time = (in.plane_v - dot(n, o)) / cosine;
    if (time >= t_min) { *t = time; return true; }
    return false;
}

// One instance against the world ray.  SHADOW: stop as soon as a hit below t_stop is accepted.
template <bool SHADOW>
RPTX_D bool intersect_instance(const SceneDev& sc, uint32_t idx, const RayD& ray, double t_min, double t_stop, Hit& hit) {
    const Instance* __restrict__ in = sc.instances + idx;
    const int kind = in->kind;
    const uint32_t obj = in->object_id;
    d3 lo = ray.o, ld = ray.d;
    if (in->has_transform) {
        lo = xform_point(in->inv, ray.o);
        ld = xform_dir(in->inv, ray.d);
    }
    if (kind == kMesh) {
        const TriRec* __restrict__ tris = sc.tris + in->tri_base;
        bool stop = false;
        auto leaf = [&](uint32_t first, uint32_t count, float& tbest) -> bool {
            for (uint32_t i = first; i < first + count; i++) {
                TriRec tr;
                load_tri(tris + i, &tr);
                double t, v, w;
                const bool eq_ok = obj < hit.obj;
                if (tri_test(tr, lo, ld, t_min, hit.t, eq_ok, &t, &v, &w)) {
                    hit.t = t; hit.inst = idx; hit.obj = obj; hit.prim = i; hit.v = v; hit.w = w;
                    if (SHADOW && t < t_stop) { stop = true; return true; }
                }
            }
            return false;
        };
        if (in->root_ref < 0) {  // tiny mesh: no BLAS nodes
            float dummy = 0.f;
            const uint32_t code = (uint32_t)(~in->root_ref);
            leaf(code >> 3, (code & 7u) + 1u, dummy);
            return stop;
        }
        double t0, t1;
        clip_f64(in->lmin, in->lmax, lo, ld, &t0, &t1);
        if (!(t0 <= t1) || t1 < 0.0 || t0 > hit.t) return false;
        const double shift = t0 > 0.0 ? t0 : 0.0;
        const RayF rf = make_rayf(lo + shift * ld, ld);
        float tbest = f32_up(hit.t - shift);
        tbest = tbest + fabsf(tbest) * kSlabSlack;
        auto leaf2 = [&](uint32_t first, uint32_t count, float& tb) -> bool {
            const double before = hit.t;
            const bool s = leaf(first, count, tb);
            if (hit.t != before) { tb = f32_up(hit.t - shift); tb = tb + fabsf(tb) * kSlabSlack; }
            return s;
        };
        bvh_traverse(sc.nodes + in->node_base, in->root_ref, rf, tbest, leaf2);
        return stop;
    }
    double t;
    uint32_t code = 0;
    bool ok;
    if (kind == kSphere) ok = sphere_test(lo, ld, t_min, &t);
    else if (kind == kCube) ok = cube_test(lo, ld, t_min, &t, &code);
    else ok = plane_test(*in, lo, ld, t_min, &t);
    if (ok && accept(hit, t, obj)) {
        hit.t = t; hit.inst = idx; hit.obj = obj; hit.prim = code;
        if (SHADOW && t < t_stop) return true;
    }
    return false;
}

// get_closest_hit.  hit.t must be preset to the search bound (+inf for extend rays).
template <bool SHADOW>
RPTX_D void closest_hit(const SceneDev& sc, const RayD& ray, double t_min, double t_stop, Hit& hit) {
    hit.inst = kNoHit; hit.obj = kNoHit; hit.prim = 0; hit.v = 0; hit.w = 0;
    // unbounded shapes (planes) cannot enter the BVH (src/renderer.rs:413-415)
    for (uint32_t i = 0; i < sc.n_unbounded; i++)
        if (intersect_instance<SHADOW>(sc, sc.n_bounded + i, ray, t_min, t_stop, hit)) return;
    if (sc.n_bounded == 0) return;
    double t0, t1;
    clip_f64(sc.wmin, sc.wmax, ray.o, ray.d, &t0, &t1);
    if (!(t0 <= t1) || t1 < 0.0 || t0 > hit.t) return;
    const double shift = t0 > 0.0 ? t0 : 0.0;
    const RayF rf = make_rayf(ray.o + shift * ray.d, ray.d);
    float tbest = f32_up(hit.t - shift);
    tbest = tbest + fabsf(tbest) * kSlabSlack;
    auto leaf = [&](uint32_t first, uint32_t count, float& tb) -> bool {
        for (uint32_t i = first; i < first + count; i++) {
            const double before = hit.t;
            if (intersect_instance<SHADOW>(sc, i, ray, t_min, t_stop, hit)) return true;
            if (hit.t != before) { tb = f32_up(hit.t - shift); tb = tb + fabsf(tb) * kSlabSlack; }
        }
        return false;
    };
    bvh_traverse(sc.nodes, sc.tlas_root_ref, rf, tbest, leaf);
}

// normal of the accepted hit (the per-shape normal code of the reference + Transformed's M^-T)
RPTX_D d3 hit_normal(const SceneDev& sc, const RayD& ray, const Hit& hit, uint32_t* prim_orig, uint32_t* inst_id) {
    const Instance* __restrict__ in = sc.instances + hit.inst;
    d3 lo = ray.o, ld = ray.d;
    if (in->has_transform) {
        lo = xform_point(in->inv, ray.o);
        ld = xform_dir(in->inv, ray.d);
    }
    d3 n;
    *prim_orig = 0;
    *inst_id = in->inst_id;
    if (in->kind == kMesh) {
        const TriNormals* __restrict__ tn = sc.tri_normals + in->tri_base + hit.prim;
        const double u = 1.0 - hit.v - hit.w;
        const d3 n1 = mk3(tn->n1[0], tn->n1[1], tn->n1[2]), n2 = mk3(tn->n2[0], tn->n2[1], tn->n2[2]),
                 n3 = mk3(tn->n3[0], tn->n3[1], tn->n3[2]);
        n = normalize(u * n1 + hit.v * n2 + hit.w * n3);
        *prim_orig = sc.tri_orig[in->tri_base + hit.prim];
    } else if (in->kind == kSphere) {
        n = normalize(lo + hit.t * ld);
    } else if (in->kind == kCube) {
        const uint32_t axis = hit.prim >> 1;
        const double s = (hit.prim & 1u) ? 1.0 : -1.0;
        n = mk3(axis == 0 ? s : 0.0, axis == 1 ? s : 0.0, axis == 2 ? s : 0.0);
    } else {
        const d3 pn = mk3(in->plane_n[0], in->plane_n[1], in->plane_n[2]);
        const double cosine = dot(pn, ld);
        const double sg = signbit(cosine) ? -1.0 : 1.0;
        n = (-normalize(pn)) * sg;
    }
    if (in->has_transform) n = normalize(mul3(in->nrm, n));
    return n;
}

}  // namespace rptx

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
// Pending entries of the two-level walk: a BVH2 walk holds at most one entry per level.  PLOC trees are accepted up to 56
// levels (rptx.cu: kPlocMaxDepth), host SAH trees are at most 36 deep (bvh_sah.hpp), the Karras LBVH over 30-bit Morton
// codes + 32-bit index tie-break (bvh_build.cu) at most 62: TLAS 62 + sentinel + root + BLAS 62 = 126.  Only touched
// entries cost cache lines.
constexpr int kStackSize = 128;

RPTX_D float f32_up(double x) { return __double2float_ru(x); }

RPTX_D RayF make_rayf(d3 o, d3 d) {
    RayF r;
    r.ox = (float)o.x; r.oy = (float)o.y; r.oz = (float)o.z;
    r.ix = 1.0f / (float)d.x; r.iy = 1.0f / (float)d.y; r.iz = 1.0f / (float)d.z;
    return r;
}

// conservative f32 slab test; boxes are padded at build time (see bvh_build.cu)
RPTX_D bool slab_f32(float bx0, float by0, float bz0, float bx1, float by1, float bz1, const RayF& r, float tbest,
                     float* tnear, float tlow = 0.0f) {
    const float t0x = (bx0 - r.ox) * r.ix, t1x = (bx1 - r.ox) * r.ix;
    const float t0y = (by0 - r.oy) * r.iy, t1y = (by1 - r.oy) * r.iy;
    const float t0z = (bz0 - r.oz) * r.iz, t1z = (bz1 - r.oz) * r.iz;
    float tn = fmaxf(fmaxf(fminf(t0x, t1x), fminf(t0y, t1y)), fminf(t0z, t1z));
    float tf = fminf(fminf(fmaxf(t0x, t1x), fmaxf(t0y, t1y)), fmaxf(t0z, t1z));
    tn = tn - fabsf(tn) * kSlabSlack;
    tf = tf + fabsf(tf) * kSlabSlack;
    *tnear = tn;
    return tn <= tf && tf >= tlow && tn <= tbest;
}

// f64 entry/exit of a ray against a padded box (fmin/fmax drop NaNs from 0 * inf).  Only culling and the start of
// the f32 ray depend on it and the boxes are padded by 2^-21 of their extent, so one reciprocal per axis (1 ulp off
// the quotient) serves both planes: three divisions instead of six per box entry.
RPTX_D void clip_f64(const double* bmin, const double* bmax, d3 o, d3 d, double* t0, double* t1) {
    const double ix = DDIV(1.0, d.x), iy = DDIV(1.0, d.y), iz = DDIV(1.0, d.z);
    const double ax = (bmin[0] - o.x) * ix, bx = (bmax[0] - o.x) * ix;
    const double ay = (bmin[1] - o.y) * iy, by = (bmax[1] - o.y) * iy;
    const double az = (bmin[2] - o.z) * iz, bz = (bmax[2] - o.z) * iz;
    *t0 = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmin(az, bz));
    *t1 = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmax(az, bz));
}

// Ties on t: the reference walks its objects in index order and every shape but one replaces the record only for a
// strictly smaller t, so the lowest index keeps a tie.  MonomialSurface::intersect rejects only r > record.time
// (src/shape/monomial_surface.rs:84): a monomial surface TAKES a tie from whoever holds it and, once it holds the
// record, keeps it against every later object (two monomials: the later index wins).  A monomial holder is marked by
// Hit::prim == kMonoPrim.
constexpr uint32_t kMonoPrim = 0x80000000u;
RPTX_D bool accept(const Hit& h, double t, uint32_t obj) { return t < h.t || (t == h.t && obj < h.obj && h.prim != kMonoPrim); }
RPTX_D bool accept_mono(const Hit& h, double t, uint32_t obj) { return t < h.t || (t == h.t && (h.prim != kMonoPrim || obj > h.obj)); }

// Triangle::intersect, src/shape/mesh.rs:50-83, on the prepared record
RPTX_D bool tri_test(const TriRec& tr, d3 o, d3 d, double t_min, double t_max, bool allow_eq, double* t, double* vv, double* ww) {
    const d3 pn = mk3(tr.pn[0], tr.pn[1], tr.pn[2]);
    const double cosine = dot(pn, d);
    if (fabs(cosine) < 1e-8) return false;
    const d3 v1 = mk3(tr.v1[0], tr.v1[1], tr.v1[2]);
    const double time = fdiv(dot(pn, v1 - o), cosine);
    if (time < t_min || !(time < t_max || (allow_eq && time == t_max))) return false;
    const d3 d0 = mk3(tr.d0[0], tr.d0[1], tr.d0[2]), d1 = mk3(tr.d1[0], tr.d1[1], tr.d1[2]);
    const d3 d2 = (o + time * d) - v1;
    const double d20 = dot(d2, d0), d21 = dot(d2, d1);
    const double v = DDIV(tr.d11 * d20 - tr.d01 * d21, tr.denom);
    const double w = DDIV(tr.d00 * d21 - tr.d01 * d20, tr.denom);
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
    disc = DSQRT(disc);
    const double t_minus = DDIV(-b - disc, a);
    if (t_minus < t_min) {
        const double t_plus = DDIV(-b + disc, a);
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
        double x1 = fdiv(-0.5 - oo, dd);
        double x2 = fdiv(0.5 - oo, dd);
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
    const double time = DDIV(in.plane_v - dot(n, o), cosine);
    if (time >= t_min) { *t = time; return true; }
    return false;
}

// MonomialSurface::intersect, src/shape/monomial_surface.rs:22-107: y = h (x^2 + z^2)^2 over the unit disc.
// Same operation order as the reference (powi(2), powi(3) are multiplication chains); t_best is the
// record's current time (the reference rejects r > record.time).
RPTX_DN bool monomial_test(double h, d3 o, d3 d, double t_min, double t_best, double* t_out) {
    // BoundingBox::intersect (src/kdtree.rs:57-71) on (-1,0,-1)-(1,h,1); fmin/fmax drop NaNs like f64::min/max
    double b_min, b_max;
    {
        const double x1 = (-1.0 - o.x) / d.x, x2 = (1.0 - o.x) / d.x;
        const double y1 = (0.0 - o.y) / d.y, y2 = (h - o.y) / d.y;
        const double z1 = (-1.0 - o.z) / d.z, z2 = (1.0 - o.z) / d.z;
        b_min = fmax(fmax(fmin(x1, x2), fmin(y1, y2)), fmin(z1, z2));
        b_max = fmin(fmin(fmax(x1, x2), fmax(y1, y2)), fmax(z1, z2));
    }
    if (fmax(b_min, t_min) > fmin(b_max, t_best)) return false;
    auto dist = [&](double t) {
        const double x = o.x + t * d.x, y = o.y + t * d.y, z = o.z + t * d.z;
        const double s = x * x + z * z;
        return y - h * (s * s);
    };
    const double coef0 = o.x * o.x + o.z * o.z;
    const double coef1 = 2. * (o.x * d.x + o.z * d.z);
    const double coef2 = d.x * d.x + d.z * d.z;
    double t_max;
    const bool maximize = dist(t_min) < 0.0;
    if (maximize) {
        double cur = (b_min + b_max) / 2.;
        for (int i = 0; i < 10; i++) {
            if (dist(cur) > 0.) break;
            const double der = d.y - h * (2. * coef0 * coef1 + 2. * cur * (coef1 * coef1 + 2. * coef0 * coef2) +
                                          3. * (cur * cur) * 2. * coef1 * coef2 + 4. * (cur * cur * cur) * coef2 * coef2);
            const double der2 = -h * (2. * (coef1 * coef1 + 2. * coef0 * coef2) + 3. * 2. * cur * 2. * coef1 * coef2 +
                                      4. * 3. * (cur * cur) * coef2 * coef2);
            cur -= der / der2;
        }
        t_max = cur;
        if (t_max < t_min) return false;
    } else {
        t_max = 10000.;
    }
    if ((dist(t_min) < 0.0) == (dist(t_max) < 0.0)) return false;
    double l = t_min, r = t_max;
    for (int i = 0; i < 60; i++) {
        const double m = (l + r) / 2.0;
        if ((dist(m) >= 0.0) == maximize) r = m;
        else l = m;
    }
    if (r > t_best) return false;
    const double px = o.x + r * d.x, pz = o.z + r * d.z;
    if (px * px + pz * pz > 1.0) return false;
    *t_out = r;
    return true;
}

// Conservative f32 Moller-Trumbore pre-test.  Returns false only when the ray provably misses
// the triangle: every comparison carries an absolute error bound built from L1 norms of the
// operands (an f32 dot/cross of vectors with relative input error ~2^-23 is off by at most a
// few 2^-23 * |a||b|; kPreEps = 4e-6 leaves > 10x margin).  Undecidable cases go to the exact
// f64 test.
constexpr float kPreEps = 4e-6f;
RPTX_D float l1(float x, float y, float z) { return fabsf(x) + fabsf(y) + fabsf(z); }
RPTX_D bool tri_maybe_hit_v(const float4 v0, const float4 e1, const float4 e2, float ox, float oy, float oz, float dx, float dy, float dz) {
    const float px = dy * e2.z - dz * e2.y, py = dz * e2.x - dx * e2.z, pz = dx * e2.y - dy * e2.x;  // d x e2
    const float det = e1.x * px + e1.y * py + e1.z * pz;
    const float tx = ox - v0.x, ty = oy - v0.y, tz = oz - v0.z;
    const float nd = l1(dx, dy, dz), ne1 = l1(e1.x, e1.y, e1.z), ne2 = l1(e2.x, e2.y, e2.z);
    const float mag = l1(ox, oy, oz) + l1(v0.x, v0.y, v0.z);  // bounds |tvec| and its rounding error
    const float np = 2.0f * nd * ne2;                          // >= |d x e2| (L1)
    const float e_u = kPreEps * np * mag;
    const float e_det = kPreEps * ne1 * np;
    const float u = tx * px + ty * py + tz * pz;
    const float qx = ty * e1.z - tz * e1.y, qy = tz * e1.x - tx * e1.z, qz = tx * e1.y - ty * e1.x;  // tvec x e1
    const float v = dx * qx + dy * qy + dz * qz;
    const float e_v = kPreEps * nd * 2.0f * ne1 * mag;
    const float sgn = det < 0.0f ? -1.0f : 1.0f;
    const float adet = fabsf(det);
    const float us = u * sgn, vs = v * sgn;
    if (us < -e_u || vs < -e_v) return false;
    if (us + vs > adet + e_det + e_u + e_v) return false;
    return true;  // includes NaN/inf cases (all comparisons false)
}
RPTX_D bool tri_maybe_hit(const TriF* __restrict__ tf, float ox, float oy, float oz, float dx, float dy, float dz) {
    const float4* q = reinterpret_cast<const float4*>(tf);
    return tri_maybe_hit_v(__ldg(q), __ldg(q + 1), __ldg(q + 2), ox, oy, oz, dx, dy, dz);
}

// KdTree::intersect first clips the ray against the tree's bounding box with NaN-ignoring
// f64::min/max (src/kdtree.rs:57-71,133-137).  A ray lying exactly in a face plane of that box
// (direction component 0, origin on the plane) divides 0/0 there; when the box has extent in
// that dimension the slab collapses to (+inf,+inf) or (-inf,-inf) and the reference rejects the
// whole mesh — although triangles touching that plane (a wall's bottom edge seen from inside the
// floor) are geometrically hit.  Returns true when the reference would reject.
RPTX_D bool mesh_bounds_nan_reject(const Instance* __restrict__ in, d3 lo, d3 ld) {
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const double d = comp(ld, k), o = comp(lo, k);
        if (d == 0.0 && in->bmax[k] > in->bmin[k] && (o == in->bmin[k] || o == in->bmax[k])) return true;
    }
    return false;
}

// get_closest_hit over the two-level BVH as ONE loop (while-while): every lane of a warp runs
// the same inner-node code whether it is in the instance BVH or inside a mesh BVH; entering a
// mesh pushes a sentinel, swaps the f32 ray for the mesh-local one and continues.
// hit.t must be preset to the search bound (+inf for extend rays).
// MODE kAnyBelow: stop as soon as a hit below t_stop is accepted (occlusion query).
// MODE kWindow: closest REPORTED hit with t >= t_stop (a primitive reports the t its reference
// intersect() would, with the usual t_min; reports below t_stop are ignored); nodes that end
// before t_stop are culled, so for a narrow window this is a point query.
constexpr int kSentinel = 0x7fffffff;
enum TraceMode : int { kClosest = 0, kAnyBelow = 1, kWindow = 2 };

template <int MODE>
RPTX_D void closest_hit(const SceneDev& sc, const RayD& ray, double t_min, double t_stop, Hit& hit) {
    constexpr bool SHADOW = MODE == kAnyBelow;
    constexpr bool WINDOW = MODE == kWindow;
    hit.inst = kNoHit; hit.obj = kNoHit; hit.prim = 0; hit.v = 0; hit.w = 0;
    // unbounded shapes (planes) cannot enter the BVH (src/renderer.rs:413-415)
    for (uint32_t i = 0; i < sc.n_unbounded; i++) {
        const uint32_t idx = sc.n_bounded + i;
        const Instance* __restrict__ in = sc.instances + idx;
        d3 lo = ray.o, ld = ray.d;
        if (in->has_transform) { lo = xform_point(in->inv, ray.o); ld = xform_dir(in->inv, ray.d); }
        double t;
        if (plane_test(*in, lo, ld, t_min, &t) && (!WINDOW || t >= t_stop) && accept(hit, t, in->object_id)) {
            hit.t = t; hit.inst = idx; hit.obj = in->object_id; hit.prim = 0;
            if (SHADOW && t < t_stop) return;
        }
    }
    if (sc.n_bounded == 0) return;
    // Cube::intersect divides 0/0 for a ray lying exactly in a face plane (src/shape/cube.rs:24-25);
    // the NaN then drops that slab from the test and a hit can be reported where the box is
    // geometrically missed — which the BVH would cull.  Such rays (a direction component is exactly
    // zero; e.g. photon visibility rays inside an axis-aligned wall) test those cubes directly.
    if (sc.n_cube_instances && (ray.d.x == 0.0 || ray.d.y == 0.0 || ray.d.z == 0.0)) {
        for (uint32_t c = 0; c < sc.n_cube_instances; c++) {
            const uint32_t idx = sc.cube_instances[c];
            const Instance* __restrict__ in = sc.instances + idx;
            d3 lo = ray.o, ld = ray.d;
            if (in->has_transform) { lo = xform_point(in->inv, ray.o); ld = xform_dir(in->inv, ray.d); }
            const bool in_face = (ld.x == 0.0 && fabs(lo.x) == 0.5) || (ld.y == 0.0 && fabs(lo.y) == 0.5) || (ld.z == 0.0 && fabs(lo.z) == 0.5);
            if (!in_face) continue;
            double t;
            uint32_t ccode = 0;
            if (cube_test(lo, ld, t_min, &t, &ccode) && (!WINDOW || t >= t_stop) && accept(hit, t, in->object_id)) {
                hit.t = t; hit.inst = idx; hit.obj = in->object_id; hit.prim = ccode;
                if (SHADOW && t < t_stop) return;
            }
        }
    }
    double t0, t1;
    clip_f64(sc.wmin, sc.wmax, ray.o, ray.d, &t0, &t1);
    if (!(t0 <= t1) || t1 < 0.0 || t0 > hit.t) return;
    if (WINDOW && t1 < t_stop) return;
    // the f32 ray starts at the box entry (or at the window, whichever is later)
    const double wshift = WINDOW ? fmax(fmax(t0, t_stop), 0.0) : (t0 > 0.0 ? t0 : 0.0);
    const float tlow = WINDOW ? -1e-30f : 0.0f;  // window mode: shift >= t_stop, so "ends before the window" is tf < 0
    const RayF wrf = make_rayf(ray.o + wshift * ray.d, ray.d);

    int stack[kStackSize];
    float stack_t[kStackSize];
    int sp = 0;
    const BvhNode* __restrict__ nodes = sc.nodes;
    RayF rf = wrf;
    double shift = wshift;
    float tbest = f32_up(hit.t - shift);
    tbest = tbest + fabsf(tbest) * kSlabSlack;
    bool in_blas = false;
    // mesh context while in_blas
    uint32_t cur_inst = 0, cur_obj = 0;
    d3 lo = ray.o, ld = ray.d;
    float fox = 0, foy = 0, foz = 0, fdx = 0, fdy = 0, fdz = 0;  // f32 local ray for the triangle pre-test
    const TriRec* __restrict__ tris = sc.tris;
    const TriF* __restrict__ trif = sc.trif;

    auto test_tris = [&](uint32_t first, uint32_t count) -> bool {
        bool improved = false;
        for (uint32_t i = first; i < first + count; i++) {
            if (!tri_maybe_hit(trif + i, fox, foy, foz, fdx, fdy, fdz)) continue;
            TriRec tr;
            load_tri(tris + i, &tr);
            double t, v, w;
            if (tri_test(tr, lo, ld, t_min, hit.t, cur_obj < hit.obj && hit.prim != kMonoPrim, &t, &v, &w) && (!WINDOW || t >= t_stop)) {
                hit.t = t; hit.inst = cur_inst; hit.obj = cur_obj; hit.prim = i; hit.v = v; hit.w = w;
                improved = true;
                if (SHADOW && t < t_stop) return true;
            }
        }
        if (improved) { tbest = f32_up(hit.t - shift); tbest = tbest + fabsf(tbest) * kSlabSlack; }
        return false;
    };

    int cur = sc.tlas_root_ref;
    float cur_t = 0.0f;  // f32 entry time (relative to wshift) of the node `cur`, from its parent's slab test
    for (;;) {
        // ---- inner nodes (same code for both levels) ----
        while (cur >= 0 && cur != kSentinel) {
            const float4* np = reinterpret_cast<const float4*>(nodes + cur);
            const float4 a = __ldg(np), b = __ldg(np + 1), c = __ldg(np + 2), d = __ldg(np + 3);
            float tl, tr;
            const bool hl = slab_f32(a.x, a.y, a.z, b.x, b.y, b.z, rf, tbest, &tl, tlow);
            const bool hr = slab_f32(c.x, c.y, c.z, d.x, d.y, d.z, rf, tbest, &tr, tlow);
            const int left = __float_as_int(a.w), right = __float_as_int(c.w);
            if (hl && hr) {
                const bool lf = tl <= tr;
                RPTX_ASSERT(sp < kStackSize);
                stack[sp] = lf ? right : left;
                stack_t[sp] = lf ? tr : tl;
                sp++;
                cur = lf ? left : right;
                cur_t = lf ? tl : tr;
            } else if (hl) {
                cur = left; cur_t = tl;
            } else if (hr) {
                cur = right; cur_t = tr;
            } else {
                cur = kSentinel + 0;  // placeholder, replaced by the pop below
                // pop
                for (;;) {
                    if (sp == 0) return;
                    sp--;
                    if (stack_t[sp] <= tbest) { cur = stack[sp]; cur_t = stack_t[sp]; break; }
                }
            }
        }
        // ---- leaf or sentinel ----
        if (cur == kSentinel) {
            in_blas = false;
            nodes = sc.nodes; rf = wrf; shift = wshift;
            tbest = f32_up(hit.t - shift);
            tbest = tbest + fabsf(tbest) * kSlabSlack;
        } else {
            const uint32_t code = (uint32_t)(~cur);
            const uint32_t first = code >> 3, count = (code & 7u) + 1u;
            if (in_blas) {
                if (test_tris(first, count)) return;
            } else {
                for (uint32_t idx = first; idx < first + count; idx++) {
                    const Instance* __restrict__ in = sc.instances + idx;
                    const int kind = in->kind;
                    cur_obj = in->object_id;
                    lo = ray.o; ld = ray.d;
                    if (in->has_transform) { lo = xform_point(in->inv, ray.o); ld = xform_dir(in->inv, ray.d); }
                    if (kind == kMesh) {
                        if ((ld.x == 0.0 || ld.y == 0.0 || ld.z == 0.0) && mesh_bounds_nan_reject(in, lo, ld)) continue;
                        cur_inst = idx;
                        tris = sc.tris + in->tri_base;
                        trif = sc.trif + in->tri_base;
                        if (in->root_ref < 0) {  // one-leaf mesh (a polygon): test its triangles here, no BLAS entry
                            fox = (float)lo.x; foy = (float)lo.y; foz = (float)lo.z;
                            fdx = (float)ld.x; fdy = (float)ld.y; fdz = (float)ld.z;
                            const uint32_t c2 = (uint32_t)(~in->root_ref);
                            if (test_tris(c2 >> 3, (c2 & 7u) + 1u)) return;
                            continue;
                        }
#ifdef RPTX_BLAS_CLIP
                        double l0, l1v;
                        clip_f64(in->lmin, in->lmax, lo, ld, &l0, &l1v);
                        if (!(l0 <= l1v) || l1v < 0.0 || l0 > hit.t) continue;
                        if (WINDOW && l1v < t_stop) continue;
#else
                        // The local f32 ray starts where the ray entered the instance's WORLD box: that time is the
                        // leaf's entry time of the slab test that led here (the parameter is the same in local space:
                        // the direction is not renormalised).  It lies at or before the local box entry and within
                        // the preimage of the world box, so |origin| stays within a few E of the mesh and its f32
                        // rounding inside the boxes' padding; an f64 clip against the local box (three divisions and
                        // as many instructions as a dozen node visits) only repeated what the BLAS root test does.
                        const double l0 = wshift + (double)cur_t;
#endif
                        // enter the mesh BVH (TLAS leaves hold one instance, so nothing is pending)
                        shift = WINDOW ? fmax(fmax(l0, t_stop), 0.0) : (l0 > 0.0 ? l0 : 0.0);
                        const d3 so = lo + shift * ld;
                        rf = make_rayf(so, ld);
                        fox = (float)so.x; foy = (float)so.y; foz = (float)so.z;
                        fdx = (float)ld.x; fdy = (float)ld.y; fdz = (float)ld.z;
                        tbest = f32_up(hit.t - shift);
                        tbest = tbest + fabsf(tbest) * kSlabSlack;
                        nodes = sc.nodes + in->node_base;
                        in_blas = true;
                        RPTX_ASSERT(sp + 1 < kStackSize);
                        stack[sp] = kSentinel;
                        stack_t[sp] = -CUDART_INF_F;
                        sp++;
                        stack[sp] = in->root_ref;
                        stack_t[sp] = -CUDART_INF_F;
                        sp++;
                        break;
                    }
                    double t;
                    uint32_t ccode = 0;
                    bool ok;
                    if (kind == kSphere) ok = sphere_test(lo, ld, t_min, &t);
                    else if (kind == kCube) ok = cube_test(lo, ld, t_min, &t, &ccode);
                    else ok = monomial_test(in->plane_v, lo, ld, t_min, hit.t, &t);
                    if (ok && (!WINDOW || t >= t_stop) && (kind == kMonomial ? accept_mono(hit, t, cur_obj) : accept(hit, t, cur_obj))) {
                        hit.t = t; hit.inst = idx; hit.obj = cur_obj; hit.prim = kind == kMonomial ? kMonoPrim : ccode;
                        if (SHADOW && t < t_stop) return;
                        tbest = f32_up(hit.t - shift);
                        tbest = tbest + fabsf(tbest) * kSlabSlack;
                    }
                }
            }
        }
        // pop next
        for (;;) {
            if (sp == 0) return;
            sp--;
            if (stack_t[sp] <= tbest) { cur = stack[sp]; cur_t = stack_t[sp]; break; }
        }
    }
}

// Scenes of a few analytic shapes and polygons (Cornell box, lampshade; SceneDev::small_scene):
// no hierarchy.  Two phases keep the warp together where the hierarchy cannot:
//   1. every lane walks the instance list in the same order and only runs the f32 pre-tests
//      (conservative triangle test, padded instance box), collecting its candidates in bit masks;
//   2. each lane pops ITS candidates and runs the exact f64 test, so in one pass of the loop the
//      lanes work on different primitives of the same kind.  With the hierarchy (or a plain list
//      walk) an exact triangle test ran with 3-5 live lanes: a warp's rays end on different walls.
// Same result as closest_hit: the pre-tests are the ones the hierarchy uses (a culled primitive
// is provably missed), accept() does not depend on the visiting order, and triangles of one mesh
// are still taken in ascending order.
template <int MODE>
RPTX_D void closest_hit_small(const SceneDev& sc, const RayD& ray, double t_min, double t_stop, Hit& hit) {
    constexpr bool SHADOW = MODE == kAnyBelow;
    constexpr bool WINDOW = MODE == kWindow;
    hit.inst = kNoHit; hit.obj = kNoHit; hit.prim = 0; hit.v = 0; hit.w = 0;
    // ---- phase 1: candidates ----
    unsigned long long tri_cand = 0;
    uint32_t box_cand = sc.small_unbounded_mask;
    {
        // f32 ray from the world-box entry, as in closest_hit; a ray with a zero direction component takes
        // every analytic instance (Cube::intersect's 0/0 quirk reports hits outside the box)
        double t0, t1;
        clip_f64(sc.wmin, sc.wmax, ray.o, ray.d, &t0, &t1);
        const bool world = sc.n_bounded != 0 && (t0 <= t1) && !(t1 < 0.0) && !(WINDOW && t1 < t_stop);
        const bool axis = ray.d.x == 0.0 || ray.d.y == 0.0 || ray.d.z == 0.0;
        const double wshift = WINDOW ? fmax(fmax(t0, t_stop), 0.0) : (t0 > 0.0 ? t0 : 0.0);
        const float tlow = WINDOW ? -1e-30f : 0.0f;
        const RayF rf = make_rayf(ray.o + wshift * ray.d, ray.d);
        float tbest = f32_up(hit.t - wshift);
        tbest = tbest + fabsf(tbest) * kSlabSlack;
        uint32_t slot = 0;
        const SmallTables& st = sc.small;  // kernel parameters: warp-uniform constant-bank loads
        for (uint32_t idx = 0; idx < sc.n_bounded; idx++) {
            if (st.kind[idx] != kMesh) {  // warp-uniform
                const float4 a = st.boxes[2 * idx], b = st.boxes[2 * idx + 1];
                float tn;
                if (axis || (world && slab_f32(a.x, a.y, a.z, b.x, b.y, b.z, rf, tbest, &tn, tlow))) box_cand |= 1u << idx;
                continue;
            }
            const uint32_t count = ((uint32_t)(~st.root_ref[idx]) & 7u) + 1u;
            d3 lo = ray.o, ld = ray.d;
            if (st.has_transform[idx]) { lo = xform_point(st.inv[idx], ray.o); ld = xform_dir(st.inv[idx], ray.d); }
            const bool rejected = (ld.x == 0.0 || ld.y == 0.0 || ld.z == 0.0) && mesh_bounds_nan_reject(sc.instances + idx, lo, ld);
            const float fox = (float)lo.x, foy = (float)lo.y, foz = (float)lo.z;
            const float fdx = (float)ld.x, fdy = (float)ld.y, fdz = (float)ld.z;
            for (uint32_t i = 0; i < count; i++, slot++)
                if (!rejected && tri_maybe_hit_v(st.trif[3 * slot], st.trif[3 * slot + 1], st.trif[3 * slot + 2], fox, foy, foz, fdx, fdy, fdz))
                    tri_cand |= 1ull << slot;
        }
    }
    // ---- phase 2: exact tests, each lane on its own candidates ----
    while (tri_cand) {
        const int k = __ffsll((long long)tri_cand) - 1;
        tri_cand &= tri_cand - 1;
        const uint32_t e = sc.small_slots[k];
        const uint32_t idx = e >> 3;
        const Instance* __restrict__ in = sc.instances + idx;
        const uint32_t obj = in->object_id;
        d3 lo = ray.o, ld = ray.d;
        if (in->has_transform) { lo = xform_point(in->inv, ray.o); ld = xform_dir(in->inv, ray.d); }
        const uint32_t i = ((uint32_t)(~in->root_ref) >> 3) + (e & 7u);
        TriRec tr;
        load_tri(sc.tris + in->tri_base + i, &tr);
        double t, v, w;
        if (tri_test(tr, lo, ld, t_min, hit.t, obj < hit.obj, &t, &v, &w) && (!WINDOW || t >= t_stop)) {
            hit.t = t; hit.inst = idx; hit.obj = obj; hit.prim = i; hit.v = v; hit.w = w;
            if (SHADOW && t < t_stop) return;
        }
    }
#pragma unroll
    for (int kind = 0; kind < 3; kind++) {  // kSphere, kPlane, kCube
        uint32_t m = box_cand & sc.small_kind_mask[kind];
        while (m) {
            const uint32_t idx = (uint32_t)__ffs((int)m) - 1u;
            m &= m - 1u;
            const Instance* __restrict__ in = sc.instances + idx;
            const uint32_t obj = in->object_id;
            d3 lo = ray.o, ld = ray.d;
            if (in->has_transform) { lo = xform_point(in->inv, ray.o); ld = xform_dir(in->inv, ray.d); }
            double t;
            uint32_t ccode = 0;
            bool ok;
            if (kind == kSphere) ok = sphere_test(lo, ld, t_min, &t);
            else if (kind == kCube) ok = cube_test(lo, ld, t_min, &t, &ccode);
            else ok = plane_test(*in, lo, ld, t_min, &t);
            if (ok && (!WINDOW || t >= t_stop) && accept(hit, t, obj)) {
                hit.t = t; hit.inst = idx; hit.obj = obj; hit.prim = ccode;
                if (SHADOW && t < t_stop) return;
            }
        }
    }
}

template <int MODE, bool SMALL>
RPTX_D void closest_hit_sel(const SceneDev& sc, const RayD& ray, double t_min, double t_stop, Hit& hit) {
    if (SMALL) closest_hit_small<MODE>(sc, ray, t_min, t_stop, hit);
    else closest_hit<MODE>(sc, ray, t_min, t_stop, hit);
}

// NEE visibility as the reference defines it (src/renderer.rs:347-348, :395-396): the CLOSEST
// hit of the shadow ray must lie within 1e-12 of the light sample.  Evaluated as (1) a probe of
// the window [dist - 2e-12, dist + 2e-12]: the closest hit REPORTED there must pass the test —
// a point query that fails at the root for lights floating in free space; (2) only then an
// any-hit search for a report below the window.  Same verdict as the full closest-hit search.
// probe_first is chosen per light at commit time (point lights: yes; emissive objects: no).
// PROBE = false: lights that are scene objects; the probe would nearly always succeed, so the
// segment is searched once up to the window's far end, stopping at the first report below it.
template <bool PROBE, bool SMALL = false>
RPTX_D bool nee_visible(const SceneDev& sc, const RayD& ray, double dist, unsigned long long* n_occlusion) {
    const double t_lo = dist - 2e-12;
    Hit h;
    h.t = dist + 2e-12;
    if (PROBE) {
        closest_hit_sel<kWindow, SMALL>(sc, ray, kEpsilon, t_lo, h);
        if (h.inst == kNoHit || !(fabs(h.t - dist) < kEpsilon)) return false;
        h.t = t_lo;
    }
    (*n_occlusion)++;
    closest_hit_sel<kAnyBelow, SMALL>(sc, ray, kEpsilon, t_lo, h);
    if (PROBE) return h.inst == kNoHit;
    return h.inst != kNoHit && fabs(h.t - dist) < kEpsilon;
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
    } else if (in->kind == kMonomial) {  // src/shape/monomial_surface.rs:94-104 (two-sided)
        const d3 pos = lo + hit.t * ld;
        const double h = in->plane_v, s = pos.x * pos.x + pos.z * pos.z;
        n = normalize(mk3(h * 4.0 * pos.x * s, -1.0, h * 4.0 * pos.z * s));
        if (dot(n, ld) > 0.0) n = -n;
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

// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// Restates src/shape.rs, src/shape/{sphere,plane,cube,mesh,monomial_surface}.rs and src/kdtree.rs of the
// reference.  Additions the reference lacks: HitRecord carries the triangle index
// (prim_id) and the group-child index (inst_id) of the winning primitive so that ray
// sets can be compared id-for-id.
#pragma once
#include <algorithm>
#include <cassert>
#include <memory>
#include <vector>

#include "rng.hpp"
#include "vec.hpp"

namespace orc {

// src/shape.rs:50-73
struct Ray {
    Vec3 origin, dir;
    Vec3 at(double time) const { return origin + time * dir; }
    // src/shape.rs:64-72 — direction is NOT renormalised
    Ray apply_transform(const Mat4& t) const { return {mul_h(t, origin, 1.0), mul_h(t, dir, 0.0)}; }
};

// src/shape.rs:76-98
struct HitRecord {
    double time = INF;
    Vec3 normal{0, 0, 0};
    uint32_t prim_id = 0;  // oracle addition
    uint32_t inst_id = 0;  // oracle addition
};

// src/kdtree.rs:31-99
struct BoundingBox {
    Vec3 p_min{INF, INF, INF}, p_max{-INF, -INF, -INF};
    BoundingBox merge(const BoundingBox& o) const { return {vmin(p_min, o.p_min), vmax(p_max, o.p_max)}; }
    // src/kdtree.rs:57-71 (f64::min/max semantics = fmin/fmax)
    void intersect(const Ray& ray, double* tmin, double* tmax) const {
        double x1 = (p_min.x - ray.origin.x) / ray.dir.x;
        double x2 = (p_max.x - ray.origin.x) / ray.dir.x;
        double xa = std::fmin(x1, x2), xb = std::fmax(x1, x2);
        double y1 = (p_min.y - ray.origin.y) / ray.dir.y;
        double y2 = (p_max.y - ray.origin.y) / ray.dir.y;
        double ya = std::fmin(y1, y2), yb = std::fmax(y1, y2);
        double z1 = (p_min.z - ray.origin.z) / ray.dir.z;
        double z2 = (p_max.z - ray.origin.z) / ray.dir.z;
        double za = std::fmin(z1, z2), zb = std::fmax(z1, z2);
        *tmin = std::fmax(std::fmax(xa, ya), za);
        *tmax = std::fmin(std::fmin(xb, yb), zb);
    }
    // src/kdtree.rs:74-98
    void split(int axis, double value, BoundingBox* l, BoundingBox* r) const {
        *l = *this; l->p_max[axis] = value;
        *r = *this; r->p_min[axis] = value;
    }
};

// src/shape.rs:19-26 (+ Bounded, src/kdtree.rs:12-15)
struct Shape {
    virtual ~Shape() = default;
    virtual bool intersect(const Ray& ray, double t_min, HitRecord& rec) const = 0;
    virtual void sample(const Vec3& target, Rng& rng, Vec3* v, Vec3* n, double* p) const = 0;
    virtual bool bounded() const { return true; }
    virtual BoundingBox bounding_box() const = 0;
};

// src/shape/sphere.rs:10-75
struct Sphere : Shape {
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        const double a = length2(ray.dir);
        const double b = dot(ray.dir, ray.origin);
        const double c = length2(ray.origin) - 1.0;
        double d = b * b - a * c;
        if (std::signbit(d)) return false;  // is_sign_negative (true for -0.0 and -NaN too)
        d = std::sqrt(d);
        double t;
        const double t_minus = (-b - d) / a;
        if (t_minus < t_min) {
            const double t_plus = (-b + d) / a;
            if (t_plus < t_min) return false;
            t = t_plus;
        } else {
            t = t_minus;
        }
        if (t < rec.time) {
            rec.time = t;
            rec.normal = normalize(ray.at(t));
            rec.prim_id = 0;
            return true;
        }
        return false;
    }
    void sample(const Vec3& target, Rng& rng, Vec3* v, Vec3* n, double* p) const override {
        double x, y;
        rng.unit_disc(&x, &y);
        const double z = std::sqrt(1.0 - x * x - y * y);
        const Vec3 nn = normalize(target);
        Vec3 n1;
        if (std::isnormal(nn.x)) n1 = normalize(Vec3(nn.y, -nn.x, 0.0));
        else n1 = normalize(Vec3(0.0, -nn.z, nn.y));
        const Vec3 n2 = cross(n1, nn);
        const Vec3 pt = x * n1 + y * n2 + z * nn;
        *v = pt; *n = pt; *p = z * (1.0 / PI);
    }
    BoundingBox bounding_box() const override { return {Vec3(-1, -1, -1), Vec3(1, 1, 1)}; }
};

// src/shape/plane.rs:7-37
struct Plane : Shape {
    Vec3 normal;
    double value;
    Plane(const Vec3& n, double v) : normal(n), value(v) {}
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        const double cosine = dot(normal, ray.dir);
        if (std::fabs(cosine) < 1e-8) return false;
        const double time = (value - dot(normal, ray.origin)) / cosine;
        if (time >= t_min && time < rec.time) {
            rec.time = time;
            rec.normal = (-normalize(normal)) * signum(cosine);
            rec.prim_id = 0;
            return true;
        }
        return false;
    }
    void sample(const Vec3&, Rng&, Vec3*, Vec3*, double*) const override { assert(!"unimplemented"); }
    bool bounded() const override { return false; }
    BoundingBox bounding_box() const override { return {}; }
};

// src/shape/monomial_surface.rs:8-190 — y = height * (x^2 + z^2)^2 over the unit disc.  powi(2)/powi(3) are
// restated as the multiplication chains LLVM emits for constant integer powers.
struct MonomialSurface : Shape {
    double height, exp;
    MonomialSurface(double h, double e) : height(h), exp(e) {}
    BoundingBox bounding_box() const override { return {Vec3(-1.0, 0.0, -1.0), Vec3(1.0, height, 1.0)}; }
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        double b_min, b_max;
        bounding_box().intersect(ray, &b_min, &b_max);
        if (std::fmax(b_min, t_min) > std::fmin(b_max, rec.time)) return false;  // :23-26
        auto dist = [&](double t) {
            const double x = ray.origin.x + t * ray.dir.x;
            const double y = ray.origin.y + t * ray.dir.y;
            const double z = ray.origin.z + t * ray.dir.z;
            const double s = x * x + z * z;
            return y - height * (s * s);
        };
        const double coef0 = ray.origin.x * ray.origin.x + ray.origin.z * ray.origin.z;
        const double coef1 = 2. * (ray.origin.x * ray.dir.x + ray.origin.z * ray.dir.z);
        const double coef2 = ray.dir.x * ray.dir.x + ray.dir.z * ray.dir.z;
        auto deriv = [&](double t) {
            const double dy = 2. * coef0 * coef1 + 2. * t * (coef1 * coef1 + 2. * coef0 * coef2) +
                              3. * (t * t) * 2. * coef1 * coef2 + 4. * (t * t * t) * coef2 * coef2;
            return ray.dir.y - height * dy;
        };
        auto deriv2 = [&](double t) {
            const double dy = 2. * (coef1 * coef1 + 2. * coef0 * coef2) + 3. * 2. * t * 2. * coef1 * coef2 +
                              4. * 3. * (t * t) * coef2 * coef2;
            return -height * dy;
        };
        double t_max;
        const bool maximize = dist(t_min) < 0.0;
        if (maximize) {  // Newton steps toward the maximum of dist, :51-67
            double cur_x = (b_min + b_max) / 2.;
            for (int i = 0; i < 10; i++) {
                const double f = dist(cur_x);
                if (f > 0.) break;
                const double der = deriv(cur_x);
                const double der2 = deriv2(cur_x);
                cur_x -= der / der2;
            }
            t_max = cur_x;
            if (t_max < t_min) return false;
        } else {
            t_max = 10000.;
        }
        if ((dist(t_min) < 0.0) == (dist(t_max) < 0.0)) return false;
        double l = t_min, r = t_max;
        for (int i = 0; i < 60; i++) {  // bisection, :75-82
            const double m = (l + r) / 2.0;
            if ((dist(m) >= 0.0) == maximize) r = m;
            else l = m;
        }
        if (r > rec.time) return false;
        const Vec3 pos = ray.at(r);
        if (pos.x * pos.x + pos.z * pos.z > 1.0) return false;
        rec.time = r;
        rec.normal = normalize(Vec3(height * 4.0 * pos.x * (pos.x * pos.x + pos.z * pos.z), -1.0,
                                    height * 4.0 * pos.z * (pos.x * pos.x + pos.z * pos.z)));
        if (dot(rec.normal, ray.dir) > 0.0) rec.normal = -rec.normal;
        rec.prim_id = 0;
        return true;
    }
    // :109-124; rand_distr::UnitCircle = von Neumann's rejection method, gen::<bool>() restated as U < 0.5
    void sample(const Vec3&, Rng& rng, Vec3* v, Vec3* n, double* p) const override {
        double x, z;
        rng.unit_circle(&x, &z);
        const Vec3 pos(x, height * std::pow(x * x + z * z, exp / 2.), z);
        Vec3 normal = normalize(Vec3(height * 4. * pos.x * (pos.x * pos.x + pos.z * pos.z), -1.,
                                     height * 4. * pos.z * (pos.x * pos.x + pos.z * pos.z)));
        const double AREA = 6.3406654362;
        if (rng.gen() < 0.5) normal = -normal;
        *v = pos; *n = normal; *p = 1. / (2. * AREA);
    }
};

// src/shape/cube.rs:9-90
struct Cube : Shape {
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        double lo[3], hi[3];
        Vec3 lon[3], hin[3];
        for (int dim = 0; dim < 3; dim++) {
            double x1 = (-0.5 - ray.origin[dim]) / ray.dir[dim];
            double x2 = (0.5 - ray.origin[dim]) / ray.dir[dim];
            Vec3 x1n(0, 0, 0), x2n(0, 0, 0);
            x1n[dim] = -1.0;
            x2n[dim] = 1.0;
            if (x1 > x2) { std::swap(x1, x2); std::swap(x1n, x2n); }
            lo[dim] = x1; hi[dim] = x2; lon[dim] = x1n; hin[dim] = x2n;
        }
        double start, end;
        Vec3 sn, en;
        if (lo[0] > lo[1] && lo[0] > lo[2]) { start = lo[0]; sn = lon[0]; }
        else if (lo[1] > lo[2]) { start = lo[1]; sn = lon[1]; }
        else { start = lo[2]; sn = lon[2]; }
        if (hi[0] < hi[1] && hi[0] < hi[2]) { end = hi[0]; en = hin[0]; }
        else if (hi[1] < hi[2]) { end = hi[1]; en = hin[1]; }
        else { end = hi[2]; en = hin[2]; }
        if (start > end || end < t_min) return false;
        double time; Vec3 nrm;
        if (start < t_min) { time = end; nrm = en; } else { time = start; nrm = sn; }
        if (time < rec.time) {
            rec.time = time; rec.normal = nrm; rec.prim_id = 0;
            return true;
        }
        return false;
    }
    void sample(const Vec3&, Rng& rng, Vec3* v, Vec3* n, double* p) const override {
        const double a = rng.gen() - 0.5;
        const double b = rng.gen() - 0.5;
        switch (rng.gen_index(6)) {
            case 0: *v = Vec3(a, b, 0.5); *n = Vec3(0, 0, 1); break;
            case 1: *v = Vec3(a, b, -0.5); *n = Vec3(0, 0, -1); break;
            case 2: *v = Vec3(a, 0.5, b); *n = Vec3(0, 1, 0); break;
            case 3: *v = Vec3(a, -0.5, b); *n = Vec3(0, -1, 0); break;
            case 4: *v = Vec3(0.5, a, b); *n = Vec3(1, 0, 0); break;
            default: *v = Vec3(-0.5, a, b); *n = Vec3(-1, 0, 0); break;
        }
        *p = 1.0 / 6.0;
    }
    BoundingBox bounding_box() const override { return {Vec3(-0.5, -0.5, -0.5), Vec3(0.5, 0.5, 0.5)}; }
};

// src/shape/mesh.rs:9-100
struct Triangle : Shape {
    Vec3 v1, v2, v3, n1, n2, n3;
    static Triangle from_vertices(const Vec3& a, const Vec3& b, const Vec3& c) {
        Triangle t;
        const Vec3 n = normalize(cross(b - a, c - a));
        t.v1 = a; t.v2 = b; t.v3 = c; t.n1 = n; t.n2 = n; t.n3 = n;
        return t;
    }
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        const Vec3 d0 = v2 - v1, d1 = v3 - v1;
        const Vec3 plane_normal = normalize(cross(d0, d1));
        const double cosine = dot(plane_normal, ray.dir);
        if (std::fabs(cosine) < 1e-8) return false;
        const double time = dot(plane_normal, v1 - ray.origin) / cosine;
        if (time < t_min || time >= rec.time) return false;
        const Vec3 d2 = ray.at(time) - v1;
        const double d00 = dot(d0, d0), d01 = dot(d0, d1), d11 = dot(d1, d1);
        const double d20 = dot(d2, d0), d21 = dot(d2, d1);
        const double denom = d00 * d11 - d01 * d01;
        const double v = (d11 * d20 - d01 * d21) / denom;
        const double w = (d00 * d21 - d01 * d20) / denom;
        const double u = 1.0 - v - w;
        if (u >= 0.0 && v >= 0.0 && w >= 0.0) {
            rec.time = time;
            rec.normal = normalize(u * n1 + v * n2 + w * n3);
            return true;
        }
        return false;
    }
    void sample(const Vec3&, Rng& rng, Vec3* v, Vec3* n, double* p) const override {
        double u = rng.gen(), vv = rng.gen();
        while (u + vv > 1.0) { u = rng.gen(); vv = rng.gen(); }
        const double w = 1.0 - u - vv;
        const double area = 0.5 * length(cross(v2 - v1, v3 - v1));
        *v = u * v1 + vv * v2 + w * v3;
        *n = normalize(u * n1 + vv * n2 + w * n3);
        *p = 1.0 / area;
    }
    BoundingBox bounding_box() const override {
        return {vmin(vmin(v1, v2), v3), vmax(vmax(v1, v2), v3)};
    }
};

// src/shape.rs:102-178
struct Transformed : Shape {
    std::shared_ptr<const Shape> shape;
    Mat4 transform, inverse_transform;
    Mat3 linear, normal_transform;
    double scale;
    Transformed(std::shared_ptr<const Shape> s, const Mat4& t) : shape(std::move(s)), transform(t) {
        inverse_transform = inverse(t);
        linear = mat4_to_mat3(t);
        scale = determinant(linear);
        normal_transform = inverse_transpose(linear);
    }
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        const Ray local = ray.apply_transform(inverse_transform);
        if (shape->intersect(local, t_min, rec)) {
            rec.normal = normalize(mul(normal_transform, rec.normal));
            return true;
        }
        return false;
    }
    void sample(const Vec3& target, Rng& rng, Vec3* v, Vec3* n, double* p) const override {
        const Vec3 tl = mul_h(inverse_transform, target, 1.0);
        Vec3 lv, ln; double lp;
        shape->sample(tl, rng, &lv, &ln, &lp);
        const Vec3 new_normal = normalize(mul(normal_transform, ln));
        const double height = dot(mul(linear, ln), new_normal);
        const double base = scale / height;
        *v = mul_h(transform, lv, 1.0);
        *n = new_normal;
        *p = lp / base;
    }
    BoundingBox bounding_box() const override {
        const BoundingBox b = shape->bounding_box();
        BoundingBox out;
        const double xs[2] = {b.p_min.x, b.p_max.x}, ys[2] = {b.p_min.y, b.p_max.y}, zs[2] = {b.p_min.z, b.p_max.z};
        for (int i = 0; i < 2; i++)
            for (int j = 0; j < 2; j++)
                for (int k = 0; k < 2; k++) {
                    const Vec3 c = mul_h(transform, Vec3(xs[i], ys[j], zs[k]), 1.0);
                    out.p_min = vmin(out.p_min, c);
                    out.p_max = vmax(out.p_max, c);
                }
        return out;
    }
};

// src/kdtree.rs:228-234
struct KdNode {
    int axis = -1;  // -1 leaf, 0/1/2 split
    double value = 0;
    std::unique_ptr<KdNode> left, right;
    std::vector<size_t> indices;
};

// src/kdtree.rs:350-358
inline double kd_median(const std::vector<double>& s) {
    assert(!s.empty());
    if (s.size() % 2 == 1) return s[s.size() / 2];
    const size_t mid = s.size() / 2;
    return (s[mid] + s[mid - 1]) / 2.0;
}

// src/kdtree.rs:238-348
inline std::unique_ptr<KdNode> kd_construct(const std::vector<BoundingBox>& all_boxes, std::vector<size_t> indices) {
    auto node = std::make_unique<KdNode>();
    if (indices.size() < 16) { node->indices = std::move(indices); return node; }
    std::vector<double> xs, ys, zs;
    std::vector<BoundingBox> bboxs;
    for (size_t index : indices) {
        const BoundingBox& bb = all_boxes[index];
        xs.push_back(bb.p_min.x); xs.push_back(bb.p_max.x);
        ys.push_back(bb.p_min.y); ys.push_back(bb.p_max.y);
        zs.push_back(bb.p_min.z); zs.push_back(bb.p_max.z);
        bboxs.push_back(bb);
    }
    std::sort(xs.begin(), xs.end()); std::sort(ys.begin(), ys.end()); std::sort(zs.begin(), zs.end());
    const double med[3] = {kd_median(xs), kd_median(ys), kd_median(zs)};
    auto score = [&](int dim, double value) {
        size_t l = 0, r = 0;
        for (const auto& bb : bboxs) {
            if (bb.p_min[dim] <= value) l++;
            if (bb.p_max[dim] >= value) r++;
        }
        return std::max(l, r);
    };
    const size_t sx = score(0, med[0]), sy = score(1, med[1]), sz = score(2, med[2]);
    const size_t threshold = (size_t)((double)indices.size() * 0.85);
    if (std::min(std::min(sx, sy), sz) >= threshold) { node->indices = std::move(indices); return node; }
    int split_dir = -1;
    BoundingBox bounds;
    for (const auto& bb : bboxs) bounds = bounds.merge(bb);
    const Vec3 extent = bounds.p_max - bounds.p_min;
    if (extent.x > extent.y && extent.x > extent.z) { if (sx < threshold) split_dir = 0; }
    else if (extent.y > extent.z) { if (sy < threshold) split_dir = 1; }
    else if (sz < threshold) split_dir = 2;
    if (split_dir == -1) {
        if (sx < sy && sx < sz) split_dir = 0;
        else if (sy < sz) split_dir = 1;
        else split_dir = 2;
    }
    std::vector<size_t> left, right;
    for (size_t i = 0; i < indices.size(); i++) {
        if (bboxs[i].p_min[split_dir] <= med[split_dir]) left.push_back(indices[i]);
        if (bboxs[i].p_max[split_dir] >= med[split_dir]) right.push_back(indices[i]);
    }
    node->axis = split_dir;
    node->value = med[split_dir];
    node->left = kd_construct(all_boxes, std::move(left));
    node->right = kd_construct(all_boxes, std::move(right));
    return node;
}

// src/kdtree.rs:103-226.  `group` selects which id a leaf hit records: triangle index
// (Mesh = KdTree<Triangle>) or child index (KdTree<Box<dyn Bounded>>).
struct KdTree : Shape {
    std::unique_ptr<KdNode> root;
    std::vector<std::shared_ptr<const Shape>> objects;
    BoundingBox bounds;
    bool group;
    size_t n_nodes = 0, n_leaves = 0, n_leaf_items = 0;
    KdTree(std::vector<std::shared_ptr<const Shape>> objs, bool is_group) : objects(std::move(objs)), group(is_group) {
        std::vector<BoundingBox> boxes;
        std::vector<size_t> indices;
        for (size_t i = 0; i < objects.size(); i++) {
            boxes.push_back(objects[i]->bounding_box());
            bounds = bounds.merge(boxes.back());
            indices.push_back(i);
        }
        root = kd_construct(boxes, std::move(indices));
        count(root.get());
    }
    void count(const KdNode* n) {
        n_nodes++;
        if (n->axis < 0) { n_leaves++; n_leaf_items += n->indices.size(); return; }
        count(n->left.get()); count(n->right.get());
    }
    bool intersect(const Ray& ray, double t_min, HitRecord& rec) const override {
        double b_min, b_max;
        bounds.intersect(ray, &b_min, &b_max);
        if (std::fmax(b_min, t_min) > std::fmin(b_max, rec.time)) return false;
        return intersect_subtree(root.get(), bounds, ray, t_min, rec);
    }
    bool intersect_subtree(const KdNode* node, const BoundingBox& bbox, const Ray& ray, double t_min, HitRecord& rec) const {
        double b_min, b_max;
        bbox.intersect(ray, &b_min, &b_max);
        if (node->axis < 0) {
            bool result = false;
            for (size_t index : node->indices) {
                if (objects[index]->intersect(ray, t_min, rec)) {
                    result = true;
                    if (group) rec.inst_id = (uint32_t)index; else rec.prim_id = (uint32_t)index;
                }
            }
            return result;
        }
        const int ax = node->axis;
        const double value = node->value;
        const double t_split = (value - ray.origin[ax]) / ray.dir[ax];
        const bool left_first = (ray.origin[ax] < value) || (ray.origin[ax] == value && ray.dir[ax] <= 0.0);
        BoundingBox bl, br;
        bbox.split(ax, value, &bl, &br);
        const KdNode* first = left_first ? node->left.get() : node->right.get();
        const KdNode* second = left_first ? node->right.get() : node->left.get();
        const BoundingBox& b0 = left_first ? bl : br;
        const BoundingBox& b1 = left_first ? br : bl;
        if (t_split > std::fmin(b_max, rec.time) || t_split <= 0.0) {
            return intersect_subtree(first, b0, ray, t_min, rec);
        } else if (t_split < std::fmax(b_min, t_min)) {
            return intersect_subtree(second, b1, ray, t_min, rec);
        } else {
            const bool h1 = intersect_subtree(first, b0, ray, t_min, rec);
            if (h1 && rec.time < t_split) return true;
            const bool h2 = intersect_subtree(second, b1, ray, t_split, rec);
            return h1 || h2;
        }
    }
    // src/kdtree.rs:141-146
    void sample(const Vec3& target, Rng& rng, Vec3* v, Vec3* n, double* p) const override {
        const size_t num = objects.size();
        const size_t index = rng.gen_index(num);
        objects[index]->sample(target, rng, v, n, p);
        *p = *p / (double)num;
    }
    BoundingBox bounding_box() const override { return bounds; }
};

}  // namespace orc

// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// Restates src/photon.rs of the reference: photon emission and tracing (:724-946), the three
// photon-map builders (:193-305), the radiance estimators (:316-627) and the render driver
// (:655-720, :950-985).
//
// Third-party algorithms whose source is NOT under /root/reference (Cargo.toml:20-21), restated
// from their published behaviour:
//   kd-tree 0.4.1  KdTree::build_by / nearests(q,k): exact k nearest by squared Euclidean
//                  distance (fewer if the map is smaller).  Any exact kNN is equivalent up
//                  to distance ties; PointKd below is a median-split kd-tree.
//   bvh 0.6.0      BVH::build / traverse: returns every shape whose f32 AABB passes the
//                  Williams slab test against the f32 ray with normalised direction and
//                  ray_max > 0 (Ray::intersects_aabb); leaves hold one shape; node boxes are
//                  exact f32 unions, so the result is exactly that per-shape set.
// PARITY UNPINNED (no reference tests or golden vectors exist for this path).
#pragma once
#include <chrono>
#include <numeric>
#include <queue>

#include "render.hpp"

namespace orc {

enum PhotonRenderKind { PhotonMapKind = 0, PhotonPointBeam = 1, PhotonBeamBeam = 2 };
constexpr uint64_t DOMAIN_THIN = 4;

// src/photon.rs:25-34
struct Photon {
    Vec3 position, direction, power, starting_position;
    uint32_t depth = 0;  // oracle addition: num_bounces at deposit (keys the thinning stream)
    uint64_t index = 0;  // oracle addition: index of the emitted photon this deposit belongs to
    uint64_t key() const { return (index << 16) | (uint64_t)(depth < 65535u ? depth : 65535u); }
};

struct PhotonList {
    std::vector<Photon> surface, volume;
};

// ---- kd-tree 0.4.1 restatement: exact kNN ----
struct PointKd {
    std::vector<Photon> items;  // reordered in place, implicit balanced tree
    void build(std::vector<Photon> pts) {
        items = std::move(pts);
        build_rec(0, items.size(), 0);
    }
    void build_rec(size_t lo, size_t hi, int axis) {
        if (hi - lo < 2) return;
        const size_t mid = lo + (hi - lo) / 2;
        std::nth_element(items.begin() + lo, items.begin() + mid, items.begin() + hi,
                         [axis](const Photon& a, const Photon& b) { return a.position[axis] < b.position[axis]; });
        build_rec(lo, mid, (axis + 1) % 3);
        build_rec(mid + 1, hi, (axis + 1) % 3);
    }
    struct Found {
        double d2;
        const Photon* item;
        bool operator<(const Found& o) const { return d2 < o.d2; }
    };
    static double dist2(const Vec3& a, const Vec3& b) {
        const double dx = a.x - b.x, dy = a.y - b.y, dz = a.z - b.z;
        return dx * dx + dy * dy + dz * dz;
    }
    void nearests(const Vec3& q, size_t k, std::vector<Found>* out) const {
        out->clear();
        if (k == 0 || items.empty()) return;
        std::priority_queue<Found> heap;
        search(0, items.size(), 0, q, k, heap);
        out->resize(heap.size());
        for (size_t i = heap.size(); i-- > 0;) { (*out)[i] = heap.top(); heap.pop(); }  // ascending
    }
    void search(size_t lo, size_t hi, int axis, const Vec3& q, size_t k, std::priority_queue<Found>& heap) const {
        if (lo >= hi) return;
        const size_t mid = lo + (hi - lo) / 2;
        const Photon& p = items[mid];
        const double d2 = dist2(p.position, q);
        if (heap.size() < k) heap.push({d2, &p});
        else if (d2 < heap.top().d2) { heap.pop(); heap.push({d2, &p}); }
        const double delta = q[axis] - p.position[axis];
        const int na = (axis + 1) % 3;
        if (delta < 0) {
            search(lo, mid, na, q, k, heap);
            if (heap.size() < k || delta * delta <= heap.top().d2) search(mid + 1, hi, na, q, k, heap);
        } else {
            search(mid + 1, hi, na, q, k, heap);
            if (heap.size() < k || delta * delta <= heap.top().d2) search(lo, mid, na, q, k, heap);
        }
    }
};

// ---- bvh 0.6.0 restatement: f32 AABB candidate filter ----
struct Aabb32 {
    float mn[3], mx[3];
};
struct Ray32 {
    float o[3], inv[3];
    int sign[3];
    Ray32(const Ray& r) {
        float d[3] = {(float)r.dir.x, (float)r.dir.y, (float)r.dir.z};
        const float len = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
        o[0] = (float)r.origin.x; o[1] = (float)r.origin.y; o[2] = (float)r.origin.z;
        for (int i = 0; i < 3; i++) {
            d[i] = d[i] / len;
            inv[i] = 1.0f / d[i];
            sign[i] = d[i] < 0.0f ? 1 : 0;
        }
    }
    bool intersects(const Aabb32& b) const {
        const float* bounds[2] = {b.mn, b.mx};
        float ray_min = (bounds[sign[0]][0] - o[0]) * inv[0];
        float ray_max = (bounds[1 - sign[0]][0] - o[0]) * inv[0];
        const float y_min = (bounds[sign[1]][1] - o[1]) * inv[1];
        const float y_max = (bounds[1 - sign[1]][1] - o[1]) * inv[1];
        if (ray_min > y_max || y_min > ray_max) return false;
        if (y_min > ray_min) ray_min = y_min;
        if (y_max < ray_max) ray_max = y_max;
        const float z_min = (bounds[sign[2]][2] - o[2]) * inv[2];
        const float z_max = (bounds[1 - sign[2]][2] - o[2]) * inv[2];
        if (ray_min > z_max || z_min > ray_max) return false;
        if (z_max < ray_max) ray_max = z_max;
        return ray_max > 0.0f;
    }
};
struct BoxBvh {
    struct Node { Aabb32 box; int left, right, item; };
    std::vector<Node> nodes;
    std::vector<Aabb32> boxes;
    void build(const std::vector<Aabb32>& b) {
        boxes = b;
        nodes.clear();
        if (b.empty()) return;
        std::vector<int> idx(b.size());
        std::iota(idx.begin(), idx.end(), 0);
        nodes.reserve(2 * b.size());
        build_rec(idx, 0, idx.size());
    }
    int build_rec(std::vector<int>& idx, size_t lo, size_t hi) {
        Node n; n.left = n.right = n.item = -1;
        for (int k = 0; k < 3; k++) { n.box.mn[k] = INFINITY; n.box.mx[k] = -INFINITY; }
        for (size_t i = lo; i < hi; i++)
            for (int k = 0; k < 3; k++) {
                n.box.mn[k] = std::fmin(n.box.mn[k], boxes[idx[i]].mn[k]);
                n.box.mx[k] = std::fmax(n.box.mx[k], boxes[idx[i]].mx[k]);
            }
        const int me = (int)nodes.size();
        nodes.push_back(n);
        if (hi - lo == 1) { nodes[me].item = idx[lo]; return me; }
        int ax = 0;
        float ext = -1;
        for (int k = 0; k < 3; k++) if (n.box.mx[k] - n.box.mn[k] > ext) { ext = n.box.mx[k] - n.box.mn[k]; ax = k; }
        const size_t mid = lo + (hi - lo) / 2;
        std::nth_element(idx.begin() + lo, idx.begin() + mid, idx.begin() + hi, [&](int a, int c) {
            return boxes[a].mn[ax] + boxes[a].mx[ax] < boxes[c].mn[ax] + boxes[c].mx[ax];
        });
        const int l = build_rec(idx, lo, mid);
        const int r = build_rec(idx, mid, hi);
        nodes[me].left = l; nodes[me].right = r;
        return me;
    }
    void traverse(const Ray32& ray, std::vector<int>* out) const {
        out->clear();
        if (nodes.empty()) return;
        if (nodes.size() == 1) { out->push_back(nodes[0].item); return; }  // single-leaf root is returned untested
        std::vector<int> stack{0};
        while (!stack.empty()) {
            const Node& n = nodes[stack.back()];
            stack.pop_back();
            if (n.item >= 0) { out->push_back(n.item); continue; }
            if (ray.intersects(nodes[n.left].box)) stack.push_back(n.left);
            if (ray.intersects(nodes[n.right].box)) stack.push_back(n.right);
        }
    }
};

// src/photon.rs:46-59
struct PhotonBeam {
    Vec3 start_position, end_position;
    double radius;
    Ray ray;
    Vec3 power;
    // src/photon.rs:76-105
    Aabb32 aabb() const {
        const double ax = start_position.x, ay = start_position.y, az = start_position.z;
        const double bx = end_position.x, by = end_position.y, bz = end_position.z;
        const double cx_sq = (ax - bx) * (ax - bx), cy_sq = (ay - by) * (ay - by), cz_sq = (az - bz) * (az - bz);
        const double sq_sum = cx_sq + cy_sq + cz_sq;
        const double kx = std::sqrt((cy_sq + cz_sq) / sq_sum);
        const double ky = std::sqrt((cx_sq + cz_sq) / sq_sum);
        const double kz = std::sqrt((cx_sq + cy_sq) / sq_sum);
        const Vec3 adj(kx * radius, ky * radius, kz * radius);
        const Vec3 mn = Vec3(std::fmin(ax, bx), std::fmin(ay, by), std::fmin(az, bz)) - adj;
        const Vec3 mx = Vec3(std::fmax(ax, bx), std::fmax(ay, by), std::fmax(az, bz)) + adj;
        return {{(float)mn.x, (float)mn.y, (float)mn.z}, {(float)mx.x, (float)mx.y, (float)mx.z}};
    }
};

// src/photon.rs:62-73
struct PhotonSphere {
    Vec3 position;
    double radius;
    Vec3 direction, power;
    // src/photon.rs:119-128
    Aabb32 aabb() const {
        const Vec3 h(radius, radius, radius);
        const Vec3 mn = position - h, mx = position + h;
        return {{(float)mn.x, (float)mn.y, (float)mn.z}, {(float)mx.x, (float)mx.y, (float)mx.z}};
    }
};

struct PhotonStats {
    uint64_t n_surface = 0, n_volume = 0;
    double emit_s = 0, build_s = 0, gather_s = 0;
};

// src/photon.rs:176-180
struct PhotonMap {
    PhotonRenderKind kind;
    PointKd surface_map, volume_map;
    BoxBvh bvh;
    std::vector<PhotonSphere> spheres;
    std::vector<PhotonBeam> beams;

    // src/photon.rs:193-305
    void build(PhotonRenderKind k, PhotonList list) {
        kind = k;
        surface_map.build(std::move(list.surface));
        if (k == PhotonMapKind) {
            volume_map.build(std::move(list.volume));
        } else if (k == PhotonPointBeam) {
            volume_map.build(list.volume);
            std::vector<PointKd::Found> found;
            std::vector<Aabb32> boxes;
            for (const Photon& p : list.volume) {
                volume_map.nearests(p.position, 10, &found);
                double m = -1.0;
                for (const auto& f : found) m = std::fmax(m, f.d2);
                spheres.push_back({p.position, std::sqrt(m), p.direction, p.power});
                boxes.push_back(spheres.back().aabb());
            }
            bvh.build(boxes);
        } else {
            std::vector<Aabb32> boxes;
            for (const Photon& p : list.volume) {
                PhotonBeam b;
                b.start_position = p.starting_position;
                b.end_position = p.position;
                b.radius = 3.0;
                b.ray = {p.starting_position, normalize(p.position - p.starting_position)};
                b.power = p.power;
                beams.push_back(b);
                boxes.push_back(b.aabb());
            }
            bvh.build(boxes);
        }
    }

    static double k2(double sq) {
        const double tmp = 1.0 - sq;
        return (3.0 / PI) * tmp * tmp;
    }

    // src/photon.rs:327-375
    Color surface_estimate(const Renderer& r, const Ray& ray, const Vec3& wo, const HitRecord& h, const Material& material) const {
        const Vec3 world_pos = ray.at(h.time);
        std::vector<PointKd::Found> near;
        surface_map.nearests(world_pos, r.gather_size, &near);
        double max_d2 = 0.0;
        for (const auto& f : near) max_d2 = std::fmax(max_d2, f.d2);
        Color color = material.emittance() * material.color();
        for (const auto& f : near) {
            const Photon& photon = *f.item;
            const Vec3 disp = world_pos - photon.position;
            const Ray vr{photon.position, normalize(disp)};
            HitRecord vh; const Object* o;
            if (r.get_closest_hit(vr, &vh, &o)) {
                if (length(disp) > vh.time) continue;
            }
            color += cmul(material.bsdf(h.normal, wo, photon.direction), photon.power) *
                     clampd(dot(photon.direction, h.normal), 0.0, 1.0);
        }
        color = color * (1.0 / (PI * max_d2));
        return color;
    }

    // src/photon.rs:379-594
    Color volume_estimate(const Renderer& r, const Ray& ray, const Vec3& wo, const Medium& medium, const HitRecord* hit,
                          const Object* object, Rng& rng) const {
        if (kind == PhotonMapKind) {
            double d, d_pdf, d_cdf;
            medium.sample_d(ray, rng, &d, &d_pdf, &d_cdf);
            if (!hit || d < hit->time) {
                const Vec3 collision = ray.at(d);
                const double abs_ = medium.absorption(collision);
                const Color medium_color = medium.color(collision);
                const double scat = medium.scattering(collision);
                const double extinction = abs_ + scat;
                Color color(0, 0, 0);
                std::vector<PointKd::Found> near;
                volume_map.nearests(collision, r.gather_size_volume, &near);
                double max_d2 = 0.0;
                for (const auto& f : near) max_d2 = std::fmax(max_d2, f.d2);
                for (const auto& f : near) color += cmul(f.item->power, medium_color) * medium.phase(wo, f.item->direction);
                color /= (4.0 / 3.0) * PI * std::pow(max_d2, 1.5);
                color /= extinction;
                color *= medium.transmittence(ray, d);
                color /= d_pdf;
                return color;
            }
            return surface_estimate(r, ray, wo, *hit, object->material) * medium.transmittence(ray, hit->time) / (1.0 - d_cdf);
        }
        const Ray32 r32(ray);
        std::vector<int> cand;
        bvh.traverse(r32, &cand);
        const Vec3 dummy(0, 0, 0);
        const Color medium_color = medium.color(dummy);
        const double extinction = medium.absorption(dummy) + medium.scattering(dummy);
        Color volume_color(0, 0, 0);
        if (kind == PhotonPointBeam) {
            for (int ci : cand) {
                const PhotonSphere& photon = spheres[ci];
                const Vec3 origin_to_center = photon.position - ray.origin;
                if (hit && length(origin_to_center) > hit->time) continue;
                const double radius_squared = photon.radius * photon.radius;
                const double disk_distance = dot(origin_to_center, ray.dir);
                const double dn = length(ray.at(disk_distance) - photon.position);
                const double distance_squared = dn * dn;
                if (disk_distance > 0.0 && distance_squared < radius_squared) {
                    const double weight = k2(distance_squared / radius_squared) / radius_squared;
                    const Vec3 wi = -photon.direction;
                    const double transmittance = std::exp(-extinction * disk_distance);
                    volume_color += transmittance * cmul(photon.power, medium_color) * medium.phase(wi, -ray.dir) * weight * 1.0;
                }
            }
            return volume_color;
        }
        for (int ci : cand) {
            const PhotonBeam& beam = beams[ci];
            const Vec3 l = beam.start_position - ray.origin;
            const Vec3 u = normalize(cross(l, beam.ray.dir));
            const Vec3 n = normalize(cross(beam.ray.dir, u));
            const double t = dot(n, l) / dot(n, ray.dir);
            const Vec3 query_collision = ray.at(t);
            if (hit && t >= hit->time) continue;
            const double dd = dot(ray.dir, beam.ray.dir);
            const double inv_sin_theta = 1.0 / std::sqrt(std::fmax(0.0, 1.0 - dd * dd));
            const double beam_t = dot(beam.ray.dir, query_collision - beam.start_position);
            const double beam_len = length(beam.end_position - beam.start_position);
            if (beam_t < 0.0 || beam_t > beam_len) continue;
            const Vec3 beam_collision = beam.ray.at(beam_t);
            const double dist = length(query_collision - beam_collision);
            if (dist >= beam.radius) continue;
            const Ray rb{beam_collision, -beam.ray.dir};
            const Color color = extinction * cmul(beam.power, medium_color) * medium.phase(-beam.ray.dir, -ray.dir) * inv_sin_theta *
                                medium.transmittence(ray, t) * medium.transmittence(rb, beam_t) * k2(dist / beam.radius) /
                                (2.0 * beam.radius);
            volume_color += color;
        }
        return volume_color;
    }

    // src/photon.rs:316-627
    Color estimate_indirect(const Renderer& r, const Ray& ray, const Medium* medium, Rng& rng) const {
        const Vec3 wo = -normalize(ray.dir);
        HitRecord h; const Object* object;
        if (!r.get_closest_hit(ray, &h, &object)) {
            if (!medium) return r.scene->environment.get_color(ray.dir);
            return volume_estimate(r, ray, wo, *medium, nullptr, nullptr, rng);
        }
        const Material material = object->material;
        if (!medium) return surface_estimate(r, ray, wo, h, material);
        if (kind == PhotonMapKind) return volume_estimate(r, ray, wo, *medium, &h, object, rng);
        const Color volume_color = volume_estimate(r, ray, wo, *medium, &h, object, rng);
        const Color surface_color = surface_estimate(r, ray, wo, h, material) * medium->transmittence(ray, h.time);
        return surface_color + volume_color;
    }
};

// src/photon.rs:803-946.  Deposits are appended deepest-first as the reference's recursion
// returns them.  Returns through `out`.
inline void trace_photon(const Renderer& r, const Ray& ray, const Vec3& power, Rng& rng, int num_bounces, PhotonList* out) {
    const Vec3 wo = -normalize(ray.dir);
    const Medium* medium = r.scene->media.empty() ? nullptr : &r.scene->media[0];

    auto trace_on_surface = [&](const HitRecord& h, const Object* object) {
        const Vec3 world_pos = ray.at(h.time);
        const Material material = object->material;
        const double p_r = 0.7;  // max(specular + diffuse) with specular = 0, diffuse = 0.7
        const double diffuse_sum = 0.7 + 0.7 + 0.7, specular_sum = 0.0;
        const double p_d = diffuse_sum / (diffuse_sum + specular_sum) * p_r;
        const double russian_roulette = rng.gen();
        if (russian_roulette < p_d) {
            Vec3 wi; double pdf;
            if (material.sample_f(h.normal, wo, rng, &wi, &pdf)) {
                const Color f = material.bsdf(h.normal, wo, wi);
                const Ray next{world_pos, wi};
                const double cosine_term = dot(wi, h.normal) > 0.0 ? dot(wi, h.normal) : 1.0;
                const Vec3 attenuated = cmul(power, f) * cosine_term / pdf / p_d;
                PhotonList deeper;
                trace_photon(r, next, attenuated, rng, num_bounces + 1, &deeper);
                if (!material.is_mirror())
                    deeper.surface.push_back({world_pos, wo, power, next.origin, (uint32_t)num_bounces});
                out->surface.insert(out->surface.end(), deeper.surface.begin(), deeper.surface.end());
                out->volume.insert(out->volume.end(), deeper.volume.begin(), deeper.volume.end());
            }
        }
    };
    auto trace_in_volume = [&](double d) {
        const Vec3 collision = ray.at(d);
        const double abs_ = medium->absorption(collision);
        const Color medium_color = medium->color(collision);
        const double scat = medium->scattering(collision);
        const double extinction = abs_ + scat;
        const Vec3 attenuated = cmul(power, medium_color) * scat / extinction;
        const double rr_prob = scat / extinction;
        if (rng.gen() < rr_prob) {
            Vec3 wi; double ph_p;
            medium->sample_ph(wo, rng, &wi, &ph_p);
            const Ray next{collision, wi};
            trace_photon(r, next, attenuated * medium->phase(wo, wi) / ph_p, rng, num_bounces + 1, out);
        }
        out->volume.push_back({collision, wo, power, ray.origin, (uint32_t)num_bounces});
    };

    HitRecord h; const Object* object;
    if (!r.get_closest_hit(ray, &h, &object)) {
        if (medium) {
            double d, d_pdf, d_cdf;
            medium->sample_d(ray, rng, &d, &d_pdf, &d_cdf);
            trace_in_volume(d);
        }
        return;
    }
    if (medium) {
        double d, d_pdf, d_cdf;
        medium->sample_d(ray, rng, &d, &d_pdf, &d_cdf);
        if (d < h.time) trace_in_volume(d);
        else trace_on_surface(h, object);
    } else {
        trace_on_surface(h, object);
    }
}

// src/photon.rs:724-799.  Returns false when no Light::Object exists (the reference panics).
inline bool shoot_photon(const Renderer& r, double power, uint64_t photon_index, PhotonRenderKind kind, PhotonList* out) {
    Rng rng(r.seed, DOMAIN_PHOTON, photon_index, 0);
    for (const Light& light : r.scene->lights) {
        if (light.kind != Light::ObjectLight) continue;
        Vec3 pos, n; double pdf;
        light.object.shape->sample(Vec3(0, 0, 0), rng, &pos, &n, &pdf);
        const double phi = 2.0 * PI * rng.gen();
        const double theta = std::acos(1.0 - rng.gen());
        const Vec3 hemi(std::sin(theta) * std::cos(phi), std::cos(theta), std::sin(theta) * std::sin(phi));
        Mat3 rot;
        if (!rotation_between(Vec3(0, 1, 0), n, &rot)) rotation_between(Vec3(0.0, 1.0, 0.00000001), n, &rot);
        const Vec3 direction = mul(rot, hemi);
        PhotonList photons;
        trace_photon(r, Ray{pos, direction}, power * light.object.material.color(), rng, 0, &photons);
        for (Photon& x : photons.surface) x.index = photon_index;
        for (Photon& x : photons.volume) x.index = photon_index;
        out->surface.insert(out->surface.end(), photons.surface.begin(), photons.surface.end());
        for (const Photon& x : photons.volume) {
            if (kind == PhotonBeamBeam) {
                const double thresh = 0.001;
                // stream layout is this project's own: one draw keyed by (photon, depth)
                Rng thin(r.seed, DOMAIN_THIN, photon_index, x.depth);
                if (thin.gen() < thresh) {
                    Photon np = x;
                    np.power /= thresh;
                    out->volume.push_back(np);
                }
            } else {
                out->volume.push_back(x);
            }
        }
        return true;
    }
    return false;
}

// photons [first, first + count) of a run of photon_count (the whole run by default)
inline bool emit_photons(const Renderer& r, uint64_t photon_count, PhotonRenderKind kind, PhotonList* list, uint64_t first = 0,
                         uint64_t count = ~0ull) {
    if (count == ~0ull) count = photon_count - first;
    int nt = r.threads > 0 ? r.threads : std::max(1, (int)std::thread::hardware_concurrency());
    if ((uint64_t)nt > count) nt = (int)std::max<uint64_t>(1, count);
    std::vector<PhotonList> parts(nt);
    std::atomic<bool> ok{true};
    const double power = r.watts / (double)photon_count;
    std::vector<std::thread> pool;
    for (int t = 0; t < nt; t++)
        pool.emplace_back([&, t] {
            const uint64_t lo = first + count * t / nt, hi = first + count * (t + 1) / nt;
            for (uint64_t i = lo; i < hi; i++)
                if (!shoot_photon(r, power, i, kind, &parts[t])) { ok = false; return; }
        });
    for (auto& th : pool) th.join();
    if (!ok) return false;
    for (auto& p : parts) {  // PhotonList::merge, src/photon.rs:153-160
        list->surface.insert(list->surface.end(), p.surface.begin(), p.surface.end());
        list->volume.insert(list->volume.end(), p.volume.begin(), p.volume.end());
    }
    return true;
}

// src/photon.rs:950-985: one gather sample
inline Color photon_sample_radiance(const Renderer& r, const PhotonMap& map, uint32_t x, uint32_t y, uint64_t s) {
    Rng rng(r.seed, DOMAIN_GATHER, (uint64_t)y * r.width + x, s);
    const Ray ray = r.primary_ray(x, y, rng);
    return map.estimate_indirect(r, ray, r.scene->media.empty() ? nullptr : &r.scene->media[0], rng);
}

// PASS 3 of photon_render, src/photon.rs:706-716
inline void photon_gather(const Renderer& r, const PhotonMap& map, Buffer& buffer) {
    std::vector<Color> colors((size_t)r.width * r.height);
    r.parallel_rows([&](uint32_t y, int) {
        for (uint32_t x = 0; x < r.width; x++) {
            Color c(0, 0, 0);
            for (uint32_t s = 0; s < r.num_samples; s++) c += photon_sample_radiance(r, map, x, y, s);
            colors[(size_t)y * r.width + x] = c / (double)r.num_samples * std::pow(2.0, r.exposure_value);
        }
    });
    buffer.add_samples(colors);
}

// src/photon.rs:655-720
inline bool photon_render(const Renderer& r, uint64_t photon_count, PhotonRenderKind kind, Buffer& buffer, PhotonStats* st) {
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    PhotonList list;
    if (!emit_photons(r, photon_count, kind, &list)) return false;
    auto t1 = clk::now();
    st->n_surface = list.surface.size();
    st->n_volume = list.volume.size();
    PhotonMap map;
    map.build(kind, std::move(list));
    auto t2 = clk::now();
    photon_gather(r, map, buffer);
    auto t3 = clk::now();
    st->emit_s = std::chrono::duration<double>(t1 - t0).count();
    st->build_s = std::chrono::duration<double>(t2 - t1).count();
    st->gather_s = std::chrono::duration<double>(t3 - t2).count();
    return true;
}

}  // namespace orc

// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// Restates src/renderer.rs (builder fields, sample/get_color/trace_ray/sample_lights/
// sample_lights_for_media/get_closest_hit) and src/buffer.rs of the reference.
// Recursion, linear object loop and one-task-per-row threading are kept as in the
// reference (rayon rows, src/renderer.rs:159-169) so the same code doubles as the timed
// CPU baseline.
#pragma once
#include <atomic>
#include <thread>
#include <vector>

#include "scene.hpp"

namespace orc {

constexpr double EPSILON = 1e-12;        // src/renderer.rs:17
constexpr double FIREFLY_CLAMP = 100.0;  // src/renderer.rs:18

// src/buffer.rs:5-108
struct Buffer {
    uint32_t width, height, filter_radius;
    std::vector<std::vector<Color>> samples;
    Buffer(uint32_t w, uint32_t h, uint32_t r) : width(w), height(h), filter_radius(r), samples((size_t)w * h) {}
    void add_samples(const std::vector<Color>& s) {
        for (size_t i = 0; i < s.size(); i++) samples[i].push_back(s[i]);
    }
    Color filtered(uint32_t x, uint32_t y) const {
        Color color(0, 0, 0);
        size_t count = 0;
        const uint32_t r = filter_radius;
        for (uint32_t i = (x >= r ? x - r : 0); i <= x + r; i++)
            for (uint32_t j = (y >= r ? y - r : 0); j <= y + r; j++)
                if (i < width && j < height) {
                    const auto& px = samples[(size_t)j * width + i];
                    Color sum(0, 0, 0);
                    for (const auto& c : px) sum += c;  // iter().sum::<Color>()
                    color += sum;
                    count += px.size();
                }
        return color / (double)count;
    }
    void image(uint8_t* out) const {
        for (uint32_t y = 0; y < height; y++)
            for (uint32_t x = 0; x < width; x++) color_bytes(filtered(x, y), out + 3 * ((size_t)y * width + x));
    }
    double variance() const {
        double variance = 0.0, count = 0.0;
        for (const auto& px : samples) {
            Color sum(0, 0, 0);
            for (const auto& c : px) sum += c;
            const Color mean = sum / (double)px.size();
            double ss = 0.0;
            for (const auto& c : px) ss += length2(c - mean);
            variance += ss / ((double)px.size() - 1.0);
            count += 1.0;
        }
        return variance / count;
    }
    // mean over batches (unfiltered) — oracle convenience for RMSE checks
    void mean(double* out) const {
        for (size_t i = 0; i < samples.size(); i++) {
            Color sum(0, 0, 0);
            for (const auto& c : samples[i]) sum += c;
            const Color m = samples[i].empty() ? Color(0, 0, 0) : sum / (double)samples[i].size();
            out[3 * i] = m.x; out[3 * i + 1] = m.y; out[3 * i + 2] = m.z;
        }
    }
};

// Ray recorder (oracle addition): collects every closest-hit query a render makes.
struct RayLog {
    std::vector<double> o, d;      // [n][3]
    std::vector<uint8_t> kind;     // 0 primary, 1 bounce, 2 shadow
    size_t cap = 0;
    void push(const Ray& r, uint8_t k) {
        if (kind.size() >= cap) return;
        o.push_back(r.origin.x); o.push_back(r.origin.y); o.push_back(r.origin.z);
        d.push_back(r.dir.x); d.push_back(r.dir.y); d.push_back(r.dir.z);
        kind.push_back(k);
    }
};

struct Counters {
    uint64_t extend_rays = 0, shadow_rays = 0, shadow_pass = 0, shadow_near = 0;
};

struct Renderer {
    const Scene* scene;
    Camera camera;
    uint32_t width = 800, height = 600;
    double exposure_value = 0.0;
    uint32_t filter_radius = 0;
    uint32_t max_bounces = 0, num_samples = 1;
    size_t gather_size = 50, gather_size_volume = 50;
    double watts = 100.0;
    uint64_t seed = 0;
    int threads = 0;  // 0 = hardware_concurrency

    // src/renderer.rs:416-425
    bool get_closest_hit(const Ray& ray, HitRecord* h, const Object** obj, uint32_t* obj_id = nullptr) const {
        *h = HitRecord();
        const Object* hit = nullptr;
        uint32_t id = 0xffffffffu;
        for (size_t i = 0; i < scene->objects.size(); i++) {
            const Object& object = scene->objects[i];
            if (object.shape->intersect(ray, EPSILON, *h)) { hit = &object; id = (uint32_t)i; }
        }
        if (obj_id) *obj_id = id;
        *obj = hit;
        return hit != nullptr;
    }

    // src/renderer.rs:362-409
    Color sample_lights(const Material& material, const Vec3& pos, const Vec3& n, const Vec3& wo, Rng& rng,
                        Counters* ctr, RayLog* log) const {
        Color color(0, 0, 0);
        for (const Light& light : scene->lights) {
            if (light.kind == Light::Ambient) {
                color += cmul(light.color, material.color());
            } else {
                Color intensity; Vec3 wi; double dist = 0.0;
                light.illuminate(pos, rng, &intensity, &wi, &dist);
                const Ray ray{pos, wi};
                HitRecord h; const Object* o;
                if (ctr) ctr->shadow_rays++;
                if (log) log->push(ray, 2);
                if (get_closest_hit(ray, &h, &o)) {
                    if (ctr && std::fabs(h.time - dist) < 1e-6) ctr->shadow_near++;
                    if (std::fabs(h.time - dist) < EPSILON) {
                        if (ctr) ctr->shadow_pass++;
                        const Color f = material.bsdf(n, wo, wi);
                        color += cmul(f, intensity) * dot(wi, n);
                    }
                }
            }
        }
        return color;
    }

    // src/renderer.rs:325-359
    Color sample_lights_for_media(const Medium& medium, const Vec3& pos, const Vec3& wo, Rng& rng, Counters* ctr,
                                  RayLog* log) const {
        Color color(0, 0, 0);
        const double scat = medium.scattering(pos);
        const double ext = medium.extinction(pos);
        const Color medium_color = medium.color(pos);
        for (const Light& light : scene->lights) {
            if (light.kind == Light::Ambient) {
                color += cmul(light.color, medium_color);
            } else {
                Color intensity; Vec3 wi; double dist = 0.0;
                light.illuminate(pos, rng, &intensity, &wi, &dist);
                const Ray ray{pos, wi};
                HitRecord h; const Object* o;
                if (ctr) ctr->shadow_rays++;
                if (log) log->push(ray, 2);
                if (get_closest_hit(ray, &h, &o)) {
                    if (ctr && std::fabs(h.time - dist) < 1e-6) ctr->shadow_near++;
                    if (std::fabs(h.time - dist) < EPSILON) {
                        if (ctr) ctr->shadow_pass++;
                        const double ph = medium.phase(wo, wi);
                        color += (scat / ext) * cmul(intensity, medium_color) * ph;
                    }
                }
            }
        }
        return color;
    }

    // src/renderer.rs:187-322
    Color trace_ray(const Ray& ray, uint32_t num_bounces, Rng& rng, Counters* ctr, RayLog* log) const {
        if (ctr) ctr->extend_rays++;
        if (log) log->push(ray, num_bounces == 0 ? 0 : 1);
        if (!scene->media.empty()) {
            const Medium& medium = scene->media[0];
            const double rr_p = 0.8;
            double d, d_pdf, d_cdf;
            medium.sample_d(ray, rng, &d, &d_pdf, &d_cdf);
            const Vec3 wo = -normalize(ray.dir);
            double max_dist; Color surface_color;
            HitRecord h; const Object* object;
            if (!get_closest_hit(ray, &h, &object)) {
                const double background_dist = 400.0;
                surface_color = d >= background_dist ? scene->environment.get_color(ray.dir) : Color(0, 0, 0);
                max_dist = background_dist;
            } else {
                if (d >= h.time) {
                    const Vec3 world_pos = ray.at(h.time);
                    const Material material = object->material;
                    Color color = num_bounces == 0 ? material.emittance() * material.color() : Color(0, 0, 0);
                    color += sample_lights(material, world_pos, h.normal, wo, rng, ctr, log);
                    if (rng.gen() < rr_p) {
                        Vec3 wi; double pdf;
                        if (material.sample_f(h.normal, wo, rng, &wi, &pdf)) {
                            const Color f = material.bsdf(h.normal, wo, wi);
                            const Ray next{world_pos, wi};
                            const Color indirect = 1.0 / (pdf * rr_p) * cmul(f, trace_ray(next, num_bounces + 1, rng, ctr, log)) *
                                                   std::fabs(dot(wi, h.normal));
                            color += indirect;
                        }
                    }
                    surface_color = color;
                } else {
                    surface_color = Color(0, 0, 0);
                }
                max_dist = h.time;
            }
            if (d < max_dist) {
                const Vec3 collision = ray.at(d);
                const double abs_ = medium.absorption(collision);
                const double emm = medium.emission(collision);
                const Color medium_color = medium.color(collision);
                const double scat = medium.scattering(collision);
                const double extinction = abs_ + scat;
                Color color = num_bounces == 0 ? emm * medium_color : Color(0, 0, 0);
                color += sample_lights_for_media(medium, collision, wo, rng, ctr, log);
                if (rng.gen() < rr_p) {
                    Vec3 wi; double ph_p;
                    medium.sample_ph(wo, rng, &wi, &ph_p);
                    const Ray next{collision, wi};
                    Color indirect = (scat / extinction) * trace_ray(next, num_bounces + 1, rng, ctr, log);
                    indirect /= ph_p;
                    indirect = cmul(indirect, medium_color) * medium.phase(wo, wi);
                    indirect /= rr_p;
                    color += indirect;
                }
                return color;
            }
            return surface_color;
        }
        HitRecord h; const Object* object;
        if (!get_closest_hit(ray, &h, &object)) return scene->environment.get_color(ray.dir);
        const Vec3 world_pos = ray.at(h.time);
        const Material material = object->material;
        const Vec3 wo = -normalize(ray.dir);
        Color color = num_bounces == 0 ? material.emittance() * material.color() : Color(0, 0, 0);
        color += sample_lights(material, world_pos, h.normal, wo, rng, ctr, log);
        if (num_bounces < max_bounces) {
            Vec3 wi; double pdf;
            if (material.sample_f(h.normal, wo, rng, &wi, &pdf)) {
                const Color f = material.bsdf(h.normal, wo, wi);
                const Ray next{world_pos, wi};
                const Color indirect =
                    1.0 / pdf * cmul(f, trace_ray(next, num_bounces + 1, rng, ctr, log)) * std::fabs(dot(wi, h.normal));
                color.x += std::fmin(indirect.x, FIREFLY_CLAMP);  // f64::min: NaN-ignoring like fmin
                color.y += std::fmin(indirect.y, FIREFLY_CLAMP);
                color.z += std::fmin(indirect.z, FIREFLY_CLAMP);
            }
        }
        return color;
    }

    // pixel mapping + jitter of src/renderer.rs:173-181 for ONE sample; the per-(pixel,sample)
    // counter-based stream replaces the per-row StdRng.
    Ray primary_ray(uint32_t x, uint32_t y, Rng& rng) const {
        const double dim = (double)std::max(width, height);
        const double xn = ((double)(2 * x + 1) - (double)width) / dim;
        const double yn = ((double)(2 * (height - y) - 1) - (double)height) / dim;
        const double dx = rng.gen_range(-1.0 / dim, 1.0 / dim);
        const double dy = rng.gen_range(-1.0 / dim, 1.0 / dim);
        return camera.cast_ray(xn + dx, yn + dy, rng);
    }

    Color sample_radiance(uint32_t x, uint32_t y, uint64_t sample_index, Counters* ctr = nullptr, RayLog* log = nullptr) const {
        Rng rng(seed, DOMAIN_CAMERA, (uint64_t)y * width + x, sample_index);
        return trace_ray(primary_ray(x, y, rng), 0, rng, ctr, log);
    }

    // src/renderer.rs:173-184 with samples [first, first+iterations)
    Color get_color(uint32_t x, uint32_t y, uint64_t first, uint32_t iterations, Counters* ctr, RayLog* log) const {
        Color color(0, 0, 0);
        for (uint32_t i = 0; i < iterations; i++) color += sample_radiance(x, y, first + i, ctr, log);
        return color / (double)iterations * std::pow(2.0, exposure_value);
    }

    template <class F>
    void parallel_rows(F&& f) const {
        int nt = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
        if (nt < 1) nt = 1;
        std::atomic<uint32_t> next{0};
        std::vector<std::thread> pool;
        for (int t = 0; t < nt; t++)
            pool.emplace_back([&, t] {
                for (;;) {
                    const uint32_t y = next.fetch_add(1);
                    if (y >= height) break;
                    f(y, t);
                }
            });
        for (auto& th : pool) th.join();
    }

    // src/renderer.rs:158-171
    void sample(uint64_t first, uint32_t iterations, Buffer& buffer, Counters* total = nullptr, RayLog* log = nullptr) const {
        std::vector<Color> colors((size_t)width * height);
        const int nt = threads > 0 ? threads : std::max(1, (int)std::thread::hardware_concurrency());
        std::vector<Counters> ctrs(nt);
        if (log) {  // ray logging is serial so the dump order is deterministic
            for (uint32_t y = 0; y < height; y++)
                for (uint32_t x = 0; x < width; x++)
                    colors[(size_t)y * width + x] = get_color(x, y, first, iterations, &ctrs[0], log);
        } else {
            parallel_rows([&](uint32_t y, int t) {
                for (uint32_t x = 0; x < width; x++)
                    colors[(size_t)y * width + x] = get_color(x, y, first, iterations, total ? &ctrs[t] : nullptr, nullptr);
            });
        }
        if (total)
            for (const auto& c : ctrs) {
                total->extend_rays += c.extend_rays; total->shadow_rays += c.shadow_rays;
                total->shadow_pass += c.shadow_pass; total->shadow_near += c.shadow_near;
            }
        buffer.add_samples(colors);
    }
};

}  // namespace orc

// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// Restates src/material.rs:1-290, src/medium.rs, src/light.rs, src/camera.rs,
// src/environment.rs, src/color.rs, src/object.rs, src/scene.rs of the reference.
#pragma once
#include <memory>
#include <vector>

#include "shape.hpp"

namespace orc {

using Color = Vec3;
constexpr double SRGB_GAMMA = 2.2;

// src/color.rs:10-15
inline Color hex_color(uint32_t x) {
    const double r = (double)((x >> 16) & 0xff) / 255.0;
    const double g = (double)((x >> 8) & 0xff) / 255.0;
    const double b = (double)(x & 0xff) / 255.0;
    return {std::pow(r, SRGB_GAMMA), std::pow(g, SRGB_GAMMA), std::pow(b, SRGB_GAMMA)};
}
// src/color.rs:18-24 (`as u8` truncates and saturates)
inline void color_bytes(const Color& c, uint8_t out[3]) {
    for (int i = 0; i < 3; i++) {
        const double v = std::pow(std::fmin(std::fmax(c[i], 0.0), 1.0), 1.0 / SRGB_GAMMA) * 255.0;
        out[i] = (uint8_t)(v != v ? 0 : (v >= 255.0 ? 255 : (v <= 0.0 ? 0 : (int)v)));
    }
}

// src/material.rs:8-141
struct Material {
    enum Kind { Lambertian = 0, Phong = 1, Mirror = 2, Transmissive = 3 } kind = Lambertian;
    Color albedo{0.5, 0.5, 0.5};
    double emittance_ = 0.0, shininess = 0.0, ior = 1.0;

    double emittance() const { return (kind == Lambertian || kind == Phong) ? emittance_ : 0.0; }
    Color color() const { return (kind == Lambertian || kind == Phong) ? albedo : Color(0, 0, 0); }
    bool is_mirror() const { return kind == Mirror || kind == Transmissive; }

    // src/material.rs:144-162
    static double snell_solve(double ni, double nt, double c) {
        const double r = ni / nt;
        return std::sqrt(1.0 - (r * r) * (1.0 - c * c));
    }
    static Vec3 refract_ray(double ni, double nt, const Vec3& w_i, double ci, double ct, const Vec3& n) {
        return (ni / nt) * (-w_i) + ((ni / nt) * ci - ct) * n;
    }
    static double schlick(double ni, double nt, double c) {
        const double q = (ni - nt) / (ni + nt);
        const double r0 = q * q;
        const double m = 1.0 - c;
        const double m2 = m * m;
        return r0 + (1.0 - r0) * (m * (m2 * m2));  // powi(5) by binary exponentiation (compiler-rt __powidf2)
    }

    // src/material.rs:166-263.  Returns false for None (total internal reflection).
    bool sample_f(const Vec3& normal, const Vec3& wo, Rng& rng, Vec3* wi, double* pdf) const {
        switch (kind) {
            case Lambertian: {
                const double r1 = rng.gen();
                const double r2 = rng.gen();
                const double phi = 2.0 * PI * r1;
                const double theta = std::acos(std::sqrt(r2));
                *pdf = std::cos(theta) / PI;
                const Vec3 hemi(std::sin(theta) * std::cos(phi), std::cos(theta), std::sin(theta) * std::sin(phi));
                Mat3 rot;
                if (!rotation_between(Vec3(0, 1, 0), normal, &rot)) {
                    const bool ok = rotation_between(Vec3(0.0, 1.0, 0.00000001), normal, &rot);
                    (void)ok;
                }
                *wi = normalize(mul(rot, hemi));
                return true;
            }
            case Phong: {
                const double r1 = rng.gen();
                const double r2 = rng.gen();
                const double phi = 2.0 * PI * r1;
                const double theta = std::acos(std::pow(r2, 1.0 / (shininess + 1.0)));
                *pdf = (shininess + 1.0) / (2.0 * PI) * std::pow(std::cos(theta), shininess);
                const Vec3 hemi(std::sin(theta) * std::cos(phi), std::cos(theta), std::sin(theta) * std::sin(phi));
                const Vec3 reflected = -reflect_vec(wo, normal);
                const Quat q = quat_rotation(Vec3(0, 1, 0), reflected);
                *wi = normalize(quat_rotate_vec3(q, hemi));
                return true;
            }
            case Mirror:
                *wi = -reflect_vec(wo, normalize(normal));
                *pdf = 1.0;
                return true;
            case Transmissive: {
                const bool inside = dot(normal, wo) < 0.0;
                const double ci = clampd(dot(wo, inside ? -normal : normal), 0.0, 1.0);
                const double ni = inside ? ior : 1.0, nt = inside ? 1.0 : ior;
                const double ratio = clampd(schlick(ni, nt, ci), 0.0, 1.0);
                if (rng.gen() < ratio) {
                    *wi = -reflect_vec(wo, normal);
                    *pdf = 1.0;
                    return true;
                }
                const double ct = snell_solve(ni, nt, ci);
                if (ct != ct) return false;
                *wi = refract_ray(ni, nt, wo, ci, ct, inside ? -normal : normal);
                *pdf = 1.0;
                return true;
            }
        }
        return false;
    }

    // src/material.rs:266-289
    Color bsdf(const Vec3& normal, const Vec3& wo, const Vec3& wi) const {
        const double n_dot_wi = dot(normal, wi), n_dot_wo = dot(normal, wo);
        if (std::signbit(n_dot_wi) || std::signbit(n_dot_wo)) return {0, 0, 0};
        switch (kind) {
            case Lambertian: return (1.0 / PI) * albedo;
            case Phong: {
                const Color normalization = albedo * ((shininess + 2.0) / (2.0 * PI));
                const Vec3 reflected = normalize(-reflect_vec(wi, normal));
                return normalization * std::pow(clampd(dot(reflected, wo), 0.0, 1.0), shininess);
            }
            default: return {1, 1, 1};
        }
    }
};

// src/object.rs:10-31
struct Object {
    std::shared_ptr<const Shape> shape;
    Material material;
};

// src/medium.rs:9-147 (closed set: the two public constructors)
struct Medium {
    enum Kind { HomogeneousIsotropic = 1, ColoredGlowingFog = 2 } kind = HomogeneousIsotropic;
    double absorption_ = 0, scattering_ = 0;
    double absorption(const Vec3&) const { return absorption_; }
    double scattering(const Vec3&) const { return scattering_; }
    double emission(const Vec3&) const { return kind == ColoredGlowingFog ? 10.0 : 0.0; }
    Color color(const Vec3& p) const {
        if (kind == ColoredGlowingFog) return p.y > 250.0 ? hex_color(0xFF0000) : hex_color(0x0000FF);
        return hex_color(0xD2B48C);
    }
    double extinction(const Vec3& p) const { return absorption(p) + scattering(p); }
    // src/medium.rs:86,114: 1/(4 pi) vs the precedence slip 1.0/4.0*pi
    double phase(const Vec3&, const Vec3&) const {
        return kind == ColoredGlowingFog ? 1.0 / 4.0 * PI : 1.0 / (4.0 * PI);
    }
    // src/medium.rs:87-94,115-121: normalised point of the cube [-1,1)^3
    void sample_ph(const Vec3&, Rng& rng, Vec3* wi, double* p) const {
        Vec3 w;
        w.x = rng.gen_range(-1.0, 1.0);
        w.y = rng.gen_range(-1.0, 1.0);
        w.z = rng.gen_range(-1.0, 1.0);
        *wi = normalize(w);
        *p = phase(w, w);
    }
    // src/medium.rs:126-130
    double transmittence(const Ray& ray, double t_max) const { return std::exp(-(extinction(ray.origin) * t_max)); }
    // src/medium.rs:133-146
    void sample_d(const Ray& ray, Rng& rng, double* dist, double* pdf, double* cdf) const {
        const double random = rng.gen_range(0.0, 1.0);
        const double ext = extinction(ray.origin);
        *dist = -std::log(random) / ext;
        const double tr = transmittence(ray, *dist);
        *pdf = ext * tr;
        *cdf = 1.0 - tr;
    }
};

// src/light.rs:7-47
struct Light {
    enum Kind { Point = 0, Ambient = 1, Directional = 2, ObjectLight = 3 } kind = Ambient;
    Color color;
    Vec3 vec;
    Object object;
    void illuminate(const Vec3& world_pos, Rng& rng, Color* intensity, Vec3* wi, double* dist) const {
        switch (kind) {
            case Ambient: *intensity = color; *wi = Vec3(0, 0, 0); *dist = 0.0; return;
            case Point: {
                const Vec3 disp = vec - world_pos;
                const double len = length(disp);
                *intensity = color / (len * len); *wi = disp / len; *dist = len;
                return;
            }
            case Directional: *intensity = color; *wi = -normalize(vec); *dist = INF; return;
            case ObjectLight: {
                Vec3 v, n; double p;
                object.shape->sample(world_pos, rng, &v, &n, &p);
                const Vec3 disp = v - world_pos;
                const double len = length(disp);
                const double cosine = std::fmax(-dot(disp, n), 0.0) / len;
                const double surface_area = std::fmax(cosine, 0.0) / (len * len);
                *intensity = object.material.color() * object.material.emittance() * surface_area / p;
                *wi = disp / len; *dist = len;
                return;
            }
        }
    }
};

// src/camera.rs:9-82
struct Camera {
    Vec3 eye{0, 0, 10}, direction{0, 0, -1}, up{0, 1, 0};
    double fov = PI / 6.0, aperture = 0.0, focal_distance = 0.0;
    Ray cast_ray(double x, double y, Rng& rng) const {
        const double d = 1.0 / std::tan(fov / 2.0);
        const Vec3 right = normalize(cross(direction, up));
        Vec3 origin = eye;
        Vec3 new_dir = d * direction + x * right + y * up;
        if (aperture > 0.0) {
            const Vec3 focal_point = origin + normalize(new_dir) * focal_distance;
            double ax, ay;
            rng.unit_disc(&ax, &ay);
            origin += (ax * right + ay * up) * aperture;
            new_dir = focal_point - origin;
        }
        return {origin, normalize(new_dir)};
    }
};

// src/environment.rs:5-77
struct Environment {
    bool is_hdri = false;
    Color color{0, 0, 0};
    uint32_t width = 0, height = 0;
    std::vector<Color> buf;
    static Color mix(const Color& a, const Color& b, double t) { return a * (1.0 - t) + b * t; }  // glm::mix
    Color get_color(const Vec3& dir_in) const {
        if (!is_hdri) return color;
        const Vec3 dir = normalize(dir_in);
        const double azimuth = std::atan2(dir.z, dir.x) + PI;
        const double polar = std::acos(dir.y);
        const double x = azimuth / (2.0 * PI) * (double)(width - 1);
        const double y = polar / PI * (double)(height - 1);
        const uint32_t x0 = std::min((uint32_t)x, width - 1), y0 = std::min((uint32_t)y, height - 1);
        const double ax = x - (double)x0, ay = y - (double)y0;
        auto at = [&](size_t i) { return i < buf.size() ? buf[i] : Color(0, 0, 0); };  // the reference would panic out of range
        return mix(mix(at((size_t)y0 * width + x0), at((size_t)y0 * width + x0 + 1), ax),
                   mix(at((size_t)(y0 + 1) * width + x0), at((size_t)(y0 + 1) * width + x0 + 1), ax), ay);
    }
};

// src/scene.rs:12-24
struct Scene {
    std::vector<Object> objects;
    std::vector<Light> lights;
    std::vector<Medium> media;
    Environment environment;
};

}  // namespace orc

// f64 device restatement of the shading maths on rpt's path-tracing hot path:
// Material::{sample_f,bsdf,emittance,color} (src/material.rs:100-289), Light::illuminate
// (src/light.rs:23-47) with Shape::sample (src/shape/{sphere,cube,mesh}.rs, src/kdtree.rs:141-146,
// src/shape.rs:140-151), Medium (src/medium.rs:80-146), Camera::cast_ray (src/camera.rs:65-82),
// Environment::get_color (src/environment.rs:25-52,72-77).  RNG draw order is part of the
// contract with the oracle (same seed => same sample).
#pragma once
#include "common.cuh"
#include "rng.cuh"

namespace rptx {

RPTX_D d3 mat_color(const Material& m) {
    return (m.kind == 0 || m.kind == 1) ? mk3(m.albedo[0], m.albedo[1], m.albedo[2]) : mk3(0, 0, 0);
}
RPTX_D double mat_emittance(const Material& m) { return (m.kind == 0 || m.kind == 1) ? m.emittance : 0.0; }
RPTX_D bool mat_is_mirror(const Material& m) { return m.kind == 2 || m.kind == 3; }

struct Rot3 {
    double m[9];
};
RPTX_D d3 rot_apply(const Rot3& r, d3 v) {
    return mk3(r.m[0] * v.x + r.m[1] * v.y + r.m[2] * v.z, r.m[3] * v.x + r.m[4] * v.y + r.m[5] * v.z,
               r.m[6] * v.x + r.m[7] * v.y + r.m[8] * v.z);
}
// Trigonometry.  The reference goes through angles (theta = acos(x); cos(theta), sin(theta);
// rotation_between = from_axis_angle(axis, acos(a.b))).  Here cos(acos(x)) is x, sin(acos(x)) is
// sqrt(1 - x^2) and the sine of the angle between unit vectors is |a x b|: the same values to a
// few ulps — the size of the CUDA-vs-glibc libm differences the parity tests already allow —
// without the f64 acos/sin/cos code (each several KB of SASS; k_shade was 105-143 KB and
// instruction-cache bound).
RPTX_D double sin_from_cos(double c) { return DSQRT(fmax(0.0, 1.0 - c * c)); }

// nalgebra Rotation3::from_axis_angle with (sin, cos) of the angle
RPTX_D Rot3 rot_identity() {
    Rot3 r;
    r.m[0] = 1; r.m[1] = 0; r.m[2] = 0; r.m[3] = 0; r.m[4] = 1; r.m[5] = 0; r.m[6] = 0; r.m[7] = 0; r.m[8] = 1;
    return r;
}
RPTX_D Rot3 rot_from_axis_sincos(d3 u, double s, double c) {
    Rot3 r;
    const double sqx = u.x * u.x, sqy = u.y * u.y, sqz = u.z * u.z;
    const double omc = 1.0 - c;
    r.m[0] = sqx + (1.0 - sqx) * c;
    r.m[1] = u.x * u.y * omc - u.z * s;
    r.m[2] = u.x * u.z * omc + u.y * s;
    r.m[3] = u.x * u.y * omc + u.z * s;
    r.m[4] = sqy + (1.0 - sqy) * c;
    r.m[5] = u.y * u.z * omc - u.x * s;
    r.m[6] = u.x * u.z * omc - u.y * s;
    r.m[7] = u.y * u.z * omc + u.x * s;
    r.m[8] = sqz + (1.0 - sqz) * c;
    return r;
}
// nalgebra Rotation3::rotation_between; false = None (antiparallel)
RPTX_D bool rotation_between(d3 a, d3 b, Rot3* out) {
    const double la = length(a), lb = length(b);
    if (la > 0.0 && lb > 0.0) {
        const d3 na = a / la, nb = b / lb;
        const d3 c = cross(na, nb);
        const double sq = dot(c, c);
        const double eps = 2.220446049250313e-16;
        if (sq > eps * eps) {
            // angle = acos(na.nb): its cosine is the dot product and its sine follows from the
            // cosine alone, as in the reference (|na x nb| would be MORE accurate for nearly
            // (anti)parallel vectors, e.g. the (0,1,1e-8) fallback against -Y, and so differ from it)
            const double cs = dot(na, nb);
            *out = rot_from_axis_sincos(c / DSQRT(sq), sin_from_cos(cs), cs);
            return true;
        }
        if (dot(na, nb) < 0.0) return false;
    }
    *out = rot_identity();
    return true;
}
// the reference's +Y -> n rotation with its antiparallel fallback (src/material.rs:186-194)
RPTX_D Rot3 rotation_from_y(d3 n) {
    Rot3 r;
    if (!rotation_between(mk3(0, 1, 0), n, &r)) rotation_between(mk3(0.0, 1.0, 0.00000001), n, &r);
    return r;
}

// Material::sample_f.  false = None.
RPTX_DN bool mat_sample_f(const Material& m, d3 normal, d3 wo, Rng& rng, d3* wi, double* pdf) {
    if (m.kind == 0) {
        const double r1 = rng.gen();
        const double r2 = rng.gen();
        const double phi = 2.0 * kPi * r1;
        const double ct = DSQRT(r2), st = sin_from_cos(ct);  // theta = acos(sqrt(r2))
        *pdf = DDIV(ct, kPi);
        double sp, cp;
        sincos(phi, &sp, &cp);
        const d3 hemi = mk3(st * cp, ct, st * sp);
        const Rot3 rot = rotation_from_y(normal);
        *wi = normalize(rot_apply(rot, hemi));
        return true;
    }
    if (m.kind == 1) {
        const double r1 = rng.gen();
        const double r2 = rng.gen();
        const double phi = 2.0 * kPi * r1;
        const double ct = pow(r2, 1.0 / (m.shininess + 1.0)), st = sin_from_cos(ct);  // theta = acos(r2^(1/(s+1)))
        *pdf = (m.shininess + 1.0) / (2.0 * kPi) * pow(ct, m.shininess);
        double sp, cp;
        sincos(phi, &sp, &cp);
        const d3 hemi = mk3(st * cp, ct, st * sp);
        const d3 reflected = -reflect_vec(wo, normal);
        // glm::quat_rotation(+Y, reflected) then quat_rotate_vec3
        double qw = 1.0;
        d3 qv = mk3(0, 0, 0);
        const double lb = length(reflected);
        if (lb > 0.0) {
            const d3 na = mk3(0, 1, 0), nb = reflected / lb;
            const d3 c = cross(na, nb);
            const double sq = dot(c, c);
            const double eps = 2.220446049250313e-16;
            if (sq > eps * eps) {
                const double cs = dot(na, nb);
                if (cs > -1.0 && cs < 1.0) {
                    // half = acos(cs) / 2: cos(half) = sqrt((1 + cs) / 2), sin(half) = sqrt((1 - cs) / 2)
                    qw = sqrt((1.0 + cs) * 0.5);
                    qv = (c / sqrt(sq)) * sqrt((1.0 - cs) * 0.5);
                }
            }
        }
        const d3 t = cross(qv, hemi) * 2.0;
        const d3 cc = cross(qv, t);
        *wi = normalize(t * qw + cc + hemi);
        return true;
    }
    if (m.kind == 2) {
        *wi = -reflect_vec(wo, normalize(normal));
        *pdf = 1.0;
        return true;
    }
    const bool inside = dot(normal, wo) < 0.0;
    const d3 nn = inside ? -normal : normal;
    const double ci = clampd(dot(wo, nn), 0.0, 1.0);
    const double ni = inside ? m.ior : 1.0, nt = inside ? 1.0 : m.ior;
    const double q = (ni - nt) / (ni + nt);
    const double r0 = q * q;
    const double mm = 1.0 - ci;
    const double m2 = mm * mm;
    const double ratio = clampd(r0 + (1.0 - r0) * (mm * (m2 * m2)), 0.0, 1.0);
    if (rng.gen() < ratio) {
        *wi = -reflect_vec(wo, normal);
        *pdf = 1.0;
        return true;
    }
    const double rr = ni / nt;
    const double ct = sqrt(1.0 - (rr * rr) * (1.0 - ci * ci));
    if (ct != ct) return false;
    *wi = (ni / nt) * (-wo) + ((ni / nt) * ci - ct) * nn;
    *pdf = 1.0;
    return true;
}

// Material::bsdf
RPTX_D d3 mat_bsdf(const Material& m, d3 normal, d3 wo, d3 wi) {
    const double ndwi = dot(normal, wi), ndwo = dot(normal, wo);
    if (signbit(ndwi) || signbit(ndwo)) return mk3(0, 0, 0);
    const d3 albedo = mk3(m.albedo[0], m.albedo[1], m.albedo[2]);
    if (m.kind == 0) return (1.0 / kPi) * albedo;
    if (m.kind == 1) {
        const d3 normalization = albedo * ((m.shininess + 2.0) / (2.0 * kPi));
        const d3 reflected = normalize(-reflect_vec(wi, normal));
        return normalization * pow(clampd(dot(reflected, wo), 0.0, 1.0), m.shininess);
    }
    return mk3(1, 1, 1);
}

// ---- medium (closed set of the two public constructors) ----
RPTX_D d3 hex_color_dev(uint32_t x) {
    const double r = (double)((x >> 16) & 0xff) / 255.0, g = (double)((x >> 8) & 0xff) / 255.0, b = (double)(x & 0xff) / 255.0;
    return mk3(pow(r, 2.2), pow(g, 2.2), pow(b, 2.2));
}
struct MediumDev {
    int kind;
    double sigma_a, sigma_s;
    d3 color_lo, color_hi;  // colour below / above y = 250 (same for homogeneous_isotropic)
    double emission, phase;
    RPTX_D d3 color(d3 p) const { return (kind == 2 && p.y > 250.0) ? color_hi : color_lo; }
    RPTX_D double extinction() const { return sigma_a + sigma_s; }
};
RPTX_D void medium_sample_ph(Rng& rng, d3* wi) {
    d3 w;
    w.x = rng.gen_range(-1.0, 1.0);
    w.y = rng.gen_range(-1.0, 1.0);
    w.z = rng.gen_range(-1.0, 1.0);
    *wi = normalize(w);
}
RPTX_D double medium_sample_d(const MediumDev& md, Rng& rng) {
    const double random = rng.gen_range(0.0, 1.0);
    return DDIV(-log(random), md.extinction());
}

// ---- light sampling ----
RPTX_D void sample_light_shape(const SceneDev& sc, const Light& l, d3 target, Rng& rng, d3* v, d3* n, double* p) {
    d3 tl = target;
    if (l.has_transform) tl = xform_point(l.inv, target);
    d3 lv, ln;
    double lp;
    if (l.shape_kind == kMesh) {
        const uint64_t idx = rng.gen_index(l.ntri);
        const RawTri* __restrict__ t = sc.light_tris + l.tri_base + idx;
        double u = rng.gen(), vv = rng.gen();
        while (u + vv > 1.0) { u = rng.gen(); vv = rng.gen(); }
        const double w = 1.0 - u - vv;
        const d3 v1 = mk3(t->v[0], t->v[1], t->v[2]), v2 = mk3(t->v[3], t->v[4], t->v[5]), v3 = mk3(t->v[6], t->v[7], t->v[8]);
        const d3 n1 = mk3(t->n[0], t->n[1], t->n[2]), n2 = mk3(t->n[3], t->n[4], t->n[5]), n3 = mk3(t->n[6], t->n[7], t->n[8]);
        const double area = 0.5 * length(cross(v2 - v1, v3 - v1));
        lv = u * v1 + vv * v2 + w * v3;
        ln = normalize(u * n1 + vv * n2 + w * n3);
        lp = DDIV(DDIV(1.0, area), (double)l.ntri);
    } else if (l.shape_kind == kCube) {
        const double a = rng.gen() - 0.5;
        const double b = rng.gen() - 0.5;
        const uint64_t f = rng.gen_index(6);
        if (f == 0) { lv = mk3(a, b, 0.5); ln = mk3(0, 0, 1); }
        else if (f == 1) { lv = mk3(a, b, -0.5); ln = mk3(0, 0, -1); }
        else if (f == 2) { lv = mk3(a, 0.5, b); ln = mk3(0, 1, 0); }
        else if (f == 3) { lv = mk3(a, -0.5, b); ln = mk3(0, -1, 0); }
        else if (f == 4) { lv = mk3(0.5, a, b); ln = mk3(1, 0, 0); }
        else { lv = mk3(-0.5, a, b); ln = mk3(-1, 0, 0); }
        lp = 1.0 / 6.0;
    } else if (l.shape_kind == kMonomial) {  // src/shape/monomial_surface.rs:109-124
        double x, z;
        rng.unit_circle(&x, &z);
        const double h = l.vec[0];
        lv = mk3(x, h * pow(x * x + z * z, l.vec[1] / 2.), z);
        const double s = lv.x * lv.x + lv.z * lv.z;
        ln = normalize(mk3(h * 4. * lv.x * s, -1., h * 4. * lv.z * s));
        if (rng.gen() < 0.5) ln = -ln;
        lp = 1. / (2. * 6.3406654362);
    } else {  // sphere, src/shape/sphere.rs:53-65
        double x, y;
        rng.unit_disc(&x, &y);
        const double z = sqrt(1.0 - x * x - y * y);
        const d3 nn = normalize(tl);
        d3 n1;
        const double ax = fabs(nn.x);
        const bool is_normal = ax >= 2.2250738585072014e-308 && ax < CUDART_INF;  // f64::is_normal
        if (is_normal) n1 = normalize(mk3(nn.y, -nn.x, 0.0));
        else n1 = normalize(mk3(0.0, -nn.z, nn.y));
        const d3 n2 = cross(n1, nn);
        lv = x * n1 + y * n2 + z * nn;
        ln = lv;
        lp = z * (1.0 / kPi);
    }
    if (l.has_transform) {
        const d3 new_normal = normalize(mul3(l.nrm, ln));
        const double height = dot(mul3(l.lin, ln), new_normal);
        const double base = DDIV(l.scale, height);
        *v = xform_point(l.fwd, lv);
        *n = new_normal;
        *p = DDIV(lp, base);
    } else {
        *v = lv; *n = ln; *p = lp;
    }
}

// Light::illuminate for non-ambient lights
RPTX_DN void illuminate(const SceneDev& sc, const Light& l, d3 pos, Rng& rng, d3* intensity, d3* wi, double* dist) {
    const d3 color = mk3(l.color[0], l.color[1], l.color[2]);
    if (l.kind == 0) {  // Point
        const d3 disp = mk3(l.vec[0], l.vec[1], l.vec[2]) - pos;
        const double len = length(disp);
        *intensity = color / (len * len);
        *wi = disp / len;
        *dist = len;
    } else if (l.kind == 2) {  // Directional
        *intensity = color;
        *wi = -normalize(mk3(l.vec[0], l.vec[1], l.vec[2]));
        *dist = CUDART_INF;
    } else {  // Object
        d3 v, n;
        double p;
        sample_light_shape(sc, l, pos, rng, &v, &n, &p);
        const d3 disp = v - pos;
        const double len = length(disp);
        const double cosine = DDIV(fmax(-dot(disp, n), 0.0), len);
        const double surface_area = DDIV(fmax(cosine, 0.0), len * len);
        *intensity = mat_color(l.material) * mat_emittance(l.material) * surface_area / p;
        *wi = disp / len;
        *dist = len;
    }
}

// Camera::cast_ray with the pixel mapping and jitter of Renderer::get_color (src/renderer.rs:173-181)
RPTX_D void primary_ray(const Camera& cam, uint32_t width, uint32_t height, uint32_t x, uint32_t y, Rng& rng, d3* o, d3* d) {
    const double dim = (double)(width > height ? width : height);
    const double xn = DDIV((double)(2 * x + 1) - (double)width, dim);
    const double yn = DDIV((double)(2 * (height - y) - 1) - (double)height, dim);
    const double jl = DDIV(-1.0, dim), jh = DDIV(1.0, dim);
    const double dx = rng.gen_range(jl, jh);
    const double dy = rng.gen_range(jl, jh);
    const double px = xn + dx, py = yn + dy;
    const double dd = cam.inv_tan_half_fov;
    const d3 right = normalize(cross(cam.direction, cam.up));
    d3 origin = cam.eye;
    d3 new_dir = dd * cam.direction + px * right + py * cam.up;
    if (cam.aperture > 0.0) {
        const d3 focal_point = origin + normalize(new_dir) * cam.focal_distance;
        double ax, ay;
        rng.unit_disc(&ax, &ay);
        origin = origin + (ax * right + ay * cam.up) * cam.aperture;
        new_dir = focal_point - origin;
    }
    *o = origin;
    *d = normalize(new_dir);
}

// Environment::get_color
RPTX_D d3 env_color(const SceneDev& sc, d3 dir_in) {
    if (!sc.env_is_hdri) return mk3(sc.env_color[0], sc.env_color[1], sc.env_color[2]);
    const d3 dir = normalize(dir_in);
    const double azimuth = atan2(dir.z, dir.x) + kPi;
    const double polar = acos(dir.y);
    const double x = azimuth / (2.0 * kPi) * (double)(sc.env_w - 1);
    const double y = polar / kPi * (double)(sc.env_h - 1);
    uint32_t x0 = (uint32_t)x, y0 = (uint32_t)y;
    if (x0 > sc.env_w - 1) x0 = sc.env_w - 1;
    if (y0 > sc.env_h - 1) y0 = sc.env_h - 1;
    const double ax = x - (double)x0, ay = y - (double)y0;
    const size_t npx = (size_t)sc.env_w * sc.env_h;
    auto at = [&](size_t i) {
        if (i >= npx) return mk3(0, 0, 0);  // the reference would panic here
        return mk3(sc.env_pixels[3 * i], sc.env_pixels[3 * i + 1], sc.env_pixels[3 * i + 2]);
    };
    const d3 a = at((size_t)y0 * sc.env_w + x0) * (1.0 - ax) + at((size_t)y0 * sc.env_w + x0 + 1) * ax;
    const d3 b = at((size_t)(y0 + 1) * sc.env_w + x0) * (1.0 - ax) + at((size_t)(y0 + 1) * sc.env_w + x0 + 1) * ax;
    return a * (1.0 - ay) + b * ay;
}

}  // namespace rptx

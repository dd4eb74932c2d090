// ORACLE — TEST INFRASTRUCTURE ONLY (see vec.hpp header).
// The BASELINE scenes built NATIVELY from the literals of the reference's example sources, with the reference's
// own construction path restated: polygon() fan triangulation (src/shape.rs:304-311), the Transformable chain
// (`glm::translate(&identity, v) * self.transform`, src/shape.rs:238-289), SceneAdd for (Mesh, Material) = object AND
// Light::Object copy (src/scene.rs:57-65), KdTree::new over boxed Transformed<Arc<Mesh>> (examples/fractal_teapots.rs).
// Nothing here goes through the Python host's lowering (rpt_b200/world.py, scenes.py): tests/test_oracle_native.py
// asserts that closest hits and renders on these scenes equal the ones on the Python-lowered scenes, which closes the
// common-mode hole of feeding one lowering to both the checker and the product.
//   cornell            examples/cornell.rs:15-83
//   teapot             examples/teapot.rs:10-27
//   fractal_teapots    examples/fractal_teapots.rs:8-81
//   lampshade_*        examples/volumetric_{pathtrace,photonphoton,beamphoton,beambeam}_lampshade.rs
#pragma once
#include <cmath>
#include <memory>
#include <string>
#include <vector>

#include "scene.hpp"

namespace orc {
namespace builtin {

inline Mat4 identity4() {
    Mat4 m;
    for (int r = 0; r < 4; r++)
        for (int c = 0; c < 4; c++) m.m[r][c] = r == c ? 1.0 : 0.0;
    return m;
}
// nalgebra matrix product: column j of the result accumulates a's columns in order k = 0..3
inline Mat4 matmul(const Mat4& a, const Mat4& b) {
    Mat4 out;
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            double s = 0.0;
            for (int k = 0; k < 4; k++) s += a.m[i][k] * b.m[k][j];
            out.m[i][j] = s;
        }
    return out;
}
inline Mat4 translation(const Vec3& v) {  // glm::translate(&identity, v)
    Mat4 m = identity4();
    m.m[0][3] = v.x; m.m[1][3] = v.y; m.m[2][3] = v.z;
    return m;
}
inline Mat4 scaling(const Vec3& v) {  // glm::scale(&identity, v)
    Mat4 m = identity4();
    m.m[0][0] = v.x; m.m[1][1] = v.y; m.m[2][2] = v.z;
    return m;
}
inline Mat4 rotation_y(double angle) {  // glm::rotate_y(&identity, angle) = Rotation3::from_axis_angle(+Y, angle).to_homogeneous()
    const Mat3 r = rot_from_axis_angle(normalize(Vec3(0.0, 1.0, 0.0)), angle);
    Mat4 m = identity4();
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) m.m[i][j] = r.m[i][j];
    return m;
}

// a shape under construction: Transformable's chain keeps ONE matrix (src/shape.rs:238-289)
struct Xf {
    std::shared_ptr<const Shape> shape;
    Mat4 m = identity4();
    bool any = false;
    Xf scale(const Vec3& v) const { Xf r = *this; r.m = matmul(scaling(v), m); r.any = true; return r; }
    Xf rotate_y(double a) const { Xf r = *this; r.m = matmul(rotation_y(a), m); r.any = true; return r; }
    Xf translate(const Vec3& v) const { Xf r = *this; r.m = matmul(translation(v), m); r.any = true; return r; }
    std::shared_ptr<const Shape> build() const {
        if (!any) return shape;
        return std::make_shared<Transformed>(shape, m);
    }
};
inline Xf xf(std::shared_ptr<const Shape> s) { Xf r; r.shape = std::move(s); return r; }

// polygon(), src/shape.rs:304-311: a fan of Triangle::from_vertices
inline std::shared_ptr<const Shape> polygon(const std::vector<Vec3>& v) {
    std::vector<std::shared_ptr<const Shape>> tris;
    for (size_t i = 1; i + 1 < v.size(); i++) {
        auto t = std::make_shared<Triangle>();
        *t = Triangle::from_vertices(v[0], v[i], v[i + 1]);
        tris.push_back(t);
    }
    return std::make_shared<KdTree>(std::move(tris), false);
}

inline std::shared_ptr<const Shape> mesh_from(const double* v, const double* n, size_t ntri) {
    std::vector<std::shared_ptr<const Shape>> tris;
    tris.reserve(ntri);
    for (size_t i = 0; i < ntri; i++) {
        const double *p = v + 9 * i, *q = n + 9 * i;
        auto t = std::make_shared<Triangle>();
        t->v1 = Vec3(p[0], p[1], p[2]); t->v2 = Vec3(p[3], p[4], p[5]); t->v3 = Vec3(p[6], p[7], p[8]);
        t->n1 = Vec3(q[0], q[1], q[2]); t->n2 = Vec3(q[3], q[4], q[5]); t->n3 = Vec3(q[6], q[7], q[8]);
        tris.push_back(t);
    }
    return std::make_shared<KdTree>(std::move(tris), false);
}

inline Material diffuse(const Color& c) { Material m; m.kind = Material::Lambertian; m.albedo = c; return m; }              // src/material.rs:32-41
inline Material light(const Color& c, double e) { Material m = diffuse(c); m.emittance_ = e; return m; }                   // :90-97
inline Material phong(const Color& c, double rough) { Material m; m.kind = Material::Phong; m.albedo = c; m.shininess = rough; return m; }  // :44-50,82-88

inline void add(Scene& s, std::shared_ptr<const Shape> sh, const Material& m) { s.objects.push_back({std::move(sh), m}); }
// SceneAdd<(Mesh, Material)>, src/scene.rs:57-65: the object, then a Light::Object holding a copy
inline void add_light_object(Scene& s, std::shared_ptr<const Shape> sh, const Material& m) {
    s.objects.push_back({sh, m});
    Light l;
    l.kind = Light::ObjectLight;
    l.object = {sh, m};
    s.lights.push_back(l);
}

inline void cornell_walls(Scene& s, const Material& white, const Material& red, const Material& green) {
    add(s, polygon({{0.0, 0.0, 0.0}, {0.0, 0.0, 559.2}, {556.0, 0.0, 559.2}, {556.0, 0.0, 0.0}}), white);                 // floor
    add(s, polygon({{0.0, 548.9, 0.0}, {556.0, 548.9, 0.0}, {556.0, 548.9, 559.2}, {0.0, 548.9, 559.2}}), white);         // ceiling
    add(s, polygon({{0.0, 0.0, 559.2}, {0.0, 548.9, 559.2}, {556.0, 548.9, 559.2}, {556.0, 0.0, 559.2}}), white);         // back wall
    add(s, polygon({{556.0, 0.0, 0.0}, {556.0, 0.0, 559.2}, {556.0, 548.9, 559.2}, {556.0, 548.9, 0.0}}), red);           // left wall
    add(s, polygon({{0.0, 0.0, 0.0}, {0.0, 548.9, 0.0}, {0.0, 548.9, 559.2}, {0.0, 0.0, 559.2}}), green);                 // right wall
}

// examples/cornell.rs:15-83 (its `small_box` is a sphere)
inline void cornell(Scene& s) {
    const Material white = diffuse(hex_color(0xAAAAAA)), red = diffuse(hex_color(0xBC0000)), green = diffuse(hex_color(0x00BC00));
    cornell_walls(s, white, red, green);
    add(s, xf(std::make_shared<Cube>()).scale({165.0, 330.0, 165.0}).rotate_y(2.0 * PI * (-253.0 / 360.0)).translate({368.0, 165.0, 351.0}).build(), white);
    add(s, xf(std::make_shared<Sphere>()).scale({80.0, 80.0, 80.0}).rotate_y(2.0 * PI * (-197.0 / 360.0)).translate({150.0, 82.5, 450.0}).build(), white);
    add_light_object(s, polygon({{343.0, 548.8, 227.0}, {343.0, 548.8, 332.0}, {213.0, 548.8, 332.0}, {213.0, 548.8, 227.0}}), light(hex_color(0xFFFEFA), 100.0));
}

// examples/volumetric_*_lampshade.rs; variant 0 pathtrace, 1 photonphoton, 2 beamphoton, 3 beambeam
inline void lampshade(Scene& s, int variant) {
    const double absorb[4] = {0.00005, 0.0008, 0.0001, 0.0001}, scat[4] = {0.003, 0.0008, 0.001, 0.001};
    const double emit[4] = {150.0, 120.0, 200000.0 / (130.0 * 105.0), 200000.0 / (130.0 * 105.0)};
    const Material white = diffuse(hex_color(0xAAAAAA)), red = diffuse(hex_color(0xBC0000)), yellow = diffuse(hex_color(0xBCBC00)),
                   green = diffuse(hex_color(0x00BC00));
    const double height = 140.0, depth = 105.0, width = 130.0, off = 10.0;
    const Vec3 center(213.0 + 65.0, 548.0, 227.0 + 55.0);
    const Vec3 front(center.x, center.y, center.z + depth / 2.0), left(center.x - width / 2.0, center.y, center.z),
               back(center.x, center.y, center.z - depth / 2.0), right(center.x + width / 2.0, center.y, center.z);
    auto cube = [] { return xf(std::make_shared<Cube>()); };
    cornell_walls(s, white, red, green);
    add(s, cube().scale({165.0, 330.0, 165.0}).rotate_y(2.0 * PI * (-253.0 / 360.0)).translate({368.0, 165.0, 351.0}).build(), white);
    add(s, cube().scale({165.0, 165.0, 165.0}).rotate_y(2.0 * PI * (-197.0 / 360.0)).translate({185.0, 82.5, 169.0}).build(), white);
    add(s, cube().scale({off, height, 105.0 + off * 2.0}).translate(right).build(), yellow);
    add(s, cube().scale({off, height, 105.0 + off * 2.0}).translate(left).build(), yellow);
    add(s, cube().scale({130.0 + off * 2.0, height, off}).translate(front).build(), yellow);
    add(s, cube().scale({130.0 + off * 2.0, height, off}).translate(back).build(), yellow);
    add_light_object(s, polygon({{330.0, 548.8, 240.0}, {330.0, 548.8, 319.0}, {226.0, 548.8, 319.0}, {226.0, 548.8, 240.0}}),
                     light(hex_color(0xFFFEFA), emit[variant]));
    Medium m;
    m.kind = Medium::HomogeneousIsotropic; m.absorption_ = absorb[variant]; m.scattering_ = scat[variant];
    s.media.push_back(m);
}

inline Light ambient(const Color& c) { Light l; l.kind = Light::Ambient; l.color = c; return l; }
inline Light point(const Color& c, const Vec3& p) { Light l; l.kind = Light::Point; l.color = c; l.vec = p; return l; }
inline Light directional(const Color& c, const Vec3& d) { Light l; l.kind = Light::Directional; l.color = c; l.vec = d; return l; }

// examples/teapot.rs:10-27
inline void teapot(Scene& s, std::shared_ptr<const Shape> mesh) {
    add(s, xf(mesh).scale({0.5, 0.5, 0.5}).translate({0.0, -1.0, 0.0}).build(), phong(hex_color(0xFF0000), 0.4));
    add(s, std::make_shared<Plane>(Vec3(0.0, 1.0, 0.0), -1.0), diffuse(hex_color(0xAAAAAA)));
    s.lights.push_back(ambient({0.02, 0.02, 0.02}));
    s.lights.push_back(point({60.0, 60.0, 60.0}, {0.0, 5.0, 5.0}));
}

// examples/fractal_teapots.rs:8-81: one KdTree of boxed transformed teapots per recursion depth
inline void fractal_gen(std::vector<std::vector<std::shared_ptr<const Shape>>>& groups, const std::shared_ptr<const Shape>& mesh, Vec3 p, double rad,
                        size_t depth, int last_dir) {
    groups[depth].push_back(xf(mesh).scale({0.5, 0.5, 0.5}).scale({rad, rad, rad}).translate(p).build());
    if (depth + 1 == groups.size()) return;
    const double disp = rad * 7.0 / 5.0;
    const double dx[6] = {disp, -disp, 0.0, 0.0, 0.0, 0.0}, dy[6] = {0.0, 0.0, disp, -disp, 0.0, 0.0}, dz[6] = {0.0, 0.0, 0.0, 0.0, disp, -disp};
    for (int i = 0; i < 6; i++)
        if (last_dir < 0 || i != (last_dir ^ 1)) fractal_gen(groups, mesh, Vec3(p.x + dx[i], p.y + dy[i], p.z + dz[i]), rad * 2.0 / 5.0, depth + 1, i);
}
inline void fractal_teapots(Scene& s, std::shared_ptr<const Shape> mesh, std::vector<bool>* object_is_group) {
    const uint32_t colors[5] = {0x264653, 0x2A9D8F, 0xE9C46A, 0xF4A261, 0xE76F51};
    std::vector<std::vector<std::shared_ptr<const Shape>>> groups(5);
    fractal_gen(groups, mesh, Vec3(0.0, 0.0, 0.0), 1.0, 0, -1);
    for (size_t i = 0; i < groups.size(); i++) {
        add(s, std::make_shared<KdTree>(std::move(groups[i]), true), phong(hex_color(colors[i]), 0.25));
        object_is_group->push_back(true);
    }
    add(s, std::make_shared<Plane>(Vec3(0.0, 0.0, 1.0), -6.0), diffuse(hex_color(0xFFCCCC)));
    object_is_group->push_back(false);
    s.lights.push_back(ambient({0.02, 0.02, 0.02}));
    s.lights.push_back(directional({0.6, 0.6, 0.6}, normalize(Vec3(0.0, -0.65, -1.0))));
    s.lights.push_back(point({100.0, 100.0, 100.0}, {0.0, 5.0, 5.0}));
}

}  // namespace builtin
}  // namespace orc

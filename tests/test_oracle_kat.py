"""CPU: pin the ORACLE against known answers derived from the reference's own formulas
(SURVEY.md section 8c).  The reference holds no golden vectors for this path (its three unit
tests are src/color.rs:31-38, monomial_surface.rs:194, ode/particle_system.rs:136), so the one
reusable reference test (colors_work) plus analytic cases are what can be asserted."""
import ctypes as C
import math

import numpy as np

import orc
from rpt_b200 import Camera, Light, Material, Medium, Object, Scene, cube, hex_color, monomial_surface, polygon, scenes, sphere, plane


def _scene(*nodes):
    s = Scene()
    for n in nodes:
        s.add(n)
    return s


def test_colors_work_reference_unit_test():
    """src/color.rs:31-38: hex_color -> color_bytes round-trips the 8-bit channels."""
    L = orc.lib()
    for x, want in [(0xFF0000, (255, 0, 0)), (0x00FF00, (0, 255, 0)), (0x0000FF, (0, 0, 255)), (0xD2B48C, (0xD2, 0xB4, 0x8C)),
                    (0x123456, (0x12, 0x34, 0x56))]:
        c = (C.c_double * 3)()
        L.orc_hex_color(x, c)
        b = (C.c_uint8 * 3)()
        L.orc_color_bytes(c, b)
        # truncating encode (src/color.rs:18-24) may land one below after the pow round trip
        assert all(abs(int(b[k]) - want[k]) <= 1 for k in range(3)), (hex(x), tuple(b))
        assert np.allclose(list(c), hex_color(x), rtol=1e-15)


def test_unit_sphere_and_cube_hits():
    """sphere.rs:14-46: from (0,0,10) toward -z => t=9, n=(0,0,1); cube.rs:22-74 => t=9.5."""
    osc = orc.OracleScene(_scene(Object(sphere())))
    t, obj, prim, inst, n = osc.intersect_batch([[0, 0, 10], [0, 0, 10], [0, 0, 0]], [[0, 0, -1], [0, 1, 0], [0, 0, 1]])
    assert t[0] == 9.0 and obj[0] == 0 and np.array_equal(n[0], [0, 0, 1])
    assert np.isinf(t[1]) and obj[1] == 0xFFFFFFFF
    assert t[2] == 1.0  # from inside: t_plus
    osc = orc.OracleScene(_scene(Object(cube())))
    t, obj, _, _, n = osc.intersect_batch([[0, 0, 10], [0.2, 0.1, 0]], [[0, 0, -1], [0, 0, 1]])
    assert t[0] == 9.5 and np.array_equal(n[0], [0, 0, 1])
    assert t[1] == 0.5  # exit face from inside


def test_transformed_keeps_world_scale_t():
    """shape.rs:129-138: the local ray is not renormalised, so t stays in world units."""
    osc = orc.OracleScene(_scene(Object(sphere().scale((2, 2, 2)).translate((0, 0, -5)))))
    t, obj, _, _, n = osc.intersect_batch([[0, 0, 10]], [[0, 0, -1]])
    assert abs(t[0] - 13.0) < 1e-12 and np.allclose(n[0], [0, 0, 1])


def test_plane_and_polygon_normals():
    """plane.rs:17-32 flips the normal toward the ray; the Cornell light faces -y."""
    osc = orc.OracleScene(_scene(Object(plane((0, 1, 0), -1.0))))
    t, _, _, _, n = osc.intersect_batch([[0, 0, 0], [0, -3, 0]], [[0, -1, 0], [0, 1, 0]])
    assert t[0] == 1.0 and np.array_equal(n[0], [0, 1, 0])
    assert t[1] == 2.0 and np.array_equal(n[1], [0, -1, 0])
    scene, cam, d = scenes.cornell()
    osc = orc.OracleScene(scene)
    t, obj, prim, _, n = osc.intersect_batch([[278, 300, 279.5]], [[0, 1, 0]])
    assert np.allclose(n[0], [0, -1, 0]) and abs(t[0] - 248.8) < 1e-9  # light polygon at y = 548.8 (examples/cornell.rs:42-47)


def test_first_object_wins_ties():
    """renderer.rs:419-423: strict < over the ordered object loop."""
    a, b = Object(sphere()), Object(sphere())
    osc = orc.OracleScene(_scene(a, b))
    _, obj, *_ = osc.intersect_batch([[0, 0, 10]], [[0, 0, -1]])
    assert obj[0] == 0


def test_rng_stream_is_the_documented_splitmix():
    L = orc.lib()
    G = 0x9E3779B97F4A7C15
    M = (1 << 64) - 1

    def mix(z):
        z &= M
        z ^= z >> 30; z = z * 0xBF58476D1CE4E5B9 & M
        z ^= z >> 27; z = z * 0x94D049BB133111EB & M
        return z ^ (z >> 31)

    for seed, dom, a, b in [(0, 0, 0, 0), (7, 1, 123456, 9), (2**63, 2, 5, 2**40)]:
        s = mix(seed + (dom + 1) * G)
        s = mix(s + a)
        st = mix(s ^ b)
        for skip in range(3):
            st = (st + G) & M
            assert L.orc_rng_u64(seed, dom, a, b, skip) == mix(st)


def test_direct_light_of_a_facing_patch_matches_the_closed_form():
    """One diffuse floor + a small square light above it, max_bounces 0: the pixel looking at the
    point under the light converges to albedo/pi * E * area * cos*cos / d^2 (light.rs:34-45,
    renderer.rs:362-409) within Monte Carlo error; the light itself shows E*colour."""
    s = Scene()
    white = Material.diffuse((0.8, 0.8, 0.8))
    s.add(Object(polygon([(-50, 0, -50), (-50, 0, 50), (50, 0, 50), (50, 0, -50)])).material(white))
    lm = Material.light((1.0, 1.0, 1.0), 10.0)
    half = 0.5
    s.add((polygon([(-half, 10, -half), (half, 10, -half), (half, 10, half), (-half, 10, half)]), lm))
    cam = Camera.look_at((0, 5, 0.001), (0, 0, 0), (0, 0, -1), math.pi / 200)
    r = scenes.make_renderer(s, cam, dict(width=8, height=8, max_bounces=0, num_samples=1), seed=5)
    osc = orc.OracleScene(s)
    of = orc.OracleFilm(osc, cam, r.params())
    of.sample(2000)
    got = of.mean()[3:5, 3:5].mean()
    want = 0.8 / math.pi * 10.0 * 1.0 / 100.0  # cos = 1 both ends, d = 10, area = 1
    assert abs(got - want) / want < 0.02, (got, want)


def test_film_encode_truncates_and_box_filters():
    """buffer.rs:43-56,75-94 + color.rs:18-24 on a constant-environment render (miss everywhere)."""
    s = Scene()
    s.environment = (0.25, 0.5, 1.5)
    s.add(Object(sphere().translate((0, 0, 1000))))
    cam = Camera()
    r = scenes.make_renderer(s, cam, dict(width=6, height=4, max_bounces=0, num_samples=1))
    of = orc.OracleFilm(orc.OracleScene(s), cam, r.params())
    of.sample(1)
    img = of.image()
    want = [int((min(max(c, 0.0), 1.0) ** (1 / 2.2)) * 255.0) for c in (0.25, 0.5, 1.5)]
    assert (img.reshape(-1, 3) == want).all()
    assert math.isnan(of.variance())  # one batch: 0/0 (buffer.rs:59-73)
    of.sample(1)
    assert of.variance() == 0.0


def test_medium_free_flight_statistics():
    """medium.rs:133-146: d = -ln(U)/sigma_t.  With an empty black scene the medium path tracer
    returns 0 radiance (no emission, no lights) but must terminate: RR 0.8 (renderer.rs:193)."""
    s = Scene()
    s.add(Object(sphere().translate((0, 0, 5000))))
    s.add(Medium.homogeneous_isotropic(0.001, 0.004))
    cam = Camera()
    r = scenes.make_renderer(s, cam, dict(width=8, height=8, max_bounces=3, num_samples=1))
    of = orc.OracleFilm(orc.OracleScene(s), cam, r.params())
    of.sample(50)
    assert np.array_equal(of.mean(), np.zeros((8, 8, 3)))
    c = of.counters()
    # segments per path: background_dist 400 (renderer.rs:199), P(scatter before 400) = 1-exp(-2) = .8647;
    # continue w.p. 0.8 => mean segments = 1/(1 - .8647*.8) = 3.243
    per_path = c["extend_rays"] / (8 * 8 * 50)
    assert abs(per_path - 1.0 / (1.0 - (1 - math.exp(-2.0)) * 0.8)) < 0.15, per_path


def test_monomial_surface_known_answers():
    """src/shape/monomial_surface.rs:22-107 on y = 2 (x^2+z^2)^2: the axis ray from above meets the apex at
    t = 3 with normal +y (two-sided flip); a horizontal ray at height y0 enters at radius (y0/2)^(1/4); rays
    outside the unit disc miss (second equation, :88-91); the surface satisfies its own equation at every hit."""
    osc = orc.OracleScene(_scene(Object(monomial_surface(2.0, 4.0))))
    y0 = 0.5
    r0 = (y0 / 2.0) ** 0.25
    o = [[0, 3, 0], [-3, y0, 0], [0, 3, 1.5], [0.3, 3, -0.2], [0, -1, 0]]
    d = [[0, -1, 0], [1, 0, 0], [0, -1, 0], [0, -1, 0], [0, 1, 0]]
    t, obj, prim, inst, n = osc.intersect_batch(o, d)
    assert abs(t[0] - 3.0) < 1e-9 and np.allclose(n[0], [0, 1, 0], atol=1e-12)
    assert abs(t[1] - (3.0 - r0)) < 1e-9 and n[1][0] < 0 and abs(np.linalg.norm(n[1]) - 1) < 1e-12
    assert np.isinf(t[2]) and obj[2] == 0xFFFFFFFF
    s = 0.3 ** 2 + 0.2 ** 2
    assert abs(t[3] - (3.0 - 2.0 * s * s)) < 1e-9
    assert abs(t[4] - 1.0) < 1e-9 and np.allclose(n[4], [0, -1, 0], atol=1e-12)
    rng = np.random.default_rng(3)
    o = np.stack([rng.uniform(-2, 2, 4000), rng.uniform(2.5, 4, 4000), rng.uniform(-2, 2, 4000)], axis=1)
    dd = np.stack([rng.uniform(-0.5, 0.5, 4000), -np.ones(4000), rng.uniform(-0.5, 0.5, 4000)], axis=1)
    t, obj, *_ = osc.intersect_batch(o, dd)
    hit = obj == 0
    assert 0.05 < hit.mean() < 0.9
    p = o[hit] + t[hit, None] * dd[hit]
    s = p[:, 0] ** 2 + p[:, 2] ** 2
    assert (s <= 1.0).all() and np.abs(p[:, 1] - 2.0 * s * s).max() < 1e-9


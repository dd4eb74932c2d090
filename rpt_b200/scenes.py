"""The BASELINE.json workloads, transcribed through the mirror API from the reference's
examples (scene content only; sizes/spp come from BASELINE.json):
  cornell()            examples/cornell.rs:15-83
  teapot()             examples/teapot.rs:10-27
  fractal_teapots()    examples/fractal_teapots.rs:8-81
  lampshade(variant)   examples/volumetric_{pathtrace,photonphoton,beamphoton,beambeam}_lampshade.rs
Each returns (scene, camera, renderer_defaults)."""
import math
import os

import numpy as np

from .scene import KdTree, Mesh, cube, hex_color, plane, polygon, sphere
from .world import Camera, Light, Material, Medium, Object, Scene

_TEAPOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets", "teapot_tris.npz")


def teapot_mesh(path=None):
    """The examples' teapot.obj as a Mesh.  Read from the packaged triangle asset rpt_b200/assets/teapot_tris.npz (derived
    from the reference asset by tests/golden/make_fixtures.py) or from an .obj given by path /
    $RPTX_TEAPOT_OBJ."""
    path = path or os.environ.get("RPTX_TEAPOT_OBJ")
    if path:
        from .io import load_obj

        return load_obj(path)
    z = np.load(_TEAPOT)
    return Mesh(vertices=z["vertices"], normals=z["normals"])


def _cornell_walls(scene, white, red, green):
    floor = polygon([(0.0, 0.0, 0.0), (0.0, 0.0, 559.2), (556.0, 0.0, 559.2), (556.0, 0.0, 0.0)])
    ceiling = polygon([(0.0, 548.9, 0.0), (556.0, 548.9, 0.0), (556.0, 548.9, 559.2), (0.0, 548.9, 559.2)])
    back_wall = polygon([(0.0, 0.0, 559.2), (0.0, 548.9, 559.2), (556.0, 548.9, 559.2), (556.0, 0.0, 559.2)])
    right_wall = polygon([(0.0, 0.0, 0.0), (0.0, 548.9, 0.0), (0.0, 548.9, 559.2), (0.0, 0.0, 559.2)])
    left_wall = polygon([(556.0, 0.0, 0.0), (556.0, 0.0, 559.2), (556.0, 548.9, 559.2), (556.0, 548.9, 0.0)])
    scene.add(Object(floor).material(white))
    scene.add(Object(ceiling).material(white))
    scene.add(Object(back_wall).material(white))
    scene.add(Object(left_wall).material(red))
    scene.add(Object(right_wall).material(green))


def _cornell_camera():
    return Camera(eye=(278.0, 273.0, -800.0), direction=(0.0, 0.0, 1.0), up=(0.0, 1.0, 0.0), fov=0.686)


def cornell():
    """C1: examples/cornell.rs (the example's `small_box` really is a sphere)."""
    scene = Scene()
    white = Material.diffuse(hex_color(0xAAAAAA))
    red = Material.diffuse(hex_color(0xBC0000))
    green = Material.diffuse(hex_color(0x00BC00))
    light_mtl = Material.light(hex_color(0xFFFEFA), 100.0)
    light_rect = polygon([(343.0, 548.8, 227.0), (343.0, 548.8, 332.0), (213.0, 548.8, 332.0), (213.0, 548.8, 227.0)])
    large_box = cube().scale((165.0, 330.0, 165.0)).rotate_y(2.0 * math.pi * (-253.0 / 360.0)).translate((368.0, 165.0, 351.0))
    small_box = sphere().scale((80.0, 80.0, 80.0)).rotate_y(2.0 * math.pi * (-197.0 / 360.0)).translate((150.0, 82.5, 450.0))
    _cornell_walls(scene, white, red, green)
    scene.add(Object(large_box).material(white))
    scene.add(Object(small_box).material(white))
    scene.add((light_rect, light_mtl))
    return scene, _cornell_camera(), dict(width=512, height=512, filter_radius=1, max_bounces=5, num_samples=16)


def teapot(mesh=None):
    """C2a: examples/teapot.rs"""
    mesh = mesh or teapot_mesh()
    scene = Scene()
    scene.add(Object(mesh.scale((0.5, 0.5, 0.5)).translate((0.0, -1.0, 0.0))).material(Material.metallic(hex_color(0xFF0000), 0.4)))
    scene.add(Object(plane((0.0, 1.0, 0.0), -1.0)).material(Material.diffuse(hex_color(0xAAAAAA))))
    scene.add(Light.Ambient((0.02, 0.02, 0.02)))
    scene.add(Light.Point((60.0, 60.0, 60.0), (0.0, 5.0, 5.0)))
    return scene, Camera(), dict(width=1920, height=1080, max_bounces=0, num_samples=256)


def fractal_teapots(mesh=None, levels=5):
    """C2b: examples/fractal_teapots.rs — kd-tree of kd-trees: 1+6+30+150+750 instances."""
    mesh = mesh or teapot_mesh()
    colors = [0x264653, 0x2A9D8F, 0xE9C46A, 0xF4A261, 0xE76F51][:levels]
    groups = [[] for _ in colors]

    def gen(p, rad, depth, last_dir):
        groups[depth].append(mesh.scale((0.5, 0.5, 0.5)).scale((rad, rad, rad)).translate(p))
        if depth == len(groups) - 1:
            return
        disp = rad * 7.0 / 5.0
        dx = [disp, -disp, 0.0, 0.0, 0.0, 0.0]
        dy = [0.0, 0.0, disp, -disp, 0.0, 0.0]
        dz = [0.0, 0.0, 0.0, 0.0, disp, -disp]
        for i in range(6):
            if last_dir is None or i != (last_dir ^ 1):
                gen((p[0] + dx[i], p[1] + dy[i], p[2] + dz[i]), rad * 2.0 / 5.0, depth + 1, i)

    gen((0.0, 0.0, 0.0), 1.0, 0, None)
    scene = Scene()
    for i, g in enumerate(groups):
        scene.add(Object(KdTree(g)).material(Material.specular(hex_color(colors[i]), 0.25)))
    scene.add(Object(plane((0.0, 0.0, 1.0), -6.0)).material(Material.diffuse(hex_color(0xFFCCCC))))
    scene.add(Light.Ambient((0.02, 0.02, 0.02)))
    n = math.sqrt(0.65 * 0.65 + 1.0)
    scene.add(Light.Directional((0.6, 0.6, 0.6), (0.0 / n, -0.65 / n, -1.0 / n)))
    scene.add(Light.Point((100.0, 100.0, 100.0), (0.0, 5.0, 5.0)))
    d = (-0.285714, -0.5, -1.0)
    dn = math.sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2])
    u = (0.0, 1.0, -0.5)
    un = math.sqrt(u[1] * u[1] + u[2] * u[2])
    camera = Camera(eye=(2.0, 3.5, 7.0), direction=(d[0] / dn, d[1] / dn, d[2] / dn), up=(u[0] / un, u[1] / un, u[2] / un), fov=math.pi / 6.0)
    return scene, camera, dict(width=1920, height=1080, max_bounces=0, num_samples=256)


def lampshade(variant="pathtrace"):
    """C3-C5: the volumetric lampshade Cornell box.  variant in
    {pathtrace, photonphoton, beamphoton, beambeam}."""
    cfg = {
        "pathtrace": dict(absorb=0.00005, scat=0.003, emit=150.0),
        "photonphoton": dict(absorb=0.0008, scat=0.0008, emit=120.0),
        "beamphoton": dict(absorb=0.0001, scat=0.001, emit=200000.0 / (130.0 * 105.0)),
        "beambeam": dict(absorb=0.0001, scat=0.001, emit=200000.0 / (130.0 * 105.0)),
    }[variant]
    scene = Scene()
    white = Material.diffuse(hex_color(0xAAAAAA))
    red = Material.diffuse(hex_color(0xBC0000))
    yellow = Material.diffuse(hex_color(0xBCBC00))
    green = Material.diffuse(hex_color(0x00BC00))
    light_rect = polygon([(330.0, 548.8, 240.0), (330.0, 548.8, 319.0), (226.0, 548.8, 319.0), (226.0, 548.8, 240.0)])
    height, depth, width = 140.0, 105.0, 130.0
    center = (213.0 + 65.0, 548.0, 227.0 + 55.0)
    front_offset = (center[0], center[1], center[2] + depth / 2.0)
    left_offset = (center[0] - width / 2.0, center[1], center[2])
    back_offset = (center[0], center[1], center[2] - depth / 2.0)
    right_offset = (center[0] + width / 2.0, center[1], center[2])
    off = 10.0
    front_shade = cube().scale((130.0 + off * 2.0, height, off)).translate(front_offset)
    left_shade = cube().scale((off, height, 105.0 + off * 2.0)).translate(left_offset)
    back_shade = cube().scale((130.0 + off * 2.0, height, off)).translate(back_offset)
    right_shade = cube().scale((off, height, 105.0 + off * 2.0)).translate(right_offset)
    large_box = cube().scale((165.0, 330.0, 165.0)).rotate_y(2.0 * math.pi * (-253.0 / 360.0)).translate((368.0, 165.0, 351.0))
    small_box = cube().scale((165.0, 165.0, 165.0)).rotate_y(2.0 * math.pi * (-197.0 / 360.0)).translate((185.0, 82.5, 169.0))
    _cornell_walls(scene, white, red, green)
    scene.add(Object(large_box).material(white))
    scene.add(Object(small_box).material(white))
    scene.add(Object(right_shade).material(yellow))
    scene.add(Object(left_shade).material(yellow))
    scene.add(Object(front_shade).material(yellow))
    scene.add(Object(back_shade).material(yellow))
    scene.add((light_rect, Material.light(hex_color(0xFFFEFA), cfg["emit"])))
    scene.add(Medium.homogeneous_isotropic(cfg["absorb"], cfg["scat"]))
    defaults = {
        "pathtrace": dict(width=1920, height=1080, max_bounces=10, num_samples=1024),
        "photonphoton": dict(width=128, height=128, max_bounces=10, num_samples=100, gather_size=100, gather_size_volume=30, watts=1e7),
        "beamphoton": dict(width=128, height=128, max_bounces=10, num_samples=50, gather_size=20, gather_size_volume=3, watts=cfg["emit"]),
        "beambeam": dict(width=128, height=128, max_bounces=10, num_samples=50, gather_size=20, gather_size_volume=3, watts=cfg["emit"]),
    }[variant]
    return scene, _cornell_camera(), defaults


def monomial_glass(environment=(0.6, 0.7, 0.9)):
    """examples/monomial_glass.rs:29-82.  The example downloads its HDRI; a solid environment colour (or an
    Hdri passed in) stands in for it."""
    from .scene import monomial_surface

    scene = Scene()
    scene.environment = environment
    scene.add(Object(monomial_surface(2.0, 4.0).translate((0.0, -1.0, 0.0))).material(Material.metallic(hex_color(0xFFFFFF), 0.0001)))
    scene.add(Object(cube().rotate_y(math.pi / 6.0).scale((0.5, 0.3, 0.4)).translate((0.4, -0.8, 4.0))).material(Material.specular(hex_color(0xFF00FF), 0.5)))
    scene.add(Object(sphere().scale((0.5, 0.5, 0.5)).translate((1.5, -0.5, 1.0))).material(Material.specular(hex_color(0x0000FF), 0.1)))
    scene.add(Object(sphere().scale((0.5, 0.5, 0.5)).translate((-1.5, -0.5, 1.0))).material(Material.specular(hex_color(0x00FF00), 0.1)))
    scene.add(Object(plane((0.0, 1.0, 0.0), -1.0)).material(Material.specular(hex_color(0xAAAAAA), 0.5)))
    scene.add(Light.Ambient((0.01, 0.01, 0.01)))
    scene.add(Light.Point((100.0, 100.0, 100.0), (0.0, 5.0, 5.0)))
    return scene, Camera(), dict(width=800, height=600, max_bounces=1, num_samples=100)


def make_renderer(scene, camera, defaults, **overrides):
    """Renderer builder preloaded with a workload's parameters."""
    from .renderer import Renderer

    r = Renderer(scene, camera)
    p = dict(defaults)
    p.update(overrides)
    for k, v in p.items():
        r._set(k, v)
    return r

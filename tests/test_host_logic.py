"""CPU: the host-side mirror of rpt's API (no device needed): builder defaults, scene lowering,
SceneAdd semantics, transform composition, the OBJ loader, the committed teapot fixture."""
import io
import math
import os

import numpy as np
import pytest

from rpt_b200 import Camera, KdTree, Light, Material, Medium, Object, Renderer, Scene, cube, load_obj, plane, polygon, sphere
from rpt_b200 import _abi as A
from rpt_b200 import scenes
from rpt_b200.world import lower_scene

HERE = os.path.dirname(os.path.abspath(__file__))


def test_renderer_builder_defaults_match_the_reference():
    """src/renderer.rs:60-76"""
    p = Renderer.new(Scene(), Camera()).params()
    assert (p.width, p.height, p.max_bounces, p.num_samples) == (800, 600, 0, 1)
    assert (p.exposure_value, p.filter_radius, p.gather_size, p.gather_size_volume, p.watts) == (0.0, 0, 50, 50, 100.0)
    assert (p.shard_count, p.shard_index) == (1, 0)
    p = Renderer(Scene(), Camera()).width(3).height(4).max_bounces(5).num_samples(6).gather_size(7).gather_size_volume(8).watts(9.5).params()
    assert (p.width, p.height, p.max_bounces, p.num_samples, p.gather_size, p.gather_size_volume, p.watts) == (3, 4, 5, 6, 7, 8, 9.5)


def test_camera_defaults_and_look_at():
    """src/camera.rs:29-63"""
    c = Camera()
    assert c.eye == (0.0, 0.0, 10.0) and c.direction == (0.0, 0.0, -1.0) and abs(c.fov - math.pi / 6) < 1e-16
    c = Camera.look_at((0, 0, 5), (0, 0, 0), (0, 1, 1), 0.5)
    assert np.allclose(c.direction, (0, 0, -1)) and np.allclose(c.up, (0, 1, 0))
    c.focus((0, 0, 1), 0.1)
    assert abs(c.focal_distance - 4.0) < 1e-15 and c.aperture == 0.1


def test_scene_add_of_shape_material_tuple_adds_object_and_light():
    """src/scene.rs:57-75"""
    s = Scene()
    s.add((polygon([(0, 0, 0), (1, 0, 0), (1, 0, 1), (0, 0, 1)]), Material.light((1, 1, 1), 5.0)))
    assert len(s.objects) == 1 and len(s.lights) == 1 and s.lights[0].kind == A.LIGHT_OBJECT
    low = lower_scene(s)
    assert len(low.meshes) == 1  # the mesh is shared by the object and the light
    assert low.lights[0].shape.kind == A.SHAPE_MESH and low.lights[0].material.emittance == 5.0


def test_lowering_transforms_groups_and_media():
    s = Scene()
    s.add(Object(sphere().scale((2, 3, 4)).translate((1, 2, 3))))
    s.add(Object(plane((0, 1, 0), -1.0)))
    m = polygon([(0, 0, 0), (1, 0, 0), (0, 1, 0)])
    s.add(Object(KdTree([m.translate((5, 0, 0)), m.translate((6, 0, 0)), cube()])))
    s.add(Medium.homogeneous_isotropic(0.1, 0.2))
    s.add(Medium.colored_glowing_fog(9, 9))  # ignored: only media[0] is used (renderer.rs:190)
    s.add(Light.Point((1, 2, 3), (4, 5, 6)))
    low = lower_scene(s)
    d0 = low.objects[0][0]
    M = np.array(list(d0.transform)).reshape(4, 4)
    # scale first, then translate: M = T * S
    assert d0.has_transform == 1 and np.allclose(M, [[2, 0, 0, 1], [0, 3, 0, 2], [0, 0, 4, 3], [0, 0, 0, 1]])
    assert low.objects[1][0].kind == A.SHAPE_PLANE and low.objects[1][0].plane_value == -1.0
    g = low.objects[2][0]
    assert g.kind == A.SHAPE_GROUP and g.n_children == 3
    assert [g.children[i].kind for i in range(3)] == [A.SHAPE_MESH, A.SHAPE_MESH, A.SHAPE_CUBE]
    assert g.children[0].mesh_id == g.children[1].mesh_id and len(low.meshes) == 1  # instancing keeps one mesh
    assert low.medium == (A.MEDIUM_HOMOGENEOUS_ISOTROPIC, 0.1, 0.2)
    assert list(low.lights[0].vec) == [4, 5, 6]


def test_closed_shape_set_is_enforced():
    with pytest.raises(TypeError):
        Object(object())
    with pytest.raises(TypeError):
        KdTree([plane((0, 1, 0), 0.0)])


def test_material_constructors():
    """src/material.rs:26-98: specular/metallic store `roughness` as the Phong exponent."""
    assert Material.specular((1, 0, 0), 0.25).shininess == 0.25 and Material.specular((1, 0, 0), 0.25).kind == A.MAT_PHONG
    assert Material.mirror().kind == A.MAT_MIRROR and Material.clear(1.5).ior == 1.5
    assert Material.light((1, 1, 1), 7.0).lower().emittance == 7.0


OBJ = """# fan + negative indices + missing normals
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
vn 0 0 1
f 1//1 2//1 3//1 4//1
f -4 -3 -2
"""


def test_load_obj_semantics():
    """src/io.rs:11-19,28-74,152-201"""
    m = load_obj(io.StringIO(OBJ))
    assert len(m) == 3
    assert np.array_equal(m.vertices[0], [[0, 0, 0], [1, 0, 0], [1, 1, 0]])
    assert np.array_equal(m.vertices[1], [[0, 0, 0], [1, 1, 0], [0, 1, 0]])  # fan triangulation
    assert np.array_equal(m.vertices[2], [[0, 0, 0], [1, 0, 0], [1, 1, 0]])  # negative = relative to the end
    assert np.allclose(m.normals[2], [[0, 0, 1]] * 3)  # face normal filled in for the corner-less face


def test_load_stl_binary_and_ascii(tmp_path):
    """src/io.rs:264-364: binary = 80-byte header, u32 count, 50-byte records (f32 normal + 3 vertices + u16), values
    widened from f32, facet normal on all three corners; ASCII = `solid `, then facets until end of file."""
    import struct

    from rpt_b200 import load_stl

    tris = [((0, 0, 1), (0, 0, 0), (1, 0, 0), (0, 1, 0)), ((1, 0, 0), (0.1, 0.2, 0.3), (0, 1, 0), (0, 0, 1))]
    blob = b"\0" * 80 + struct.pack("<I", len(tris))
    for t in tris:
        blob += struct.pack("<12f", *[c for v in t for c in v]) + b"\0\0"
    p = tmp_path / "b.stl"
    p.write_bytes(blob)
    m = load_stl(str(p))
    assert len(m) == 2
    assert np.array_equal(m.vertices[1][0], np.array([0.1, 0.2, 0.3], np.float32).astype(np.float64))
    assert np.array_equal(m.normals[0], [[0, 0, 1]] * 3) and np.array_equal(m.normals[1], [[1, 0, 0]] * 3)
    body = "solid t\n" + "".join(
        f"facet normal {n[0]} {n[1]} {n[2]}\nouter loop\nvertex {a[0]} {a[1]} {a[2]}\nvertex {b[0]} {b[1]} {b[2]}\n"
        f"vertex {c[0]} {c[1]} {c[2]}\nendloop\nendfacet\n" for n, a, b, c in tris)
    m2 = load_stl(io.BytesIO(body.encode()))
    assert len(m2) == 2 and np.array_equal(m2.vertices[1][0], [0.1, 0.2, 0.3]) and np.array_equal(m2.normals[1], [[1, 0, 0]] * 3)
    # the reference's ASCII loop treats EVERY remaining line as a facet (src/io.rs:294-299): `endsolid` is malformed
    with pytest.raises(ValueError, match="facet normal"):
        load_stl(io.BytesIO((body + "endsolid t\n").encode()))
    with pytest.raises(ValueError, match="too short"):
        load_stl(io.BytesIO(b"solid"))
    with pytest.raises(ValueError, match="could not determine format"):
        load_stl(io.BytesIO(b"not an stl file at all"))


def test_load_obj_with_mtl_groups_by_usemtl():
    """src/io.rs:84-150,203-258: default material per `newmtl`; one Object per run of a material; unknown names are
    InvalidData; any MTL property line hits the reference's `panic!("MTL loading not implemented")`."""
    from rpt_b200 import load_mtl, load_obj_with_mtl

    obj = "v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nf 1 2 3\nusemtl a\nf 1 2 3 4\nusemtl a\nf 1 3 4\nusemtl b\nf 2 3 4\n"
    objs = load_obj_with_mtl(io.StringIO(obj), io.StringIO("# lib\nnewmtl a\nnewmtl b\n"))
    assert [len(o.shape) for o in objs] == [1, 3, 1]
    assert all(o.mat.kind == Material.default().kind for o in objs)
    with pytest.raises(ValueError, match="usemtl c"):
        load_obj_with_mtl(io.StringIO(obj.replace("usemtl b", "usemtl c")), io.StringIO("newmtl a\nnewmtl b\n"))
    with pytest.raises(RuntimeError, match="not implemented"):
        load_mtl(io.StringIO("newmtl a\nKd 1 0 0\n"))
    with pytest.raises(ValueError, match="newmtl"):
        load_mtl(io.StringIO("Kd 1 0 0\n"))


def test_monomial_surface_lowering():
    """src/shape.rs:292-295 -> RPTX_SHAPE_MONOMIAL with height / exp in the POD record."""
    from rpt_b200 import monomial_surface

    s, _, _ = scenes.monomial_glass()
    low = lower_scene(s)
    d = low.objects[0][0]
    assert d.kind == A.SHAPE_MONOMIAL and d.mono_height == 2.0 and d.mono_exp == 4.0 and d.has_transform == 1
    assert monomial_surface(1.0, 4.0).kind == A.SHAPE_MONOMIAL


def test_teapot_fixture_is_the_reference_asset():
    z = np.load(os.path.join(HERE, "..", "rpt_b200", "assets", "teapot_tris.npz"))
    assert z["vertices"].shape == (2256, 3, 3) and z["normals"].shape == (2256, 3, 3)
    if os.path.exists("/root/reference/examples/teapot.obj"):  # build container only
        m = load_obj("/root/reference/examples/teapot.obj")
        assert np.array_equal(m.vertices, z["vertices"]) and np.array_equal(m.normals, z["normals"])


def test_benchmark_scenes_have_the_reference_inventory():
    """object / light counts of the BASELINE configs."""
    s, _, d = scenes.cornell()
    assert len(s.objects) == 8 and len(s.lights) == 1 and d["max_bounces"] == 5 and (d["width"], d["num_samples"]) == (512, 16)
    s, _, d = scenes.fractal_teapots()
    low = lower_scene(s)
    assert sum(o[0].n_children for o in low.objects if o[0].kind == A.SHAPE_GROUP) + sum(
        1 for o in low.objects if o[0].kind == A.SHAPE_MESH) == 937
    assert len(low.meshes) == 1 and (d["width"], d["height"], d["num_samples"]) == (1920, 1080, 256)
    s, _, d = scenes.lampshade("pathtrace")
    assert len(s.objects) == 12 and len(s.media) == 1 and (d["width"], d["height"], d["num_samples"]) == (1920, 1080, 1024)


def test_sah_builder_invariants(tmp_path):
    """rpt_b200/csrc/bvh_sah.hpp on the CPU: permutation, containment, leaf size, bounded depth (tests/native/sah_check.cu)."""
    import shutil
    import subprocess

    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    exe = tmp_path / "sah_check"
    src = os.path.join(os.path.dirname(os.path.abspath(__file__)), "native", "sah_check.cu")
    subprocess.check_call([nvcc, "-std=c++17", "-O2", "-Wno-deprecated-gpu-targets", "-x", "cu", "-o", str(exe), src])
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout[-2000:]
    assert "FAIL" not in out.stdout and out.stdout.count("ok ") >= 18

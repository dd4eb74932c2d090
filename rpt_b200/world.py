"""Material / Object / Light / Medium / Camera / Scene mirrors and the lowering of a Scene
into the POD records of include/rptx.h (see scene.py for the shape half)."""
import ctypes as C
import math

import numpy as np

from . import _abi as A
from .scene import Cube, KdTree, Mesh, MonomialSurface, Plane, Shape, Sphere, Transformed, _dot, _normalize, _sub, vec3


class Material:
    """src/material.rs:8-98"""

    def __init__(self, kind, albedo=(0.5, 0.5, 0.5), emittance=0.0, shininess=0.0, ior=1.0):
        self.kind, self.albedo, self.emittance, self.shininess, self.ior = kind, vec3(*albedo), float(emittance), float(shininess), float(ior)

    @staticmethod
    def default():
        return Material(A.MAT_LAMBERTIAN, (0.5, 0.5, 0.5), 0.0)

    @staticmethod
    def diffuse(color):
        return Material(A.MAT_LAMBERTIAN, color, 0.0)

    @staticmethod
    def specular(color, roughness):
        return Material(A.MAT_PHONG, color, 0.0, shininess=roughness)

    @staticmethod
    def metallic(color, roughness):
        return Material(A.MAT_PHONG, color, 0.0, shininess=roughness)

    @staticmethod
    def mirror():
        return Material(A.MAT_MIRROR)

    @staticmethod
    def transmissive(ior):
        return Material(A.MAT_TRANSMISSIVE, (0.0, 0.0, 0.0), ior=ior)

    @staticmethod
    def clear(index, _roughness=0.0):
        return Material(A.MAT_TRANSMISSIVE, (0.0, 0.0, 0.0), ior=index)

    @staticmethod
    def transparent(color, index, _roughness=0.0):
        return Material(A.MAT_TRANSMISSIVE, color, ior=index)

    @staticmethod
    def light(color, emittance):
        return Material(A.MAT_LAMBERTIAN, color, emittance)

    def lower(self):
        m = A.MaterialDesc()
        m.kind = self.kind
        m.albedo[:] = self.albedo
        m.emittance, m.shininess, m.ior = self.emittance, self.shininess, self.ior
        return m


class Object:
    """src/object.rs:10-31"""

    def __init__(self, shape):
        if not isinstance(shape, Shape):
            raise TypeError("Object::new needs a shape the device can lower (sphere, cube, plane, Mesh, KdTree, Transformed)")
        self.shape = shape
        self.mat = Material.default()

    def material(self, material):
        self.mat = material
        return self


class Light:
    """src/light.rs:7-19"""

    def __init__(self, kind, color=(0, 0, 0), vec=(0, 0, 0), obj=None):
        self.kind, self.color, self.vec, self.object = kind, vec3(*color), vec3(*vec), obj

    @staticmethod
    def Point(color, location):
        return Light(A.LIGHT_POINT, color, location)

    @staticmethod
    def Ambient(color):
        return Light(A.LIGHT_AMBIENT, color)

    @staticmethod
    def Directional(color, direction):
        return Light(A.LIGHT_DIRECTIONAL, color, direction)

    @staticmethod
    def Object(obj):
        return Light(A.LIGHT_OBJECT, obj=obj)


class Medium:
    """src/medium.rs:80-121 — the reference's only two constructors."""

    def __init__(self, kind, absorption, scattering):
        self.kind, self.absorption, self.scattering = kind, float(absorption), float(scattering)

    @staticmethod
    def homogeneous_isotropic(absorption, scattering):
        return Medium(A.MEDIUM_HOMOGENEOUS_ISOTROPIC, absorption, scattering)

    @staticmethod
    def colored_glowing_fog(absorption, scattering):
        return Medium(A.MEDIUM_COLORED_GLOWING_FOG, absorption, scattering)


class Camera:
    """src/camera.rs:9-63"""

    def __init__(self, eye=(0.0, 0.0, 10.0), direction=(0.0, 0.0, -1.0), up=(0.0, 1.0, 0.0), fov=math.pi / 6.0, aperture=0.0, focal_distance=0.0):
        self.eye, self.direction, self.up = vec3(*eye), vec3(*direction), vec3(*up)
        self.fov, self.aperture, self.focal_distance = float(fov), float(aperture), float(focal_distance)

    @staticmethod
    def look_at(eye, center, up, fov):
        direction = _normalize(_sub(center, eye))
        k = _dot(up, direction)
        up2 = _normalize(_sub(up, (k * direction[0], k * direction[1], k * direction[2])))
        return Camera(eye, direction, up2, fov)

    def focus(self, focal_point, aperture):
        self.focal_distance = _dot(_sub(focal_point, self.eye), self.direction)
        self.aperture = float(aperture)
        return self

    def lower(self):
        c = A.CameraDesc()
        c.eye[:], c.direction[:], c.up[:] = self.eye, self.direction, self.up
        c.fov, c.aperture, c.focal_distance = self.fov, self.aperture, self.focal_distance
        return c


class Hdri:
    """src/environment.rs:5-23"""

    def __init__(self, width, height, buf):
        self.width, self.height = int(width), int(height)
        self.buf = np.ascontiguousarray(buf, dtype=np.float64).reshape(self.height, self.width, 3)


class Scene:
    """src/scene.rs:12-81.  `add` dispatches on the argument like the SceneAdd<T> impls."""

    def __init__(self):
        self.objects, self.lights, self.media = [], [], []
        self.environment = (0.0, 0.0, 0.0)  # Environment::Color default (src/environment.rs:64-68)

    @staticmethod
    def new():
        return Scene()

    def add(self, node):
        if isinstance(node, Object):
            self.objects.append(node)
        elif isinstance(node, Light):
            self.lights.append(node)
        elif isinstance(node, Medium):
            self.media.append(node)
        elif isinstance(node, tuple) and len(node) == 2 and isinstance(node[1], Material):
            # SceneAdd<(Mesh, Material)> / <(Transformed<Cube>, Material)>: object AND light
            shape, material = node
            self.add(Object(shape).material(material))
            self.add(Light.Object(Object(shape).material(material)))
        else:
            raise TypeError(f"Scene::add: unsupported node {type(node)}")
        return self


# ------------------------------------------------------------------ lowering
class Lowered:
    """The POD form of a Scene: mesh arrays + ctypes records (kept alive here)."""

    def __init__(self):
        self.meshes = []  # (vertices [n,3,3] f64, normals or None)
        self.mesh_ids = {}
        self.objects = []  # (ShapeDesc, MaterialDesc)
        self.lights = []  # LightDesc
        self.medium = (A.MEDIUM_NONE, 0.0, 0.0)
        self.environment = None
        self._keep = []


def _lower_shape(low, shape, allow_group=True):
    d = A.ShapeDesc()
    base = shape
    if isinstance(shape, Transformed):
        d.has_transform = 1
        flat = [x for row in shape.matrix for x in row]
        d.transform[:] = flat
        base = shape.shape
    if isinstance(base, Sphere):
        d.kind = A.SHAPE_SPHERE
    elif isinstance(base, Cube):
        d.kind = A.SHAPE_CUBE
    elif isinstance(base, MonomialSurface):
        d.kind = A.SHAPE_MONOMIAL
        d.mono_height, d.mono_exp = base.height, base.exp
    elif isinstance(base, Plane):
        d.kind = A.SHAPE_PLANE
        d.plane_normal[:] = base.normal
        d.plane_value = base.value
    elif isinstance(base, Mesh):
        d.kind = A.SHAPE_MESH
        key = id(base)
        if key not in low.mesh_ids:
            low.mesh_ids[key] = len(low.meshes)
            low.meshes.append((base.vertices, base.normals))
            low._keep.append(base)
        d.mesh_id = low.mesh_ids[key]
    elif isinstance(base, KdTree):
        if not allow_group:
            raise TypeError("nested KdTree groups are not lowered")
        d.kind = A.SHAPE_GROUP
        kids = (A.ShapeDesc * len(base.objects))()
        for i, c in enumerate(base.objects):
            kids[i] = _lower_shape(low, c, allow_group=False)
        low._keep.append(kids)
        d.children = C.cast(kids, C.POINTER(A.ShapeDesc))
        d.n_children = len(base.objects)
    else:
        raise TypeError(f"shape {type(base)} cannot be lowered to the device (closed shape set)")
    return d


def lower_scene(scene):
    low = Lowered()
    for o in scene.objects:
        low.objects.append((_lower_shape(low, o.shape), o.mat.lower()))
    for l in scene.lights:
        ld = A.LightDesc()
        ld.kind = l.kind
        ld.color[:] = l.color
        ld.vec[:] = l.vec
        if l.kind == A.LIGHT_OBJECT:
            ld.shape = _lower_shape(low, l.object.shape, allow_group=False)
            ld.material = l.object.mat.lower()
        low.lights.append(ld)
    if scene.media:
        m = scene.media[0]  # only media[0] is used (src/renderer.rs:190)
        low.medium = (m.kind, m.absorption, m.scattering)
    low.environment = scene.environment
    return low


def upload_scene(low, backend):
    """Drive a backend exposing the rptx scene calls (the CUDA library, or in tests the oracle)."""
    for k, (v, n) in enumerate(low.meshes):
        mid = backend.add_mesh(v, n)
        if mid != k:
            raise RuntimeError("backend returned an unexpected mesh id")
    for sd, md in low.objects:
        backend.add_object(sd, md)
    for ld in low.lights:
        backend.add_light(ld)
    backend.set_medium(*low.medium)
    env = low.environment
    if isinstance(env, Hdri):
        backend.set_environment_hdri(env.width, env.height, env.buf)
    else:
        backend.set_environment_color(env)
    backend.commit()

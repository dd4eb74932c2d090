"""Host-side mirror of rpt's scene description API (pure data, no compute).

Mirrors, with the same names and argument meaning:
  shapes       sphere/cube/plane/polygon/Mesh + Transformable (src/shape.rs:180-314)
  KdTree       group of bounded shapes (src/kdtree.rs:103-122)
  Material     src/material.rs:8-98        Object   src/object.rs:10-31
  Light        src/light.rs:7-19           Medium   src/medium.rs:80-121
  Camera       src/camera.rs:9-63          Scene    src/scene.rs:12-81
  hex_color    src/color.rs:10-15          load_obj src/io.rs:28-74,152-201
Vectors are 3-sequences of float; matrices are 4x4 nested lists (row major).
"""
import math

from . import _abi as A


def vec3(x, y, z):
    return (float(x), float(y), float(z))


def _sub(a, b):
    return (a[0] - b[0], a[1] - b[1], a[2] - b[2])


def _dot(a, b):
    return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]


def _cross(a, b):
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])


def _normalize(a):
    n = math.sqrt(_dot(a, a))
    return (a[0] / n, a[1] / n, a[2] / n)


def hex_color(x):
    """src/color.rs:10-15"""
    r = ((x >> 16) & 0xFF) / 255.0
    g = ((x >> 8) & 0xFF) / 255.0
    b = (x & 0xFF) / 255.0
    return (r**2.2, g**2.2, b**2.2)


# ---- 4x4 helpers (glm::translate / scale / rotate on the identity) ----
def _identity():
    return [[1.0 if i == j else 0.0 for j in range(4)] for i in range(4)]


def _matmul(a, b):
    return [[sum(a[i][k] * b[k][j] for k in range(4)) for j in range(4)] for i in range(4)]


def _translation(v):
    m = _identity()
    m[0][3], m[1][3], m[2][3] = float(v[0]), float(v[1]), float(v[2])
    return m


def _scaling(v):
    m = _identity()
    m[0][0], m[1][1], m[2][2] = float(v[0]), float(v[1]), float(v[2])
    return m


def _rotation(angle, axis):
    """nalgebra Rotation3::from_axis_angle(normalize(axis), angle) as a homogeneous matrix."""
    ux, uy, uz = _normalize(axis)
    s, c = math.sin(angle), math.cos(angle)
    omc = 1.0 - c
    m = _identity()
    m[0][0] = ux * ux + (1.0 - ux * ux) * c
    m[0][1] = ux * uy * omc - uz * s
    m[0][2] = ux * uz * omc + uy * s
    m[1][0] = ux * uy * omc + uz * s
    m[1][1] = uy * uy + (1.0 - uy * uy) * c
    m[1][2] = uy * uz * omc - ux * s
    m[2][0] = ux * uz * omc - uy * s
    m[2][1] = uy * uz * omc + ux * s
    m[2][2] = uz * uz + (1.0 - uz * uz) * c
    return m


class Shape:
    """Base of the closed shape set the device understands (trait Shape, src/shape.rs:19-26)."""

    kind = None
    bounded = True

    # Transformable<T> (src/shape.rs:180-236): wraps self in Transformed
    def transform(self, m):
        return Transformed(self, [list(map(float, row)) for row in m])

    def translate(self, v):
        return self.transform(_translation(v))

    def scale(self, v):
        return self.transform(_scaling(v))

    def rotate(self, angle, axis):
        return self.transform(_rotation(angle, axis))

    def rotate_x(self, angle):
        return self.rotate(angle, (1.0, 0.0, 0.0))

    def rotate_y(self, angle):
        return self.rotate(angle, (0.0, 1.0, 0.0))

    def rotate_z(self, angle):
        return self.rotate(angle, (0.0, 0.0, 1.0))


class Transformed(Shape):
    """src/shape.rs:102-178; chained transforms compose into one matrix (src/shape.rs:238-289)."""

    def __init__(self, shape, matrix):
        if isinstance(shape, Transformed):
            raise TypeError("use the chaining methods; Transformed<Transformed<T>> is not lowered")
        self.shape = shape
        self.matrix = matrix
        self.bounded = shape.bounded

    def transform(self, m):
        return Transformed(self.shape, _matmul([list(map(float, row)) for row in m], self.matrix))


class Sphere(Shape):
    kind = A.SHAPE_SPHERE


class Cube(Shape):
    kind = A.SHAPE_CUBE


class MonomialSurface(Shape):
    """src/shape/monomial_surface.rs:8-19: y = height * sqrt(x^2 + z^2)^exp over the unit disc (exp must be 4)."""

    kind = A.SHAPE_MONOMIAL

    def __init__(self, height, exp):
        self.height, self.exp = float(height), float(exp)


class Plane(Shape):
    kind = A.SHAPE_PLANE
    bounded = False

    def __init__(self, normal, value):
        self.normal = vec3(*normal)
        self.value = float(value)


class Triangle:
    """src/shape/mesh.rs:9-38"""

    def __init__(self, v1, v2, v3, n1=None, n2=None, n3=None):
        self.v = (vec3(*v1), vec3(*v2), vec3(*v3))
        self.n = None if n1 is None else (vec3(*n1), vec3(*n2), vec3(*n3))

    @staticmethod
    def from_vertices(v1, v2, v3):
        return Triangle(v1, v2, v3)


class Mesh(Shape):
    """Mesh = KdTree<Triangle> (src/shape/mesh.rs:103).  Triangles without normals get face
    normals on the device (Triangle::from_vertices)."""

    kind = A.SHAPE_MESH

    def __init__(self, triangles=None, vertices=None, normals=None):
        import numpy as np

        if vertices is not None:
            self.vertices = np.ascontiguousarray(vertices, dtype=np.float64).reshape(-1, 3, 3)
            self.normals = None if normals is None else np.ascontiguousarray(normals, dtype=np.float64).reshape(-1, 3, 3)
        else:
            tris = list(triangles)
            self.vertices = np.array([t.v for t in tris], dtype=np.float64).reshape(-1, 3, 3)
            if all(t.n is not None for t in tris) and tris:
                self.normals = np.array([t.n for t in tris], dtype=np.float64).reshape(-1, 3, 3)
            elif any(t.n is not None for t in tris):
                # mixed: resolve face normals on the host with the reference's formula
                ns = []
                for t in tris:
                    if t.n is not None:
                        ns.append(t.n)
                    else:
                        n = _normalize(_cross(_sub(t.v[1], t.v[0]), _sub(t.v[2], t.v[0])))
                        ns.append((n, n, n))
                self.normals = np.array(ns, dtype=np.float64).reshape(-1, 3, 3)
            else:
                self.normals = None

    def __len__(self):
        return len(self.vertices)


class KdTree(Shape):
    """KdTree<Box<dyn Bounded>> (src/kdtree.rs:103-122): a group of bounded shapes."""

    kind = A.SHAPE_GROUP

    def __init__(self, objects):
        self.objects = list(objects)
        for o in self.objects:
            if not o.bounded:
                raise TypeError("KdTree children must be Bounded")


def sphere():
    return Sphere()


def cube():
    return Cube()


def monomial_surface(height, exp):
    """src/shape.rs:292-295"""
    return MonomialSurface(height, exp)


def plane(normal, value):
    return Plane(normal, value)


def polygon(verts):
    """src/shape.rs:306-313"""
    return Mesh([Triangle.from_vertices(verts[0], verts[i], verts[i + 1]) for i in range(1, len(verts) - 1)])

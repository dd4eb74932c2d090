"""Host mesh IO mirroring src/io.rs: load_obj (:28-74,152-201: 1-based / negative indices, fan
triangulation, per-vertex normals when every corner has one, else face normals), load_obj_with_mtl
(:84-150: one Object per `usemtl` run), load_mtl (:203-258) and load_stl (:264-364, binary and ASCII),
with the reference's error behaviour (io::Error(InvalidData) -> ValueError, panic -> RuntimeError)."""
import os
import struct

import numpy as np

from .scene import Mesh


def _parse_index(value, length):
    try:
        index = int(value)
    except ValueError:
        return None
    return index - 1 if index > 0 else length + index


def _parse_point(tok):
    """parse_obj_point, src/io.rs:152-162"""
    try:
        return (float(tok[1]), float(tok[2]), float(tok[3]))
    except (ValueError, IndexError):
        raise ValueError("Failed to parse vertex in .OBJ") from None


def _parse_face(tok, vertices, normals, tri_v, tri_n):
    """parse_obj_face, src/io.rs:164-201: fan triangulation about the first corner"""
    vi, vni = [], []
    for vert in tok[1:]:
        args = (vert.split("/") + ["", "", ""])[:3]
        idx = _parse_index(args[0], len(vertices))
        if idx is None:
            raise ValueError("Invalid vertex index")
        vi.append(idx)
        vni.append(_parse_index(args[2], len(normals)))
    for i in range(1, len(vi) - 1):
        a, b, c = 0, i, i + 1
        tri_v.append((vertices[vi[a]], vertices[vi[b]], vertices[vi[c]]))
        if vni[a] is None or vni[b] is None or vni[c] is None:
            tri_n.append(None)
        else:
            tri_n.append((normals[vni[a]], normals[vni[b]], normals[vni[c]]))


def _open(path_or_file, mode="r"):
    return (open(path_or_file, mode), True) if isinstance(path_or_file, (str, os.PathLike)) else (path_or_file, False)


def load_obj(path_or_file):
    """Load a mesh geometry from a Wavefront .OBJ file -> Mesh."""
    f, own = _open(path_or_file)
    vertices, normals, tri_v, tri_n = [], [], [], []
    try:
        for line in f:
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            tok = line.split()
            if tok[0] == "v":
                vertices.append(_parse_point(tok))
            elif tok[0] == "vn":
                normals.append(_parse_point(tok))
            elif tok[0] == "f":
                _parse_face(tok, vertices, normals, tri_v, tri_n)
    finally:
        if own:
            f.close()
    return _mesh_of(tri_v, tri_n)


def _mesh_of(tri_v, tri_n):
    v = np.array(tri_v, dtype=np.float64).reshape(-1, 3, 3)
    if all(n is not None for n in tri_n):
        return Mesh(vertices=v, normals=np.array(tri_n, dtype=np.float64).reshape(-1, 3, 3))
    if all(n is None for n in tri_n):
        return Mesh(vertices=v)
    ns = []
    for tv, tn in zip(v, tri_n):
        if tn is None:
            c = np.cross(tv[1] - tv[0], tv[2] - tv[0])
            c = c / np.sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2])
            tn = (c, c, c)
        ns.append(tn)
    return Mesh(vertices=v, normals=np.array(ns, dtype=np.float64).reshape(-1, 3, 3))


def load_mtl(path_or_file):
    """load_mtl, src/io.rs:203-258 -> {name: Material}.  The reference creates a default material per `newmtl`
    and panics ("MTL loading not implemented") on the first property line; a property before any `newmtl` is
    InvalidData."""
    from .world import Material

    f, own = _open(path_or_file)
    materials, current = {}, None
    try:
        for line in f:
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            tok = line.split()
            if tok[0] == "newmtl":
                current = tok[1]
                materials.setdefault(current, Material.default())
            elif current is None:
                raise ValueError("Material was not specified with `newmtl` before properties were added")
            else:
                raise RuntimeError("MTL loading not implemented")
    finally:
        if own:
            f.close()
    return materials


def load_obj_with_mtl(obj_file, mtl_file):
    """load_obj_with_mtl, src/io.rs:84-150 -> [Object]: a new Object starts at every `usemtl` that names a
    different material than the previous one; `mtllib` lines are ignored."""
    from .world import Material, Object

    materials = load_mtl(mtl_file)
    f, own = _open(obj_file)
    vertices, normals, tri_v, tri_n, objects = [], [], [], [], []
    current_material, last_usemtl = Material.default(), None

    def flush():
        if tri_v:
            objects.append(Object(_mesh_of(list(tri_v), list(tri_n))).material(current_material))
            tri_v.clear()
            tri_n.clear()

    try:
        for line in f:
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            tok = line.split()
            if tok[0] == "v":
                vertices.append(_parse_point(tok))
            elif tok[0] == "vn":
                normals.append(_parse_point(tok))
            elif tok[0] == "f":
                _parse_face(tok, vertices, normals, tri_v, tri_n)
            elif tok[0] == "usemtl":
                if last_usemtl is None or last_usemtl != tok[1]:
                    flush()
                    if tok[1] not in materials:
                        raise ValueError(f"Could not found `usemtl {tok[1]}` in library")
                    current_material = materials[tok[1]]
                    last_usemtl = tok[1]
    finally:
        if own:
            f.close()
    flush()
    return objects


def load_stl(path_or_file):
    """load_stl, src/io.rs:264-364 -> Mesh.  Binary when the size equals 84 + 50 n (vertices and the facet normal
    widen from f32, :336-345; the facet normal is used for all three corners), else ASCII when the file starts
    with "solid ".  The ASCII reader reproduces the reference line for line: every line after the first must be
    `facet normal`, so a file that ends with an `endsolid` line is rejected as malformed (:294-299)."""
    f, own = _open(path_or_file, "rb")
    try:
        data = f.read()
    finally:
        if own:
            f.close()
    size = len(data)
    if size < 15:
        raise ValueError("Loaded .STL file is too short")
    if size >= 84:
        (n,) = struct.unpack_from("<I", data, 80)
        if size == 84 + n * 50:
            rec = np.frombuffer(data, dtype=np.dtype([("f", "<f4", (12,)), ("attr", "<u2")]), count=n, offset=84)
            vals = rec["f"].astype(np.float64).reshape(n, 4, 3)
            return Mesh(vertices=vals[:, 1:4, :].copy(), normals=np.repeat(vals[:, 0:1, :], 3, axis=1).copy())
    if data[:6] != b"solid ":
        raise ValueError("Loaded .STL file, but could not determine format")
    lines = data[6:].decode("utf-8").split("\n")
    if lines and lines[-1] == "":
        lines.pop()
    lines = iter(lines[1:])
    tri_v, tri_n = [], []
    for line in lines:
        line = line.strip()
        if not line.startswith("facet normal "):
            raise ValueError("Malformed STL file: expected `facet normal`")
        vn = tuple(float(t) for t in line[len("facet normal "):].split())[:3]
        next(lines)  # "outer loop"
        vs = []
        for _ in range(3):
            v = next(lines).strip()
            if not v.startswith("vertex "):
                raise ValueError("Malformed STL file: expected `vertex`")
            vs.append(tuple(float(t) for t in v[len("vertex "):].split())[:3])
        next(lines)  # "endloop"
        next(lines)  # "endfacet"
        tri_v.append(tuple(vs))
        tri_n.append((vn, vn, vn))
    return _mesh_of(tri_v, tri_n)

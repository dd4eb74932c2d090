"""Host mesh IO mirroring src/io.rs:28-74,152-201 (load_obj): 1-based / negative indices,
fan triangulation, per-vertex normals when every corner has one, else face normals."""
import numpy as np

from .scene import Mesh


def _parse_index(value, length):
    try:
        index = int(value)
    except ValueError:
        return None
    return index - 1 if index > 0 else length + index


def load_obj(path_or_file):
    """Load a mesh geometry from a Wavefront .OBJ file -> Mesh."""
    f = open(path_or_file) if isinstance(path_or_file, str) else path_or_file
    vertices, normals, tri_v, tri_n, has_n = [], [], [], [], []
    try:
        for line in f:
            line = line.strip()
            if not line or line.startswith("#"):
                continue
            tok = line.split()
            if tok[0] == "v":
                vertices.append((float(tok[1]), float(tok[2]), float(tok[3])))
            elif tok[0] == "vn":
                normals.append((float(tok[1]), float(tok[2]), float(tok[3])))
            elif tok[0] == "f":
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
    finally:
        if isinstance(path_or_file, str):
            f.close()
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

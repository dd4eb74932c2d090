"""Generates rpt_b200/assets/teapot_tris.npz from the reference's mesh asset.

Run in the build container only (reads /root/reference, which the GPU box does not have):
    python tests/golden/make_fixtures.py
The fixture is DERIVED DATA (parsed triangle soup of examples/teapot.obj, the Utah teapot
asset used by examples/teapot.rs:12 and examples/fractal_teapots.rs:51), not reference source.
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from rpt_b200.io import load_obj  # noqa: E402

HERE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "rpt_b200", "assets")


def main():
    mesh = load_obj("/root/reference/examples/teapot.obj")
    assert len(mesh) == 2256 and mesh.normals is not None
    np.savez_compressed(os.path.join(HERE, "teapot_tris.npz"), vertices=mesh.vertices, normals=mesh.normals)
    print("teapot:", mesh.vertices.shape, mesh.vertices.min((0, 1)), mesh.vertices.max((0, 1)))


if __name__ == "__main__":
    main()

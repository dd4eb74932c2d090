"""CPU: the Python host's scene lowering against the oracle's NATIVE scene builders.

Every GPU parity test feeds ONE lowered scene description (rpt_b200/world.py, scenes.py) to both the CUDA library and
the oracle, so a wrong matrix composition, a missing light/object duplication or a mistyped literal in the Python
transcription would be invisible to them.  oracle/builtin_scenes.hpp builds the five BASELINE scenes inside the oracle
straight from the literals of the reference's example sources, with the reference's construction path (polygon fan,
Transformable chain, SceneAdd duplication, KdTree of boxed instances).  Here: closest hits on random and camera rays and
whole renders at the same seed must be IDENTICAL between the two routes."""
import numpy as np
import pytest

import orc
from rpt_b200 import scenes


def _teapot():
    m = scenes.teapot_mesh()
    return m.vertices, m.normals


CASES = [
    ("cornell", lambda: scenes.cornell(), None),
    ("lampshade_pathtrace", lambda: scenes.lampshade("pathtrace"), None),
    ("lampshade_photonphoton", lambda: scenes.lampshade("photonphoton"), None),
    ("lampshade_beamphoton", lambda: scenes.lampshade("beamphoton"), None),
    ("lampshade_beambeam", lambda: scenes.lampshade("beambeam"), None),
    ("teapot", lambda: scenes.teapot(), _teapot),
    ("fractal_teapots", lambda: scenes.fractal_teapots(), _teapot),
]


def _rays(cam, n, rng, lo, hi):
    """half camera-like rays (from the eye into the scene), half random segments inside the scene's extent"""
    eye = np.array(cam.eye, float)
    o1 = np.tile(eye, (n // 2, 1))
    target = rng.uniform(lo, hi, size=(n // 2, 3))
    d1 = target - o1
    d1 /= np.linalg.norm(d1, axis=1, keepdims=True)
    o2 = rng.uniform(lo, hi, size=(n - n // 2, 3))
    d2 = rng.normal(size=(n - n // 2, 3))
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    return np.vstack([o1, o2]), np.vstack([d1, d2])


@pytest.mark.parametrize("name,make,teapot", CASES, ids=[c[0] for c in CASES])
def test_python_lowering_equals_native_scene(name, make, teapot):
    scene, cam, d = make()
    lowered = orc.OracleScene(scene)
    native = orc.OracleScene.builtin(name, teapot() if teapot else None)
    rng = np.random.default_rng(7)
    big = name.startswith(("cornell", "lampshade"))
    lo, hi = ((-50.0, -50.0, -100.0), (600.0, 600.0, 600.0)) if big else ((-7.0, -7.0, -7.0), (7.0, 7.0, 7.0))
    o, dd = _rays(cam, 6000 if name != "fractal_teapots" else 3000, rng, lo, hi)
    a = lowered.intersect_batch(o, dd)
    b = native.intersect_batch(o, dd)
    assert (a[1] != 0xFFFFFFFF).mean() > 0.3  # the ray set does hit the scene
    for x, y, what in zip(a, b, ("t", "object", "triangle", "instance", "normal")):
        assert np.array_equal(x, y), (name, what)
    # a whole render at the same seed: lights, materials, medium and their order are the same too
    r = scenes.make_renderer(scene, cam, d, width=24, height=16, num_samples=2, seed=3, max_bounces=2)
    fa, fb = orc.OracleFilm(lowered, cam, r.params()), orc.OracleFilm(native, cam, r.params())
    fa.sample(2); fb.sample(2)
    assert np.array_equal(fa.mean(), fb.mean()), name

"""rptx_intersect_batch_static (include/rptx.h): the allocation-free variant of rptx_intersect_batch, whose device
staging is a fixed `static __device__` array of the library.  No counterpart in the reference,
so the checker here is the existing entry point rptx_intersect_batch — itself checked
against the oracle in test_gpu_intersect.py — at that file's tolerances: ids exact, t bit-exact, normals to 1e-12.

Collected last on purpose.  Passes alone in a fresh process and after the
other tests: it builds its own context and scenes and depends on no state they leave behind."""
import os
import re

import numpy as np
import pytest

import orc
from rpt_b200 import _abi as A
from rpt_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _rays(scene, cam, defaults, w, h, spp, extra, seed):
    r = scenes.make_renderer(scene, cam, defaults, width=w, height=h, num_samples=spp)
    of = orc.OracleFilm(orc.OracleScene(scene), cam, r.params())
    o, d, _ = of.dump_rays(spp, 40000)
    rng = np.random.default_rng(seed)
    lo, hi = o.min(0), o.max(0)
    o2 = rng.uniform(lo - 1.0, hi + 1.0, size=(extra, 3))
    d2 = rng.normal(size=(extra, 3))
    d2 /= np.linalg.norm(d2, axis=1, keepdims=True)
    return r, np.concatenate([o, o2]), np.concatenate([d, d2])


def _same(ds, o, d):
    t, obj, prim, inst, nrm = ds.intersect_batch(o, d)
    ts, objs, prims, insts, nrms, hits = ds.intersect_batch_static(o, d)
    assert np.array_equal(objs, obj), f"object id mismatches: {(objs != obj).sum()} of {len(o)}"
    assert np.array_equal(prims, prim)
    assert np.array_equal(insts, inst)
    hit = obj != 0xFFFFFFFF
    assert np.array_equal(ts[hit], t[hit]), "t not bit-exact"
    assert np.isinf(ts[~hit]).all()
    assert np.allclose(nrms, nrm, rtol=0, atol=1e-12)
    assert hits == int(hit.sum())
    return hits


@pytest.mark.gpu
def test_static_scratch_batch_equals_the_allocating_one_on_two_calls_in_a_row():
    scene, cam, dflt = scenes.cornell()
    r, o, d = _rays(scene, cam, dflt, 48, 48, 2, 8000, seed=5)
    assert 10000 < len(o) <= A.INTERSECT_STATIC_MAX_RAYS
    ds = r.device_scene()
    # first call: a large batch; second call: a smaller, different batch over the same (now dirty) scratch —
    # stale rows of the first call and its hit count must not leak into the second
    h1 = _same(ds, o, d)
    cut = len(o) // 3
    h2 = _same(ds, o[-cut:][::-1].copy(), d[-cut:][::-1].copy())
    assert h1 > 0 and 0 < h2 < h1
    # and the first batch again: identical to its first result
    assert _same(ds, o, d) == h1
    # a second scene (triangle meshes + analytic shapes + medium) on the same context and scratch
    scene2, cam2, dflt2 = scenes.lampshade("pathtrace")
    r2, o2, d2 = _rays(scene2, cam2, dflt2, 32, 32, 2, 6000, seed=6)
    _same(r2.device_scene(), o2[:A.INTERSECT_STATIC_MAX_RAYS], d2[:A.INTERSECT_STATIC_MAX_RAYS])
    # n = 0 and NULL outputs
    t0, *_rest, hits0 = ds.intersect_batch_static(np.zeros((0, 3)), np.zeros((0, 3)))
    assert len(t0) == 0 and hits0 == 0


@pytest.mark.gpu
def test_static_scratch_batch_rejects_more_rays_than_the_scratch_holds():
    from rpt_b200 import RptxError

    scene, cam, dflt = scenes.cornell()
    r = scenes.make_renderer(scene, cam, dflt, width=8, height=8, num_samples=1)
    ds = r.device_scene()
    n = A.INTERSECT_STATIC_MAX_RAYS + 1
    o = np.zeros((n, 3))
    d = np.tile([0.0, 0.0, 1.0], (n, 1))
    with pytest.raises(RptxError, match="RPTX_INTERSECT_STATIC_MAX_RAYS"):
        ds.intersect_batch_static(o, d)
    # exactly the maximum is accepted, and the context is still usable afterwards
    t, *_ = ds.intersect_batch_static(o[:-1], d[:-1])
    assert len(t) == n - 1


def test_static_scratch_source_allocates_nothing_and_header_declares_the_limit():
    """CPU: the entry point's translation unit calls no allocating routine; the header, the Python binding and the
    source agree on the maximum."""
    src = open(os.path.join(ROOT, "rpt_b200", "csrc", "static_batch.cu")).read()
    code = re.sub(r"//[^\n]*", "", src)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "rptx_alloc_async", "rptx_free_async", "new ", "std::vector"):
        assert banned not in code, f"static_batch.cu calls {banned}"
    assert "cudaGetSymbolAddress" in code and "static __device__ __align__(256)" in code
    hdr = open(os.path.join(ROOT, "include", "rptx.h")).read()
    m = re.search(r"#define RPTX_INTERSECT_STATIC_MAX_RAYS (\d+)", hdr)
    assert m and int(m.group(1)) == A.INTERSECT_STATIC_MAX_RAYS

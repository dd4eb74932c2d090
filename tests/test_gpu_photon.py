"""-m gpu: parity of the photon integrators (src/photon.rs) against the oracle, pass by pass.

The three passes are checked separately because the walks are chaotic: a 1-ulp libm difference
(CUDA vs glibc sin/cos/acos/log) can flip a Russian-roulette or hit/miss decision of one photon
walk, which changes that walk's later deposits and thereby the k-nearest sets of the pixels
near them.  So
  pass 1 (emission)  same seed -> same photon records, except for a small fraction of flipped walks;
  pass 2 (maps)      built from the ORACLE's photons -> kNN sets and sphere radii bit-exact;
  pass 3 (gather)    same map, same seed -> per-pixel agreement to 1e-11;
  whole render       same seed -> relative RMSE within a stated tolerance.
"""
import numpy as np
import pytest

import orc
from rpt_b200 import PhotonList, PhotonRenderKind, RptxError, Scene, scenes
from rpt_b200.renderer import rows_keys

pytestmark = pytest.mark.gpu

N_PHOTONS = 20000
KINDS = {"photonphoton": PhotonRenderKind.PhotonMap, "beamphoton": PhotonRenderKind.PhotonPointBeam, "beambeam": PhotonRenderKind.PhotonBeamBeam}


def _setup(variant, w=24, h=24, spp=2, seed=5, **over):
    scene, cam, d = scenes.lampshade(variant)
    r = scenes.make_renderer(scene, cam, d, width=w, height=h, num_samples=spp, seed=seed, **over)
    return scene, cam, r, orc.OracleScene(scene)


def _match_rows(got, want):
    """Rows of both lists sharing a key -> (fraction of `want` keys found, fraction of shared rows equal to 1e-9)."""
    kg, kw = rows_keys(got), rows_keys(want)
    common, ig, iw = np.intersect1d(kg, kw, return_indices=True)
    same = np.isclose(got[ig, :12], want[iw, :12], rtol=1e-9, atol=1e-12).all(axis=1)
    return len(common) / max(len(kw), 1), same.mean() if len(common) else 1.0


@pytest.mark.parametrize("variant", ["photonphoton", "beambeam"])
def test_emission_same_seed(variant):
    scene, cam, r, osc = _setup(variant)
    kind = KINDS[variant]
    n = N_PHOTONS if variant != "beambeam" else 400000  # thinning keeps 0.1 % of the volume photons
    want_s, want_v = osc.emit_photons(cam, r.params(), n, kind)
    ph = r.emit_photons(n, kind)
    ns, nv = ph.counts()
    got_s, got_v = orc.sort_rows(ph.rows(0)), orc.sort_rows(ph.rows(1))
    assert abs(ns - len(want_s)) <= 0.01 * len(want_s) and abs(nv - len(want_v)) <= max(0.05 * len(want_v), 3)
    assert len(np.unique(rows_keys(got_s))) == ns and len(np.unique(rows_keys(got_v))) == nv  # (photon, depth) is unique per list
    found, same = _match_rows(got_s, want_s)
    assert found > 0.99 and same > 0.99, (found, same)
    found, same = _match_rows(got_v, want_v)
    assert found > 0.97 and same > 0.97, (found, same)
    # energy bookkeeping: first-hit surface photons carry watts/n * light colour (src/photon.rs:763-764)
    first = got_s[(rows_keys(got_s) & 0xFFFF) == 0]
    lc = np.array(scenes.hex_color(0xFFFEFA))
    assert np.allclose(first[:, 6:9], r.params().watts / n * lc, rtol=1e-12)
    st = ph.stats()
    assert st["emitted"] == n and st["rays"] >= n


def test_emission_is_shardable_and_deterministic():
    scene, cam, r, osc = _setup("photonphoton")
    whole = r.emit_photons(6000, 0)
    a = r.emit_photons(6000, 0, first=0, count=2500)
    b = r.emit_photons(6000, 0, first=2500, count=3500)
    for which in range(2):
        w = orc.sort_rows(whole.rows(which))
        ab = orc.sort_rows(np.concatenate([a.rows(which), b.rows(which)]))
        assert np.array_equal(w.view(np.uint64), ab.view(np.uint64))  # bit-identical records
    # host/device append = PhotonList::merge
    merged = PhotonList.empty(whole.ctx)
    merged.append_host(0, a.rows(0)); merged.append_host(0, b.rows(0))
    merged.append_device(1, a.lib.rptx_photons_device_ptr(a.h, 1), a.counts()[1])
    merged.append_device(1, b.lib.rptx_photons_device_ptr(b.h, 1), b.counts()[1])
    assert merged.counts() == whole.counts()
    assert np.array_equal(orc.sort_rows(merged.rows(1)).view(np.uint64), orc.sort_rows(whole.rows(1)).view(np.uint64))


def _gpu_map_from(r, rows_s, rows_v, kind):
    pl = PhotonList.empty(r.device_scene().ctx)
    # feed the rows in a scrambled order: the map build must canonicalise
    rng = np.random.default_rng(0)
    pl.append_host(0, rows_s[rng.permutation(len(rows_s))])
    pl.append_host(1, rows_v[rng.permutation(len(rows_v))])
    return r.build_photon_map(pl, kind)


def test_knn_exact():
    scene, cam, r, osc = _setup("photonphoton")
    rows_s, rows_v = osc.emit_photons(cam, r.params(), N_PHOTONS, 0)
    gm = _gpu_map_from(r, rows_s, rows_v, 0)
    om = orc.OraclePhotonMap(osc, rows_s, rows_v, 0)
    assert gm.stats()["surface"] == len(rows_s) and gm.stats()["volume"] == len(rows_v)
    rng = np.random.default_rng(3)
    for which, rows in ((0, rows_s), (1, rows_v)):
        q = np.concatenate([rows[rng.integers(0, len(rows), 300), :3] + rng.normal(scale=2.0, size=(300, 3)),  # near photons
                            rng.uniform(-50, 600, size=(200, 3)),                                                # anywhere
                            rows[:20, :3]])                                                                      # exactly on photons
        for k in (1, 10, 30, 100, 200):
            kg, dg, cg = gm.knn(which, q, k)
            ko, do, co = om.knn(which, q, k)
            assert np.array_equal(cg, co)
            assert np.array_equal(dg, do), f"d2 differs: which={which} k={k}"
            # keys equal except where two photons tie on distance
            neq = kg != ko
            assert neq.mean() < 1e-3
    # fewer photons than k
    small = _gpu_map_from(r, rows_s[:7], rows_v[:0], 0)
    kg, dg, cg = small.knn(0, rows_s[:3, :3], 10)
    assert (cg == 7).all() and np.isinf(dg[:, 7:]).all() and (kg[:, 7:] == np.uint64(2**64 - 1)).all()
    kg, dg, cg = small.knn(1, rows_s[:3, :3], 4)
    assert (cg == 0).all()


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_sharded_build_code_path_equals_the_plain_build(kind, monkeypatch):
    """rptx_photon_map_build_sharded on ONE rank (RPTX_FORCE_SHARDED_BUILD: splitters, cuts, exchange as local copies,
    segment sort, tree): the same map as rptx_photon_map_build, bit for bit - kNN answers, sphere radii, the gather image.
    tests/test_gpu_multi.py runs it over real ranks."""
    scene, cam, r, osc = _setup(["photonphoton", "beamphoton", "beambeam"][kind])
    rows_s, rows_v = osc.emit_photons(cam, r.params(), N_PHOTONS, kind)
    plain = _gpu_map_from(r, rows_s, rows_v, kind)
    monkeypatch.setenv("RPTX_FORCE_SHARDED_BUILD", "1")
    pl = PhotonList.empty(r.device_scene().ctx)
    rng = np.random.default_rng(8)
    pl.append_host(0, rows_s[rng.permutation(len(rows_s))])
    pl.append_host(1, rows_v[rng.permutation(len(rows_v))])
    shard = r.build_photon_map(pl, kind, sharded=True)
    assert shard.stats()["surface"] == plain.stats()["surface"] and shard.stats()["bvh_items"] == plain.stats()["bvh_items"]
    q = np.concatenate([rows_s[rng.integers(0, len(rows_s), 300), :3] + rng.normal(scale=2.0, size=(300, 3)), rng.uniform(-50, 600, size=(200, 3))])
    for which in ((0, 1) if kind != 2 else (0,)):
        ka, da, ca = plain.knn(which, q, 30)
        kb, db, cb = shard.knn(which, q, 30)
        assert np.array_equal(ca, cb) and np.array_equal(da, db) and np.array_equal(ka, kb)
    if kind == 1:
        ca, ra = plain.spheres()
        cb, rb = shard.spheres()
        assert np.array_equal(ca, cb) and np.array_equal(ra, rb)
    fa, fb = r.begin_gather(), r.begin_gather()
    fa.gather(plain, 2); fa.sample_commit(2)
    fb.gather(shard, 2); fb.sample_commit(2)
    assert np.array_equal(fa.mean(), fb.mean())
    # an empty list on this rank
    monkeypatch.setenv("RPTX_FORCE_SHARDED_BUILD", "1")
    e = PhotonList.empty(r.device_scene().ctx)
    em = r.build_photon_map(e, kind, sharded=True)
    assert em.stats()["surface"] == 0


def test_sphere_radii_exact():
    scene, cam, r, osc = _setup("beamphoton")
    rows_s, rows_v = osc.emit_photons(cam, r.params(), N_PHOTONS, 1)
    gm = _gpu_map_from(r, rows_s, rows_v, 1)
    om = orc.OraclePhotonMap(osc, rows_s, rows_v, 1)
    cg, rg = gm.spheres()
    co, ro = om.spheres()
    assert len(rg) == len(ro) == len(rows_v)
    ig, io = np.lexsort(cg.T), np.lexsort(co.T)
    assert np.array_equal(cg[ig], co[io])
    assert np.array_equal(rg[ig], ro[io])  # 10-NN radius, itself included (src/photon.rs:216-225)


@pytest.mark.parametrize("variant", ["photonphoton", "beamphoton", "beambeam"])
def test_gather_same_map_same_seed(variant):
    scene, cam, r, osc = _setup(variant, w=32, h=32, spp=3)
    kind = KINDS[variant]
    n = N_PHOTONS if variant != "beambeam" else 600000
    rows_s, rows_v = osc.emit_photons(cam, r.params(), n, kind)
    gm = _gpu_map_from(r, rows_s, rows_v, kind)
    om = orc.OraclePhotonMap(osc, rows_s, rows_v, kind)
    film = r.begin_gather()
    film.gather(gm, 3)
    film.sample_commit(3)
    got = film.mean()
    _, want, _ = om.gather(cam, r.params())
    assert np.isfinite(want).all()
    # Same map, same seed, bit-exact camera rays, kNN sets and visibility rays: only libm ulps
    # (log/exp/pow) and summation order remain.  Getting here needed two NaN quirks of the
    # reference reproduced (rays lying in a cube-face plane or in a mesh bounding-box plane,
    # trace.cuh) — photon visibility rays run inside the floor plane all the time.
    match = np.isclose(got, want, rtol=1e-11, atol=1e-300).all(axis=2).mean()
    rel_rmse = np.sqrt(np.mean((got - want) ** 2)) / np.mean(want)
    assert match >= 0.995, (match, rel_rmse)
    assert rel_rmse <= 1e-3, rel_rmse
    assert film.image().shape == (32, 32, 3)
    st = film.stats()
    assert st["path_samples"] == 32 * 32 * 3


@pytest.mark.parametrize("kind", [0, 1, 2])
def test_gather_warp_and_per_thread_kernels_agree(kind, monkeypatch):
    """The camera pass has a warp-cooperative kernel (default for k <= 128: shared kNN for the point estimates, a shared
    walk of the sphere / beam BVH with dense candidate lists for the beam estimates) and the per-thread one (larger k):
    same kNN sets, same candidates (the counter is equal), same contributions, different summation order."""
    variant = ["photonphoton", "beamphoton", "beambeam"][kind]
    scene, cam, r, osc = _setup(variant, w=24, h=24, spp=3)
    ph = r.emit_photons(30000, kind)
    gm = r.build_photon_map(ph, kind)
    means, cands, queries = [], [], []
    for per_thread in (False, True):
        if per_thread:
            monkeypatch.setenv("RPTX_GATHER_PER_THREAD", "1")
        f = r.begin_gather()
        f.gather(gm, 3); f.sample_commit(3)
        means.append(f.mean())
        st = f.stats()
        if kind == 0:
            assert st["gather_queries"] == 24 * 24 * 3  # one kNN per sample: at the collision or at the surface
        queries.append(st["gather_queries"])            # beam kinds: one per camera ray that hits a surface
        cands.append(st.get("gather_candidates"))
    assert queries[0] == queries[1] > 0
    assert np.allclose(means[0], means[1], rtol=1e-11, atol=1e-300)
    if cands[0] is not None and kind != 2:
        assert cands[0] == cands[1]
        assert kind == 0 or cands[0] > 0
    if kind == 2:
        # the warp pass walks the beam SEGMENTS (BeamSegDev): it counts segments whose beam's own box is hit, far fewer tests
        # than the beams the per-thread walk is handed; without the segments the two counters are equal
        assert 0 < cands[0]
        monkeypatch.delenv("RPTX_GATHER_PER_THREAD", raising=False)
        monkeypatch.setenv("RPTX_NO_BEAM_SEGMENTS", "1")
        plain = r.build_photon_map(ph, kind)
        f = r.begin_gather()
        f.gather(plain, 3); f.sample_commit(3)
        monkeypatch.delenv("RPTX_NO_BEAM_SEGMENTS")
        assert f.stats()["gather_candidates"] == cands[1]
        assert np.allclose(f.mean(), means[1], rtol=1e-11, atol=1e-300)
    # a one-item BVH (returned untested by BVH::traverse) and an empty one
    for nvol in (1, 0):
        rows_s, rows_v = ph.rows(0), ph.rows(1)[:nvol]
        pl = PhotonList.empty(r.device_scene().ctx)
        pl.append_host(0, rows_s); pl.append_host(1, rows_v)
        small = r.build_photon_map(pl, kind)
        out = []
        for per_thread in (False, True):
            if per_thread:
                monkeypatch.setenv("RPTX_GATHER_PER_THREAD", "1")
            else:
                monkeypatch.delenv("RPTX_GATHER_PER_THREAD", raising=False)
            f = r.begin_gather()
            f.gather(small, 2); f.sample_commit(2)
            out.append(f.mean())
        monkeypatch.delenv("RPTX_GATHER_PER_THREAD", raising=False)
        # (an empty volume map makes the point-point estimate 0/0 on both paths, like the reference's division)
        assert np.allclose(out[0], out[1], rtol=1e-11, atol=1e-300, equal_nan=True), nvol


def test_gather_shards_sum_to_whole():
    import torch

    scene, cam, r, osc = _setup("photonphoton", w=16, h=16, spp=5)
    ph = r.emit_photons(5000, 0)
    gm = r.build_photon_map(ph, 0)
    whole = r.begin_gather()
    whole.gather(gm, 5); whole.sample_commit(5)
    sums = []
    for i in range(2):
        rr = scenes.make_renderer(scene, cam, dict(r._p)).shard(2, i)
        rr._dscene = r.device_scene()
        f = rr.begin_gather()
        f.gather(gm, 5)
        sums.append(f.batch_sum_tensor().cpu().numpy().reshape(16, 16, 3).copy())
    assert np.allclose((sums[0] + sums[1]) / 5, whole.mean(), rtol=1e-12, atol=1e-300)


@pytest.mark.parametrize("variant,tol", [("photonphoton", 0.05), ("beamphoton", 0.08), ("beambeam", 0.08)])
def test_photon_render_same_seed(variant, tol):
    """Whole photon_render at equal seed: flipped walks perturb a few pixels; bound the image error."""
    scene, cam, r, osc = _setup(variant, w=24, h=24, spp=2)
    kind = KINDS[variant]
    n = N_PHOTONS if variant != "beambeam" else 400000
    img, got = r.photon_render(n, kind, want_mean=True)
    oimg, want, counts, secs = osc.photon_render(cam, r.params(), n, kind)
    assert np.isfinite(want).all() and np.isfinite(got).all()
    rel_rmse = np.sqrt(np.mean((got - want) ** 2)) / np.mean(want)
    assert rel_rmse <= tol, rel_rmse
    assert abs(np.mean(got) - np.mean(want)) <= 0.01 * np.mean(want)
    assert np.mean(np.abs(img.astype(int) - oimg.astype(int)) <= 2) > 0.9


def test_surface_only_scene_and_errors():
    """No medium: estimate_indirect is the surface estimate alone (src/photon.rs:606-609)."""
    scene, cam, d = scenes.cornell()
    r = scenes.make_renderer(scene, cam, d, width=20, height=20, num_samples=2, seed=2, gather_size=40)
    osc = orc.OracleScene(scene)
    rows_s, rows_v = osc.emit_photons(cam, r.params(), 8000, 0)
    assert len(rows_v) == 0
    gm = _gpu_map_from(r, rows_s, rows_v, 0)
    om = orc.OraclePhotonMap(osc, rows_s, rows_v, 0)
    film = r.begin_gather()
    film.gather(gm, 2); film.sample_commit(2)
    _, want, _ = om.gather(cam, r.params())
    got = film.mean()
    assert np.isclose(got, want, rtol=1e-11, atol=1e-300).all(axis=2).mean() >= 0.995
    # photon mapping needs a Light::Object (src/photon.rs:798 panics)
    from rpt_b200 import Camera, Light, Material, Object, sphere

    s2 = Scene()
    s2.add(Object(sphere()).material(Material.diffuse((0.5, 0.5, 0.5))))
    s2.add(Light.Point((1.0, 1.0, 1.0), (0.0, 5.0, 0.0)))
    r2 = scenes.make_renderer(s2, Camera(), dict(width=8, height=8, num_samples=1))
    with pytest.raises(RptxError, match="non-object lights"):
        r2.photon_map_render(100)
    with pytest.raises(RptxError):
        r.emit_photons(100, 0, first=50, count=51)
    with pytest.raises(RptxError):
        film.sample(1)  # a gather film has no path pool

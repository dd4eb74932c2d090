"""-m gpu: image parity of the wavefront path tracer (rptx_render_sample) against the oracle.

The reference is not reproducible (entropy-seeded RNG), so the oracle and the CUDA path share a
counter-based stream per (pixel, sample).  With the same seed every sample must then agree to
f64 rounding: the only differences are ulp-level (CUDA vs glibc libm, forward vs recursive
accumulation order), which can flip a discontinuous decision (hit/miss, the 1e-12 shadow test,
Russian roulette) for a vanishing fraction of samples.  Tolerances are written below."""
import numpy as np
import pytest

import orc
from rpt_b200 import scenes

pytestmark = pytest.mark.gpu

RTOL = 1e-9          # per-pixel relative agreement of the batch mean at equal seed
# Fraction of pixels (<= 4 spp each) that must agree to RTOL.  It is not 1.0 because the
# reference's own tests are epsilon-fragile: the NEE visibility test |t_hit - dist| < 1e-12
# (src/renderer.rs:396) and t_min = 1e-12 self-intersection sit at the rounding noise of
# coordinates ~500 (ulp 1.1e-13), so a 1-ulp libm difference in a bounce direction re-rolls
# them for ~0.1 % of rays per bounce (measured: 0 of 4096 samples differ at max_bounces 0,
# 5 at 1, 19 at 5).  A flipped sample is another valid outcome of the same estimator.
MIN_MATCH = 0.97
MAX_REL_RMSE = 2e-2  # whole-image relative RMSE bound at equal seed (flipped samples only)


@pytest.fixture(params=["wavefront", "fused", "persist"], autouse=True)
def path_mode(request, monkeypatch):
    """Every test here runs against both kernel sets (RPTX_PATH_MODE is read at render begin)."""
    monkeypatch.setenv("RPTX_PATH_MODE", request.param)
    return request.param


def _same_seed(scene, cam, defaults, w, h, spp, batches=1, **over):
    r = scenes.make_renderer(scene, cam, defaults, width=w, height=h, num_samples=spp * batches, seed=11, **over)
    film = r.begin()
    osc = orc.OracleScene(scene)
    of = orc.OracleFilm(osc, cam, r.params())
    for _ in range(batches):
        film.sample(spp)
        of.sample(spp)
    got, want = film.mean(), of.mean()
    match = np.isclose(got, want, rtol=RTOL, atol=1e-12).all(axis=2).mean()
    rel_rmse = np.sqrt(np.mean((got - want) ** 2)) / max(np.mean(want), 1e-300)
    return film, of, match, rel_rmse


def test_cornell_same_seed():
    scene, cam, d = scenes.cornell()
    film, of, match, rmse = _same_seed(scene, cam, d, 64, 64, 4)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse
    st, oc = film.stats(), of.counters()
    assert st["path_samples"] == 64 * 64 * 4
    assert abs(st["extend_rays"] - oc["extend_rays"]) <= 2e-3 * oc["extend_rays"]  # flipped samples change path length
    assert abs(st["shadow_rays"] - oc["shadow_rays"]) <= 2e-3 * oc["shadow_rays"]


def test_cornell_direct_only_is_exact():
    """max_bounces 0: no libm call sits between the RNG and the epsilon tests -> every pixel equal."""
    scene, cam, d = scenes.cornell()
    film, of, match, rmse = _same_seed(scene, cam, d, 64, 64, 4, max_bounces=0)
    assert match == 1.0, match
    assert rmse < 1e-12


def test_cornell_batches_image_and_variance():
    scene, cam, d = scenes.cornell()
    film, of, match, rmse = _same_seed(scene, cam, d, 48, 40, 2, batches=3, max_bounces=2)
    assert match >= MIN_MATCH and rmse <= MAX_REL_RMSE
    gi, oi = film.image().astype(int), of.image().astype(int)
    assert np.mean(np.abs(gi - oi) <= 1) > 0.97  # box filter r=1 spreads a flipped sample over 9 pixels
    gv, ov = film.variance(), of.variance()
    assert abs(gv - ov) <= 2e-2 * abs(ov), (gv, ov)


def test_film_image_and_variance_exact_when_samples_equal():
    """Buffer::image / variance semantics (box filter, gamma, truncation, unbiased variance of
    batch means) on a direct-only render where every sample equals the oracle's."""
    scene, cam, d = scenes.cornell()
    film, of, match, rmse = _same_seed(scene, cam, d, 40, 30, 2, batches=3, max_bounces=0)
    assert match == 1.0
    assert np.array_equal(film.image(), of.image())
    gv, ov = film.variance(), of.variance()
    assert abs(gv - ov) <= 1e-12 * abs(ov), (gv, ov)


def test_lampshade_medium_same_seed():
    scene, cam, d = scenes.lampshade("pathtrace")
    film, of, match, rmse = _same_seed(scene, cam, d, 48, 48, 4)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse


def test_teapot_same_seed():
    scene, cam, d = scenes.teapot()
    film, of, match, rmse = _same_seed(scene, cam, d, 80, 45, 2, max_bounces=3)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse


def test_fractal_teapots_same_seed():
    scene, cam, d = scenes.fractal_teapots()
    film, of, match, rmse = _same_seed(scene, cam, d, 80, 45, 1, max_bounces=2)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse


def test_shards_sum_to_whole():
    """Multi-GPU sample sharding: two shards' batch sums add up to the unsharded batch."""
    import ctypes as C

    scene, cam, d = scenes.cornell()
    w, h, spp = 32, 32, 6
    whole = scenes.make_renderer(scene, cam, d, width=w, height=h, num_samples=spp, seed=3).begin()
    whole.sample(spp)
    parts = []
    for i in range(2):
        f = scenes.make_renderer(scene, cam, d, width=w, height=h, num_samples=spp, seed=3).shard(2, i).begin()
        f.sample_trace(spp)
        parts.append(f)
    # emulate the all-reduce on the host: copy sums, add, compare to the whole batch mean * spp
    import torch

    sums = []
    for f in parts:
        class _W:  # __cuda_array_interface__ view of the film's batch-sum buffer
            __cuda_array_interface__ = dict(shape=(h, w, 3), typestr="<f8", data=(f.batch_sum_ptr(), False), version=3)
        sums.append(torch.as_tensor(_W(), device="cuda").cpu().numpy().copy())
    total = (sums[0] + sums[1]) / spp
    assert np.allclose(total, whole.mean(), rtol=1e-12, atol=1e-14)


def test_render_api_and_errors():
    from rpt_b200 import Filter, Renderer, RptxError

    scene, cam, d = scenes.cornell()
    img = Renderer.new(scene, cam).width(40).height(30).filter(Filter.Box(1)).max_bounces(1).num_samples(2).render()
    assert img.shape == (30, 40, 3) and img.dtype == np.uint8 and img.max() > 0
    seen = []
    Renderer(scene, cam).width(16).height(16).num_samples(5).iterative_render(2, lambda it, buf: seen.append((it, buf.image().shape)))
    assert [s[0] for s in seen] == [2, 4, 5]
    with pytest.raises(RptxError):
        Renderer(scene, cam).width(0).height(16).render()


def test_fused_equals_wavefront(monkeypatch):
    """Same arithmetic, same RNG draw order: the two kernel sets give the same batch (NEE terms are
    added in a different order only when several shadow-casting lights pass)."""
    for name in ("cornell", "lampshade", "teapot"):
        scene, cam, d = getattr(scenes, name)() if name != "lampshade" else scenes.lampshade("pathtrace")
        means, modes = [], []
        for mode in ("wavefront", "fused", "persist"):
            monkeypatch.setenv("RPTX_PATH_MODE", mode)
            f = scenes.make_renderer(scene, cam, d, width=40, height=30, num_samples=3, seed=9, max_bounces=3).begin()
            f.sample(3)
            means.append(f.mean())
            modes.append(f.stats()["path_mode"])
            st = f.stats()
            assert st["path_samples"] == 40 * 30 * 3
        assert modes == [0, 1, 3]
        assert np.allclose(means[0], means[1], rtol=1e-12, atol=1e-300), name
        assert np.allclose(means[0], means[2], rtol=1e-12, atol=1e-300), name


def test_list_walk_equals_hierarchy(monkeypatch):
    """Small scenes skip the TLAS and walk the instance list (closest_hit_small); the hits, and so
    the batch, are identical bit for bit to the hierarchy's."""
    monkeypatch.setenv("RPTX_PATH_MODE", "wavefront")
    for name in ("cornell", "lampshade"):
        scene, cam, d = scenes.cornell() if name == "cornell" else scenes.lampshade("pathtrace")
        means = []
        for off in (False, True):
            if off:
                monkeypatch.setenv("RPTX_NO_SMALL_SCENE", "1")
            else:
                monkeypatch.delenv("RPTX_NO_SMALL_SCENE", raising=False)
            f = scenes.make_renderer(scene, cam, d, width=64, height=48, num_samples=4, seed=11, max_bounces=4).begin()
            f.sample(4)
            means.append(f.mean())
        monkeypatch.delenv("RPTX_NO_SMALL_SCENE", raising=False)
        assert np.array_equal(means[0], means[1]), name


def test_trim_gives_cached_memory_back_and_changes_nothing():
    """rptx_ctx_trim (Context.trim): the block cache, the pool and the spare film slab go back to the device between
    workloads; the next render of the same batch is identical, bit for bit, and photon renders still work."""
    scene, cam, d = scenes.cornell()
    r = scenes.make_renderer(scene, cam, d, width=48, height=40, num_samples=3, seed=4, max_bounces=3)
    f = r.begin()
    f.sample(3)
    a = f.mean()
    f.close()
    r.device_scene().ctx.trim()
    r.device_scene().ctx.trim()  # idempotent
    f = r.begin()
    f.sample(3)
    assert np.array_equal(a, f.mean())
    f.close()
    scene2, cam2, d2 = scenes.lampshade("photonphoton")
    r2 = scenes.make_renderer(scene2, cam2, d2, width=16, height=16, num_samples=2, seed=4)
    _, m1 = r2.photon_render(20000, 0, want_mean=True)
    r2.device_scene().ctx.trim()
    _, m2 = r2.photon_render(20000, 0, want_mean=True)
    assert np.allclose(m1, m2, rtol=1e-11, atol=1e-300, equal_nan=True)


def test_tiled_slots_equal_row_major(monkeypatch):
    """Slots walk the image in 8x4 tiles when the size allows it; the RNG is keyed by the pixel, so the
    batch is the one the row-major walk produces, bit for bit."""
    monkeypatch.setenv("RPTX_PATH_MODE", "wavefront")
    scene, cam, d = scenes.cornell()
    means = []
    for linear in (False, True):
        if linear:
            monkeypatch.setenv("RPTX_LINEAR_PIXELS", "1")
        else:
            monkeypatch.delenv("RPTX_LINEAR_PIXELS", raising=False)
        f = scenes.make_renderer(scene, cam, d, width=64, height=48, num_samples=3, seed=2, max_bounces=3).begin()
        f.sample(3)
        means.append(f.mean())
    monkeypatch.delenv("RPTX_LINEAR_PIXELS", raising=False)
    assert np.array_equal(means[0], means[1])


def test_film_slab_is_recycled_and_pool_size_does_not_matter(monkeypatch):
    """A second film of the same context reuses the first one's slab; the number of paths in flight
    (RPTX_POOL_LOG2) changes scheduling only: same batch."""
    monkeypatch.setenv("RPTX_PATH_MODE", "wavefront")
    scene, cam, d = scenes.lampshade("pathtrace")
    means = []
    for log2 in (None, "12", None):
        if log2:
            monkeypatch.setenv("RPTX_POOL_LOG2", log2)
        else:
            monkeypatch.delenv("RPTX_POOL_LOG2", raising=False)
        r = scenes.make_renderer(scene, cam, d, width=56, height=40, num_samples=6, seed=4, max_bounces=4)
        f = r.begin()
        f.sample(2)
        f.sample(4)
        means.append(f.mean())
        f.close()
    monkeypatch.delenv("RPTX_POOL_LOG2", raising=False)
    assert np.array_equal(means[0], means[2])
    assert np.allclose(means[0], means[1], rtol=1e-12, atol=1e-300)  # lane sums are added in a different order


def test_sah_trees_equal_lbvh_trees(monkeypatch):
    """Object trees are built on the GPU with the surface-area criterion (PLOC, build_ploc); RPTX_LBVH_ONLY forces the Morton
    LBVH everywhere, RPTX_BVH_BUILDER=host the binned-SAH trees from the host (round 1's default for up to 65 K items).
    The accelerator only orders the search: same hits, same batch, bit for bit."""
    monkeypatch.setenv("RPTX_PATH_MODE", "wavefront")
    for name in ("teapot", "fractal_teapots", "monomial_glass"):
        scene, cam, d = getattr(scenes, name)()
        means = []
        for env in ({}, {"RPTX_LBVH_ONLY": "1"}, {"RPTX_BVH_BUILDER": "host"}, {"RPTX_BVH_BUILDER": "ploc", "RPTX_NO_SMALL_SCENE": "1"},
                    {"RPTX_BVH_BUILDER": "ploc", "RPTX_PLOC_RADIUS": "3"}):
            for k in ("RPTX_LBVH_ONLY", "RPTX_BVH_BUILDER", "RPTX_NO_SMALL_SCENE"):
                monkeypatch.delenv(k, raising=False)
            for k, v in env.items():
                monkeypatch.setenv(k, v)
            r = scenes.make_renderer(scene, cam, d, width=96, height=64, num_samples=2, seed=5, max_bounces=2)
            f = r.begin()
            f.sample(2)
            means.append(f.mean())
            f.close()
            r.device_scene().close()
        for k in ("RPTX_LBVH_ONLY", "RPTX_BVH_BUILDER", "RPTX_NO_SMALL_SCENE"):
            monkeypatch.delenv(k, raising=False)
        for m in means[1:]:
            assert np.array_equal(means[0], m), name
    # Cornell through the hierarchy (the small-scene walk off): polygons are one-leaf meshes under a PLOC / host-SAH TLAS
    scene, cam, d = scenes.cornell()
    means = []
    for env in ({"RPTX_NO_SMALL_SCENE": "1"}, {"RPTX_NO_SMALL_SCENE": "1", "RPTX_BVH_BUILDER": "host"}):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        f = scenes.make_renderer(scene, cam, d, width=64, height=48, num_samples=2, seed=5, max_bounces=3).begin()
        f.sample(2)
        means.append(f.mean())
        f.close()
        monkeypatch.delenv("RPTX_BVH_BUILDER", raising=False)
    monkeypatch.delenv("RPTX_NO_SMALL_SCENE", raising=False)
    assert np.array_equal(means[0], means[1])


def test_monomial_glass_same_seed():
    """examples/monomial_glass.rs (MonomialSurface + Phong materials + plane, solid environment): same-seed parity,
    and the same scene lit by the surface itself as a Light::Object (MonomialSurface::sample, :109-124)."""
    from rpt_b200 import Light, Material, Object, hex_color, monomial_surface

    scene, cam, d = scenes.monomial_glass()
    film, of, match, rmse = _same_seed(scene, cam, d, 80, 60, 4, max_bounces=2)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse
    scene, cam, d = scenes.monomial_glass(environment=(0.0, 0.0, 0.0))
    glow = Object(monomial_surface(2.0, 4.0).translate((0.0, -1.0, 0.0))).material(Material.light(hex_color(0xFFFFFF), 5.0))
    scene.objects[0] = glow
    scene.add(Light.Object(glow))
    film, of, match, rmse = _same_seed(scene, cam, d, 80, 60, 4, max_bounces=1)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse
    assert film.mean().mean() > 1e-3


def test_one_vertex_kernel_equals_three_kernels(monkeypatch):
    """max_bounces == 0 (the reference's builder default): k_shade_final does shade + NEE visibility + film add +
    regeneration in one kernel.  Same batch and the same ray counters as k_shade + k_shadow + k_connect, bit for bit,
    on a BVH scene with point / directional / ambient lights and on a small scene with an area light; several samples
    per slot so that regeneration runs."""
    monkeypatch.setenv("RPTX_PATH_MODE", "wavefront")
    monkeypatch.setenv("RPTX_POOL_LOG2", "12")
    for make, size in ((scenes.fractal_teapots, (96, 56)), (scenes.cornell, (64, 48)), (scenes.monomial_glass, (64, 48))):
        scene, cam, d = make()
        out = []
        for three in (False, True):
            if three:
                monkeypatch.setenv("RPTX_NO_ONE_VERTEX", "1")
            else:
                monkeypatch.delenv("RPTX_NO_ONE_VERTEX", raising=False)
            f = scenes.make_renderer(scene, cam, d, width=size[0], height=size[1], num_samples=5, seed=4, max_bounces=0).begin()
            f.sample(5)
            st = f.stats()
            out.append((f.mean(), st["extend_rays"], st["shadow_rays"], st["shadow_traced"], st["path_samples"]))
        monkeypatch.delenv("RPTX_NO_ONE_VERTEX", raising=False)
        assert np.array_equal(out[0][0], out[1][0]), make.__name__
        assert out[0][1:] == out[1][1:], (make.__name__, out[0][1:], out[1][1:])
        assert out[0][4] == size[0] * size[1] * 5


def _feature_scene(hdri=True):
    """Everything off the BASELINE configs in one scene: HDRI environment with bilinear lookup
    (src/environment.rs:25-52), thin-lens camera (src/camera.rs:71-77), mirror / transmissive / Phong materials
    (src/material.rs:196-262), a sphere light, a point light and a plane."""
    import math

    from rpt_b200 import Camera, Hdri, Light, Material, Object, Scene, cube, hex_color, plane, sphere

    s = Scene()
    if hdri:
        h, w = 16, 32
        yy, xx = np.mgrid[0:h, 0:w]
        buf = np.stack([0.2 + 0.8 * xx / w, 0.3 + 0.5 * yy / h, 0.9 - 0.6 * xx / w], axis=-1)
        s.environment = Hdri(w, h, buf)
    else:
        s.environment = (0.05, 0.06, 0.08)
    s.add(Object(sphere().translate((-2.2, 0.0, 0.0))).material(Material.mirror()))
    s.add(Object(sphere().translate((0.0, 0.0, 0.0))).material(Material.clear(1.5)))
    s.add(Object(sphere().translate((2.2, 0.0, 0.0))).material(Material.specular(hex_color(0x2A9D8F), 8.0)))
    s.add(Object(cube().rotate_y(0.6).scale((1.0, 0.6, 1.0)).translate((0.5, -0.7, 2.0))).material(Material.transparent(hex_color(0xE9C46A), 1.3)))
    s.add(Object(plane((0.0, 1.0, 0.0), -1.0)).material(Material.diffuse(hex_color(0xAAAAAA))))
    glow = sphere().scale((0.5, 0.5, 0.5)).translate((0.0, 3.0, 1.0))
    s.add(Object(glow).material(Material.light(hex_color(0xFFFFFF), 20.0)))
    s.add(Light.Object(Object(glow).material(Material.light(hex_color(0xFFFFFF), 20.0))))
    s.add(Light.Ambient((0.02, 0.02, 0.02)))
    s.add(Light.Point((30.0, 30.0, 30.0), (3.0, 4.0, 4.0)))
    cam = Camera.look_at((0.0, 1.5, 9.0), (0.0, 0.0, 0.0), (0.0, 1.0, 0.0), math.pi / 5.0).focus((0.0, 0.0, 0.0), 0.15)
    return s, cam, dict(width=64, height=48, max_bounces=4, num_samples=4)


def test_feature_scene_same_seed():
    for hdri in (True, False):
        scene, cam, d = _feature_scene(hdri)
        film, of, match, rmse = _same_seed(scene, cam, d, 64, 48, 4)
        assert match >= MIN_MATCH, (hdri, match)
        assert rmse <= MAX_REL_RMSE, (hdri, rmse)
        assert film.mean().mean() > 1e-3


def test_colored_glowing_fog_same_seed():
    """Medium::colored_glowing_fog (src/medium.rs:99-121): emission 10, colour by y > 250, phase pi / 4 (precedence quirk)."""
    from rpt_b200 import Medium

    scene, cam, d = scenes.lampshade("pathtrace")
    scene.media[0] = Medium.colored_glowing_fog(5e-5, 3e-3)
    film, of, match, rmse = _same_seed(scene, cam, d, 48, 48, 4)
    assert match >= MIN_MATCH, match
    assert rmse <= MAX_REL_RMSE, rmse

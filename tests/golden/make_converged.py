"""Generates tests/golden/converged_<workload>.npz: the ORACLE's converged renders of the BASELINE scenes, the
reference images of the relative-RMSE check (oracle at >= 16 384 spp at a reduced size, the GPU
rendered at the same size; photon integrators: >= 1e6 photons).

    python tests/golden/make_converged.py [workload ...]        # ~15 min of 8 host cores for all of them

Each file holds `mean` ([h,w,3] f64 linear mean image, before box filter and gamma), `spp`, `seed`, `photons`
(photon workloads) and `rel_rmse_mc`: the relative RMSE an INDEPENDENT render of the same spp is expected to show
against `mean` from Monte Carlo noise alone, estimated from the two half-renders A, B the mean is made of:
sigma = rmse(A - B) / 2 is the noise of their mean, an independent render differs from it by sqrt(2) sigma.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
sys.path.insert(0, os.path.join(HERE, ".."))
import orc  # noqa: E402
from rpt_b200 import scenes  # noqa: E402

PHOTON_KINDS = {"photonphoton": 0, "beamphoton": 1, "beambeam": 2}
# workload -> (width, height, spp, photons)
SIZES = {
    "cornell": (128, 128, 16384, 0),
    "lampshade": (128, 72, 16384, 0),
    "fractal_teapots": (64, 36, 16384, 0),
    "fractal_teapots_mb5": (64, 36, 16384, 0),
    "photonphoton": (48, 48, 256, 1_000_000),
    "beamphoton": (48, 48, 128, 1_000_000),
    "beambeam": (48, 48, 128, 1_000_000),
}


def scene_of(name):
    if name == "cornell":
        return scenes.cornell()
    if name == "lampshade":
        return scenes.lampshade("pathtrace")
    if name == "fractal_teapots":
        return scenes.fractal_teapots()
    if name == "fractal_teapots_mb5":
        s, c, d = scenes.fractal_teapots()
        return s, c, dict(d, max_bounces=5)
    return scenes.lampshade(name)


def rel_rmse(a, b):
    return float(np.sqrt(np.mean((a - b) ** 2)) / np.mean(b))


def main(names):
    for name in names:
        w, h, spp, photons = SIZES[name]
        scene, cam, d = scene_of(name)
        t0 = time.time()
        osc = orc.OracleScene(scene)
        halves = []
        if photons:
            over = dict(width=w, height=h, num_samples=spp)
            if name != "photonphoton":
                over["watts"] = (200000.0 / 13650.0) * photons
            for seed in (101, 202):  # two independent photon maps + camera passes
                r = scenes.make_renderer(scene, cam, d, seed=seed, **over)
                _, mean, counts, _ = osc.photon_render(cam, r.params(), photons, PHOTON_KINDS[name])
                halves.append(mean)
        else:
            r = scenes.make_renderer(scene, cam, d, width=w, height=h, num_samples=spp, seed=101)
            of = orc.OracleFilm(osc, cam, r.params())
            for _ in range(2):  # two batches of spp / 2 consecutive sample indices
                of.sample(spp // 2)
            full = of.mean()
            r2 = scenes.make_renderer(scene, cam, d, width=w, height=h, num_samples=spp // 2, seed=101)
            of2 = orc.OracleFilm(osc, cam, r2.params())
            of2.sample(spp // 2)
            a = of2.mean()
            halves = [a, 2.0 * full - a]
        mean = 0.5 * (halves[0] + halves[1])
        sigma = rel_rmse(halves[0] - halves[1] + mean, mean) / 2.0
        out = os.path.join(HERE, f"converged_{name}.npz")
        np.savez_compressed(out, mean=mean, spp=spp, seed=101, photons=photons, rel_rmse_mc=np.sqrt(2.0) * sigma)
        print(f"{name}: {w}x{h} x {spp} spp, photons {photons}: mean {mean.mean():.6g}, expected independent relRMSE {np.sqrt(2.0) * sigma:.4g}, "
              f"{time.time() - t0:.0f} s -> {out}", flush=True)


if __name__ == "__main__":
    main(sys.argv[1:] or list(SIZES))

"""Photon passes at scale on one GPU: emit / build / gather times per kind (not a bench line)."""
import sys, os, time, json
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from rpt_b200 import scenes

def run(variant, kind, n, w, h, spp, oracle=False):
    scene, cam, d = scenes.lampshade(variant)
    if variant != "photonphoton":
        d = dict(d, watts=(200000.0 / 13650.0) * n)  # examples/volumetric_beambeam_lampshade.rs:139-148
    r = scenes.make_renderer(scene, cam, d, width=w, height=h, num_samples=spp, seed=1)
    r.device_scene()
    t0 = time.perf_counter(); ph = r.emit_photons(n, kind); t1 = time.perf_counter()
    ns, nv = ph.counts(); st = ph.stats()
    m = r.build_photon_map(ph, kind); t2 = time.perf_counter()
    ph.close()
    film = r.begin_gather(); film.gather(m, spp); film.sample_commit(spp); t3 = time.perf_counter()
    fs = film.stats(); mean = film.mean()
    out = dict(variant=variant, photons=n, surface=ns, volume=nv, emit_s=t1 - t0, emit_dev_ms=st["emit_ms"], rays=st["rays"],
               build_s=t2 - t1, build_dev_ms=m.stats()["build_ms"], bvh_items=m.stats()["bvh_items"], gather_s=t3 - t2, gather_dev_ms=fs["device_ms"],
               samples=w * h * spp, mphotons_per_s=n / (t1 - t0) / 1e6, msamples_per_s=w * h * spp / (t3 - t2) / 1e6,
               mean=float(np.nanmean(mean)), finite=bool(np.isfinite(mean).all()))
    if oracle:
        import orc
        osc = orc.OracleScene(scene)
        img, omean, counts, secs = osc.photon_render(cam, r.params(), n, kind)
        out.update(cpu_emit_s=secs[0], cpu_build_s=secs[1], cpu_gather_s=secs[2], cpu_mean=float(np.nanmean(omean)),
                   rel_rmse=float(np.sqrt(np.nanmean((mean - omean) ** 2)) / np.nanmean(omean)))
    print(json.dumps(out), flush=True)

if __name__ == "__main__":
    sizes = [int(float(x)) for x in sys.argv[1].split(",")]
    kinds = dict(photonphoton=0, beamphoton=1, beambeam=2)
    import io, contextlib
    for v, k in kinds.items():  # warm-up: module load, allocator pools
        with contextlib.redirect_stdout(io.StringIO()):
            run(v, k, 20000, 32, 32, 2)
    for n in sizes:
        for v, k in kinds.items():
            if v == "beamphoton" and n > 2e7: continue
            run(v, k, n, 128, 128, 100 if v == "photonphoton" else 50, oracle=(n <= 1e6))

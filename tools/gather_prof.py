"""Short photon gather run for ncu: python tools/gather_prof.py <variant> <photons> <spp>"""
import sys
sys.path.insert(0, ".")
from rpt_b200 import scenes
variant, n, spp = sys.argv[1], int(float(sys.argv[2])), int(sys.argv[3])
kind = dict(photonphoton=0, beamphoton=1, beambeam=2)[variant]
scene, cam, d = scenes.lampshade(variant)
r = scenes.make_renderer(scene, cam, d, width=128, height=128, num_samples=spp, seed=1)
ph = r.emit_photons(n, kind); m = r.build_photon_map(ph, kind)
film = r.begin_gather(); film.gather(m, spp); film.sample_commit(spp)
print(variant, n, spp, "gather ms", film.stats()["device_ms"], film.stats())

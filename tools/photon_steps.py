"""Per-step pass times of one photon workload (is the build time stable from step to step?):
python tools/photon_steps.py beamphoton 1e7 8"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from rpt_b200 import Context
from rpt_b200 import dist as rd
from rpt_b200.scenes import make_renderer

kindname, n, steps = sys.argv[1], int(float(sys.argv[2])), int(sys.argv[3])
kind = bench.PHOTON_KINDS[kindname]
scene, cam, d = bench.workload(kindname)
if kindname != "photonphoton":
    d = dict(d, watts=(200000.0 / 13650.0) * n)
ctx = Context(0)
r = make_renderer(scene, cam, d, seed=0)
r._ctx = ctx
r.device_scene()
for i in range(steps):
    film, pmap, tm = rd.photon_render(r, n, kind)
    film.image()
    film.close(); pmap.close()
    print(i, {k: round(tm[k], 1) for k in ("emit", "merge", "build", "gather")}, flush=True)

"""Where the end-to-end call spends its time: python tools/e2e_diag.py <workload> <spp>"""
import sys, time
sys.path.insert(0, ".")
import bench
from rpt_b200 import Context
from rpt_b200.scenes import make_renderer
name, spp = sys.argv[1], int(sys.argv[2])
scene, cam, d = bench.workload(name)
d = dict(d, num_samples=spp)
ctx = Context(0)
for it in range(3):
    t = [time.perf_counter()]
    rr = make_renderer(scene, cam, d, seed=1); rr._ctx = ctx; t.append(time.perf_counter())
    ds = rr.device_scene(); t.append(time.perf_counter())
    f2 = rr.begin(); t.append(time.perf_counter())
    f2.sample(spp); t.append(time.perf_counter())
    img = f2.image(); t.append(time.perf_counter())
    f2.close(); t.append(time.perf_counter())
    ds.close(); t.append(time.perf_counter())
    names = ["make_renderer", "lower+upload+BVH", "begin", "sample", "image", "film close", "scene close"]
    print(it, {n: round(1e3 * (b - a), 2) for n, a, b in zip(names, t, t[1:])}, "device ms", round(f2.stats()["device_ms"], 2) if False else "")

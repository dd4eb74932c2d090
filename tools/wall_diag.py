"""Where a step's wall time goes: python tools/wall_diag.py <workload> <spp>"""
import sys, time
sys.path.insert(0, ".")
import bench
from rpt_b200.scenes import make_renderer
name, spp = sys.argv[1], int(sys.argv[2])
scene, cam, d = bench.workload(name)
d = dict(d, num_samples=spp)
r = make_renderer(scene, cam, d, seed=0)
t = time.perf_counter(); film = r.begin(); print("begin s", round(time.perf_counter() - t, 4))
for i in range(4):
    t0 = time.perf_counter(); film.sample_trace(spp); t1 = time.perf_counter(); film.sample_commit(spp); t2 = time.perf_counter()
    st = film.stats()
    print(i, "trace wall ms", round(1e3 * (t1 - t0), 2), "device ms", round(st["device_ms"], 2), "commit ms", round(1e3 * (t2 - t1), 2), "waves", st["waves"],
          "k ms", round(st["extend_ms"] + st["shade_ms"] + st["shadow_ms"] + st["connect_ms"], 2))
t = time.perf_counter(); film.close(); print("close s", round(time.perf_counter() - t, 4))

import sys, os
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, collections
import orc
from rpt_b200 import scenes
scene, cam, d = scenes.lampshade("photonphoton")
r = scenes.make_renderer(scene, cam, d)
ds = r.device_scene(); osc = orc.OracleScene(scene)
rng = np.random.default_rng(5)
n = 20000
o = np.stack([rng.uniform(0, 556, n), np.zeros(n), rng.uniform(0, 559, n)], axis=1)
a = rng.uniform(0, 2 * np.pi, n)
dd = np.stack([np.cos(a), np.zeros(n), np.sin(a)], axis=1)
t, obj, prim, inst, _ = ds.intersect_batch(o, dd)
to, objo, primo, insto, _ = osc.intersect_batch(o, dd)
bad = obj != objo
print("bad", bad.sum())
print(collections.Counter(zip(obj[bad].tolist(), objo[bad].tolist())).most_common(12))
for i in np.nonzero(bad)[0][:8]:
    print(o[i], dd[i], "gpu", obj[i], t[i], "orc", objo[i], to[i])

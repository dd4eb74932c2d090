"""Worker of tests/test_gpu_multi.py (launched with torch.distributed.run, one rank per GPU): renders the same
batches sharded over the ranks through the LIBRARY's NCCL communicator (rptx_comm_init) and, on rank 0, also on
one GPU alone; writes both to an .npz for the test to compare."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from rpt_b200 import Context, scenes  # noqa: E402
from rpt_b200 import dist as rd  # noqa: E402


def main(out_path):
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = rd.init_comm(Context(local))
    assert ctx.comm_info() == (world, rank)
    out = {}
    cases = [("cornell", scenes.cornell(), dict(width=48, height=40, num_samples=7, max_bounces=3)),
             ("lampshade", scenes.lampshade("pathtrace"), dict(width=48, height=40, num_samples=6))]
    for name, (scene, cam, d), over in cases:
        for mode in ("wavefront", "persist"):
            os.environ["RPTX_PATH_MODE"] = mode
            r = scenes.make_renderer(scene, cam, d, seed=5, **over)
            r._ctx = ctx
            r.shard(world, rank)
            film = r.begin()
            rd.sample_batch(film, over["num_samples"])       # batch 1
            rd.sample_batch(film, 3)                          # batch 2: fewer samples than ranks when world > 3
            out[f"{name}_{mode}_multi"] = film.mean()
            film.close(); r.device_scene().close()
    # photon integrators: emission shards + all-gather + gather shards + all-reduce inside rptx_photon_render
    scene, cam, d = scenes.lampshade("photonphoton")
    for kind, kname in ((0, "photonphoton"), (1, "beamphoton"), (2, "beambeam")):
        r = scenes.make_renderer(scene, cam, d, seed=3, width=24, height=24, num_samples=4)
        r._ctx = ctx
        _, mean = r.photon_render(30000, kind, want_mean=True)
        out[f"{kname}_multi"] = mean
        r.device_scene().close()
    # the sharded map build itself (rptx_photon_map_build_sharded): same kNN answers, bit for bit, as the one-rank build
    os.environ["RPTX_SHARDED_MAPS_MIN_WORLD"] = "2"  # the library splits the point maps from 4 ranks on; here at any count
    rngq = np.random.default_rng(4)
    queries = np.stack([rngq.uniform(0, 556, 400), rngq.uniform(0, 548, 400), rngq.uniform(0, 559, 400)], axis=1)
    for kind, kname in ((0, "photonphoton"), (1, "beamphoton")):
        r = scenes.make_renderer(scene, cam, d, seed=3, width=24, height=24, num_samples=4)
        r._ctx = ctx
        lo, hi = rd.photon_range(50000, world, rank)
        part = r.emit_photons(50000, kind, first=lo, count=hi - lo)
        pm = r.build_photon_map(part, kind, sharded=True)
        for which in (0, 1):
            keys, d2, cnt = pm.knn(which, queries, 12)
            out[f"knn_{kname}_{which}_multi_keys"], out[f"knn_{kname}_{which}_multi_d2"] = keys, d2
        if kind == 1:
            c, rad = pm.spheres()
            out["spheres_multi_c"], out["spheres_multi_r"] = c, rad
        st = pm.stats()
        out[f"stats_{kname}_multi"] = np.array([st["surface"], st["volume"], st["bvh_items"]])
        pm.close(); part.close(); r.device_scene().close()
    # and once more through rptx_photon_render with the split maps (the renders above used the rank count's default)
    r = scenes.make_renderer(scene, cam, d, seed=3, width=24, height=24, num_samples=4)
    r._ctx = ctx
    _, mean = r.photon_render(30000, 0, want_mean=True)
    out["photonphoton_split_multi"] = mean
    r.device_scene().close()
    os.environ.pop("RPTX_SHARDED_MAPS_MIN_WORLD")
    dist.barrier()
    if rank == 0:
        solo = Context(local)  # no communicator: the whole batch on one GPU
        for name, (scene, cam, d), over in cases:
            for mode in ("wavefront", "persist"):
                os.environ["RPTX_PATH_MODE"] = mode
                r = scenes.make_renderer(scene, cam, d, seed=5, **over)
                r._ctx = solo
                film = r.begin()
                film.sample(over["num_samples"])
                film.sample(3)
                out[f"{name}_{mode}_solo"] = film.mean()
                film.close(); r.device_scene().close()
        scene, cam, d = scenes.lampshade("photonphoton")
        for kind, kname in ((0, "photonphoton"), (1, "beamphoton"), (2, "beambeam")):
            r = scenes.make_renderer(scene, cam, d, seed=3, width=24, height=24, num_samples=4)
            r._ctx = solo
            _, mean = r.photon_render(30000, kind, want_mean=True)
            out[f"{kname}_solo"] = mean
            r.device_scene().close()
        for kind, kname in ((0, "photonphoton"), (1, "beamphoton")):
            r = scenes.make_renderer(scene, cam, d, seed=3, width=24, height=24, num_samples=4)
            r._ctx = solo
            allp = r.emit_photons(50000, kind)
            pm = r.build_photon_map(allp, kind)
            for which in (0, 1):
                keys, d2, cnt = pm.knn(which, queries, 12)
                out[f"knn_{kname}_{which}_solo_keys"], out[f"knn_{kname}_{which}_solo_d2"] = keys, d2
            if kind == 1:
                c, rad = pm.spheres()
                out["spheres_solo_c"], out["spheres_solo_r"] = c, rad
            st = pm.stats()
            out[f"stats_{kname}_solo"] = np.array([st["surface"], st["volume"], st["bvh_items"]])
            pm.close(); allp.close(); r.device_scene().close()
        np.savez(out_path, **out)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1])

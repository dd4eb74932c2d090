"""-m gpu, needs >= 2 devices: an N-rank render through the library's NCCL communicator (rptx_comm_init,
rptx_render_sample_allreduce, rptx_photon_map_build_sharded) equals the 1-GPU render.  The RNG is keyed by the global
sample / photon index and every slot adds its samples in a fixed order, so the path tracer's batch sums differ only
by the association of the cross-rank sum: bit-identical is not promised by an all-reduce, 1e-13 relative is."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _devices():
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.skipif(_devices() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_render_equals_single_gpu(tmp_path):
    n = min(_devices(), 4)
    out = str(tmp_path / "mp.npz")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1", "--master-port", "29517",
           os.path.join(ROOT, "tests", "mp_parity.py"), out]
    env = dict(os.environ)
    env.pop("RPTX_PATH_MODE", None)
    res = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-4000:]
    z = np.load(out)
    for name in ("cornell", "lampshade"):
        for mode in ("wavefront", "persist"):
            a, b = z[f"{name}_{mode}_multi"], z[f"{name}_{mode}_solo"]
            assert np.allclose(a, b, rtol=1e-13, atol=1e-300), (name, mode, np.abs(a - b).max())
    # the map the ranks build together is the one-rank map: same photons found, same distances, same sphere radii
    for kname in ("photonphoton", "beamphoton"):
        assert np.array_equal(z[f"stats_{kname}_multi"], z[f"stats_{kname}_solo"]), kname
        for which in (0, 1):
            assert np.array_equal(z[f"knn_{kname}_{which}_multi_keys"], z[f"knn_{kname}_{which}_solo_keys"]), (kname, which)
            assert np.array_equal(z[f"knn_{kname}_{which}_multi_d2"], z[f"knn_{kname}_{which}_solo_d2"]), (kname, which)
    assert np.array_equal(z["spheres_multi_c"], z["spheres_solo_c"]) and np.array_equal(z["spheres_multi_r"], z["spheres_solo_r"])
    assert np.allclose(z["photonphoton_split_multi"], z["photonphoton_solo"], rtol=1e-11, atol=1e-300)
    for kname in ("photonphoton", "beamphoton", "beambeam"):
        a, b = z[f"{kname}_multi"], z[f"{kname}_solo"]
        # same photon set and same map (the build canonicalises row order by key); gather sums associate differently
        assert np.allclose(a, b, rtol=1e-11, atol=1e-300), (kname, np.abs(a - b).max())

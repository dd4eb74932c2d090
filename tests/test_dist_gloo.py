"""CPU, world_size 2 over gloo: the multi-GPU host plumbing (rpt_b200/dist.py) with the oracle
standing in for the device film.  Rank r traces samples s % 2 == r; after the all-reduce both
ranks must hold exactly the unsharded batch."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import orc
from rpt_b200 import dist as rd
from rpt_b200 import scenes

W, H, SPP = 12, 10, 5


class OracleShardFilm:
    """Test double with the Film protocol of rpt_b200.renderer (sample_trace / batch_sum_tensor /
    sample_commit); radiance per (pixel, sample) comes from the oracle."""

    def __init__(self, world, rank, seed=3):
        scene, cam, d = scenes.cornell()
        self.r = scenes.make_renderer(scene, cam, d, width=W, height=H, num_samples=SPP, seed=seed, max_bounces=2)
        self.of = orc.OracleFilm(orc.OracleScene(scene), cam, self.r.params(), 1)
        self.world, self.rank, self.base = world, rank, 0
        self.sum = torch.zeros(H * W * 3, dtype=torch.float64)
        self.means = []

    def sample_trace(self, iterations):
        mine = [s for s in range(iterations) if s % self.world == self.rank]
        assert len(mine) == rd.local_samples(iterations, self.world, self.rank)
        ys, xs = np.mgrid[0:H, 0:W]
        acc = np.zeros((H * W, 3))
        for s in mine:
            acc += self.of.sample_radiance(xs.ravel(), ys.ravel(), np.full(H * W, self.base + s, np.uint64))
        self.sum.copy_(torch.from_numpy(acc.reshape(-1)))

    def batch_sum_tensor(self):
        return self.sum

    def sample_commit(self, iterations):
        self.means.append(self.sum.numpy().reshape(H, W, 3) / iterations)
        self.base += iterations


class OraclePhotonList:
    """Test double with the PhotonList protocol used by rd.merge_photons."""

    def __init__(self, rows_s=None, rows_v=None):
        self.rows = [np.zeros((0, 13)) if rows_s is None else rows_s, np.zeros((0, 13)) if rows_v is None else rows_v]

    def device_tensor(self, which):
        return torch.from_numpy(np.ascontiguousarray(self.rows[which]))

    def append_tensor(self, which, rows):
        self.rows[which] = np.concatenate([self.rows[which], rows.numpy()])


def _photon_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        scene, cam, d = scenes.lampshade("photonphoton")
        r = scenes.make_renderer(scene, cam, d, width=8, height=8, num_samples=2, seed=4)
        osc = orc.OracleScene(scene)
        lo, hi = rd.photon_range(3001, world, rank)
        s, v = osc.emit_photons(cam, r.params(), 3001, 0, first=lo, count=hi - lo, threads=1)
        full = rd.merge_photons(OraclePhotonList(s, v), OraclePhotonList())
        out[rank] = (full.rows[0], full.rows[1])
    finally:
        dist.destroy_process_group()


def test_photon_lists_merge_to_the_unsharded_emission():
    """Rank r shoots photon_range(N, 2, r); the all-gathered lists are the single-rank emission."""
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_photon_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    scene, cam, d = scenes.lampshade("photonphoton")
    r = scenes.make_renderer(scene, cam, d, width=8, height=8, num_samples=2, seed=4)
    ws, wv = orc.OracleScene(scene).emit_photons(cam, r.params(), 3001, 0, threads=1)
    for rank in range(world):
        s, v = out[rank]
        assert np.array_equal(orc.sort_rows(s).view(np.uint64), ws.view(np.uint64))
        assert np.array_equal(orc.sort_rows(v).view(np.uint64), wv.view(np.uint64))
    assert np.array_equal(out[0][0], out[1][0])  # same order on every rank (rank-major)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        film = OracleShardFilm(world, rank)
        rd.sample_batch(film, SPP)
        rd.sample_batch(film, 3)
        # ragged all-gather of "photon records": rank r owns photon indices photon_range(11, 2, r)
        lo, hi = rd.photon_range(11, world, rank)
        rec = torch.arange(lo, hi, dtype=torch.float64).repeat_interleave(2 + rank).reshape(-1, 1) * torch.ones(1, 4, dtype=torch.float64)
        full, counts = rd.all_gather_ragged(rec)
        out[rank] = (film.means[0], film.means[1], full.numpy(), counts)
    finally:
        dist.destroy_process_group()


def test_sharded_batches_equal_the_unsharded_render():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    whole = OracleShardFilm(1, 0)
    whole.sample_trace(SPP); whole.sample_commit(SPP)
    whole.sample_trace(3); whole.sample_commit(3)
    for r in range(world):
        m0, m1, full, counts = out[r]
        assert np.allclose(m0, whole.means[0], rtol=1e-13, atol=1e-15)
        assert np.allclose(m1, whole.means[1], rtol=1e-13, atol=1e-15)
        assert counts == [5 * 2, 6 * 3] and full.shape == (28, 4)
        assert np.array_equal(full[:, 0], np.concatenate([np.repeat(np.arange(0, 5), 2), np.repeat(np.arange(5, 11), 3)]))
    assert np.array_equal(out[0][0], out[1][0])  # every rank commits the same film


def test_shard_arithmetic():
    for it in (1, 2, 5, 16, 257):
        for world in (1, 2, 3, 8):
            assert sum(rd.local_samples(it, world, r) for r in range(world)) == it
    for n in (0, 1, 7, 10**8):
        for world in (1, 2, 8):
            rs = [rd.photon_range(n, world, r) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == n and all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))

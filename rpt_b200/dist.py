"""Multi-GPU plumbing of the render path: one process per GPU, torch.distributed (NCCL over
NVLink on the box, gloo in the CPU tests).

The reference has one address space: rayon tasks per image row / per photon, a serial
`add_samples` and a serial `PhotonList::merge` (src/renderer.rs:159-171, src/photon.rs:663-675).
Here the two exchange steps of the path are
  * path tracing / gather pass: every rank traces the samples s with s % world == rank of each
    pixel into its own f64 batch-sum film; one all-reduce(sum) makes every rank hold the batch;
  * photon pass: rank r emits the photon indices [lo_r, hi_r); the deposit records are
    all-gathered (ragged) so that every rank builds the same full map.
Nothing here computes radiance; `film` / `photons` are the C-ABI objects (or test doubles with
the same methods).
"""
import os

import torch
import torch.distributed as dist


def world_rank(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_world_size(group), dist.get_rank(group)
    return 1, 0


def local_samples(iterations, world, rank):
    """How many of the batch's sample indices {0..iterations-1} satisfy s % world == rank
    (the same arithmetic as rptx_render_sample_trace)."""
    return (iterations - rank + world - 1) // world if iterations > rank else 0


def photon_range(photon_count, world, rank):
    """Contiguous photon-index range of a rank; RNG streams are keyed by the global index, so
    the union over ranks is the single-GPU photon set."""
    return photon_count * rank // world, photon_count * (rank + 1) // world


def all_reduce_sum_(t, group=None):
    if world_rank(group)[0] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


def all_gather_ragged(t, group=None):
    """Concatenate per-rank [n_r, ...] tensors in rank order (n_r differs per rank): counts are
    all-gathered first, rows padded to the maximum, gathered, trimmed."""
    world, _ = world_rank(group)
    if world == 1:
        return t, [t.shape[0]]
    n = torch.tensor([t.shape[0]], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(counts, n, group=group)
    counts = [int(c.item()) for c in counts]
    m = max(counts)
    pad = t
    if t.shape[0] < m:
        pad = torch.zeros((m,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad.contiguous(), group=group)
    return torch.cat([p[:c] for p, c in zip(parts, counts)], dim=0), counts


def init_comm(ctx, group=None):
    """Join this rank's rptx context into an NCCL communicator that the LIBRARY owns (rptx_comm_init): rank 0 makes
    the id, torch.distributed only carries its 128 bytes to the other ranks.  After this the exchanges of
    sample_batch / photon_render run inside librptx.so on its own stream; torch is not on the data path."""
    world, rank = world_rank(group)
    if world == 1 or ctx.comm_info()[0] > 1:
        return ctx
    box = [ctx.comm_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0, group=group)
    ctx.comm_init(box[0], world, rank)
    return ctx


def _lib_comm(obj):
    ctx = getattr(obj, "ctx", None)
    return ctx is not None and hasattr(ctx, "comm_info") and ctx.comm_info()[0] > 1


def sample_batch(film, iterations, group=None):
    """Renderer::sample(iterations) across ranks: trace the local shard, all-reduce the f64
    batch sums, commit the same batch mean on every rank."""
    if _lib_comm(film):
        film.sample(iterations)  # trace -> ncclAllReduce -> commit, all inside the library
        return
    film.sample_trace(iterations)
    all_reduce_sum_(film.batch_sum_tensor(), group)
    if film.batch_sum_tensor().is_cuda:
        torch.cuda.current_stream().synchronize()
    film.sample_commit(iterations)


def device_f64_view(ptr, n, device):
    """torch view (no copy) of a device buffer owned by the C library."""

    class _V:
        __cuda_array_interface__ = dict(shape=(int(n),), typestr="<f8", data=(int(ptr), False), version=3)

    return torch.as_tensor(_V(), device=device)


def merge_photons(local, empty, group=None):
    """PhotonList::merge (src/photon.rs:153-160,675) across ranks: all-gather the ragged row lists
    of `local` (protocol: device_tensor(which) -> [n,13] f64 tensor) into `empty`
    (protocol: append_tensor(which, rows)).  World size 1 returns `local` itself."""
    world, _ = world_rank(group)
    if world == 1:
        return local
    for which in (0, 1):
        rows, _ = all_gather_ragged(local.device_tensor(which), group)
        empty.append_tensor(which, rows)
    return empty


def photon_render(renderer, photon_count, kind, group=None):
    """photon_render (src/photon.rs:655-720) across ranks: emit this rank's photon range, merge
    the lists, build the (identical) map everywhere, gather this rank's sample shard, all-reduce
    the batch sums, commit.  Returns (film, map, timings_ms); `renderer` must already carry
    .shard(world, rank)."""
    import time

    from .renderer import PhotonList

    world, rank = world_rank(group)
    lo, hi = photon_range(photon_count, world, rank)
    lib_comm = _lib_comm(renderer.device_scene())
    t0 = time.perf_counter()
    local = renderer.emit_photons(photon_count, kind, first=lo, count=hi - lo)
    t1 = time.perf_counter()
    sharded_build = bool(lib_comm) and world > 1 and not os.environ.get("RPTX_REPLICATED_MAP_BUILD")
    if sharded_build:
        full = local  # the ranks split the build by Morton range; no merged row list exists (rptx_photon_map_build_sharded)
    elif lib_comm:
        full = local.allgather()
    else:
        full = merge_photons(local, PhotonList.empty(renderer.device_scene().ctx) if world > 1 else None, group)
    t2 = time.perf_counter()
    pmap = renderer.build_photon_map(full, kind, sharded=sharded_build)
    t3 = time.perf_counter()
    film = renderer.begin_gather()
    n = renderer._p["num_samples"]
    film.gather(pmap, n)
    if lib_comm:
        film.sample_allreduce()
    else:
        all_reduce_sum_(film.batch_sum_tensor(), group)
        if film.batch_sum_tensor().is_cuda:
            torch.cuda.current_stream().synchronize()
    film.sample_commit(n)
    t4 = time.perf_counter()
    st = pmap.stats()
    counts = (st["surface"], st["bvh_items"] if st["bvh_items"] else st["volume"])
    if full is not local:
        local.close()
    full.close()
    return film, pmap, dict(emit=1e3 * (t1 - t0), merge=1e3 * (t2 - t1), build=1e3 * (t3 - t2), gather=1e3 * (t4 - t3), surface=counts[0], volume=counts[1])

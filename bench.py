#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 light-transport core.

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME]

A "step" is one pass of the hot path over one batch: Renderer::sample(spp) over the whole
image (src/renderer.rs:158-171), i.e. width*height*spp path-samples per GPU.  Metric:
Mpath-samples/s (BASELINE.json).  Default workload = BASELINE.json configs[1]:
examples/fractal_teapots.rs at 1920x1080, 256 spp, the example's own max_bounces (0).

N > 1 (torchrun, one rank per GPU): weak scaling — the batch grows to N*spp samples per pixel,
rank r traces the samples s % N == r, the per-rank f64 batch sums are all-reduced with NCCL and
every rank commits the same batch mean (north_star: "per-GPU accumulation buffers are reduced").

`--impl reference` times the reference's algorithm on the host CPU cores.  rpt is Rust and
cannot be built in this image (no rustc/cargo, dependencies not vendored), so that arm runs the
oracle port (oracle/, C++ f64 restatement with rpt's linear object loop, kd-trees, recursion and
one task per image row) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "Mpath-samples/s"

# algorithmic HBM bytes of one extend-kernel ray (DESIGN.md "k_extend"): queue entry 4 B +
# ray origin/direction 48 B read; t 8 B + object id 4 B + normal 24 B written
EXTEND_BYTES_PER_RAY = 4 + 48 + 8 + 4 + 24
# dram__bytes_read.sum + dram__bytes_write.sum of one k_extend launch / its rays, from the committed
# `ncu --set full` capture of the same command (profiles/r1_v2_summary.md); None where not captured
EXTEND_DRAM_BYTES_PER_RAY = {"fractal_teapots": 88.1, "cornell": 148.7}


_OVERRIDES = {}
# photon workloads (BASELINE configs[3], configs[4]): a step = one photon_render (emit -> merge ->
# build -> gather) of --photons photons per GPU at the example's camera pass; metric Mphotons/s
PHOTON_KINDS = {"photonphoton": 0, "beamphoton": 1, "beambeam": 2}
PHOTON_ROW_BYTES = 104


def workload(name):
    from rpt_b200 import scenes

    if name == "fractal_teapots":
        scene, cam, d = scenes.fractal_teapots()
    elif name == "fractal_teapots_mb5":
        scene, cam, d = scenes.fractal_teapots()
        d = dict(d, max_bounces=5)
    elif name == "teapot":
        scene, cam, d = scenes.teapot()
    elif name == "cornell":
        scene, cam, d = scenes.cornell()
    elif name == "lampshade":
        scene, cam, d = scenes.lampshade("pathtrace")
    elif name in PHOTON_KINDS:
        scene, cam, d = scenes.lampshade(name)
    else:
        raise SystemExit(f"unknown workload {name}")
    for k, v in _OVERRIDES.items():
        if v:
            d[k] = v
    return scene, cam, d


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def cpu_sample(scene, cam, params, seconds, threads=0):
    """Time the oracle on every k-th image row of the same workload for ~`seconds`."""
    import orc

    osc = orc.OracleScene(scene)
    of = orc.OracleFilm(osc, cam, params, threads)
    h, w = params.height, params.width
    cores = threads or orc.lib().orc_hardware_threads()
    # calibrate on 1 spp over ~64 spread rows
    stride = max(1, h // 64)
    sec, ctr = of.time_rows(0, h, stride, 1)
    rows = len(range(0, h, stride))
    rate = rows * w / max(sec, 1e-9)  # path-samples/s
    spp = int(max(1, min(params.num_samples, seconds * rate / (rows * w))))
    sec, ctr = of.time_rows(0, h, stride, spp)
    paths = rows * w * spp
    return dict(value=paths / sec / 1e6, unit=METRIC, cores=cores, kind="port",
                sample=f"{rows} of {h} rows (every {stride}th) x {w} px x {spp} spp = {paths} path-samples in {sec:.2f} s; "
                       f"{(ctr[0] + ctr[1]) / sec / 1e6:.2f} Mrays/s"), sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scene, cam, d = workload(args.workload)
    from rpt_b200.scenes import make_renderer

    if args.workload in PHOTON_KINDS:
        n_total = int(args.photons) * args.gpus
        if args.workload != "photonphoton":
            d = dict(d, watts=(200000.0 / 13650.0) * n_total)
        params = make_renderer(scene, cam, d).params()
        kind = PHOTON_KINDS[args.workload]
        t0 = time.perf_counter()
        runs = [cpu_photons(scene, cam, params, kind, args.cpu_seconds) for _ in range(max(args.steps, 1))]
        secs = time.perf_counter() - t0
        last = runs[-1]
        v = sum(x["value"] for x in runs) / len(runs)
        last["value"] = v
        print(json.dumps(dict(impl="reference", metric="Mphotons/s", value=v, unit="Mphotons/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                              ms_per_step=1e3 * secs / max(args.steps, 1), higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64",
                              data="synthetic", config=dict(workload=f"{args.workload}: bounded photon_render sample, see cpu_baseline.sample", photons=n_total),
                              cpu_baseline=last, e2e=dict(value=v, unit="Mphotons/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))))
        return
    params = make_renderer(scene, cam, d).params()
    for _ in range(args.warmup):
        cpu_sample(scene, cam, params, 0.5)
    vals, secs, last = [], 0.0, None
    for _ in range(args.steps):
        last, sec = cpu_sample(scene, cam, params, args.cpu_seconds)
        vals.append(last["value"])
        secs += sec
    v = sum(vals) / len(vals)
    last["value"] = v
    line = dict(impl="reference", metric=METRIC, value=v, unit=METRIC, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=1e3 * secs / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64", data="synthetic",
                config=config_of(args, d), cpu_baseline=last,
                e2e=dict(value=v, unit=METRIC, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


def config_of(args, d):
    return dict(workload=f"{args.workload}: examples/{args.workload.split('_mb')[0]}.rs scene, {d['width']}x{d['height']}, {d['num_samples']} spp, "
                         f"max_bounces {d['max_bounces']}",
                width=d["width"], height=d["height"], spp=d["num_samples"], max_bounces=d["max_bounces"],
                l2_policy="per-step path state (13-23 GB: up to 33.5 M paths in flight) and film exceed the 126 MB L2; no explicit flush",
                parallelism=f"sample-sharded x{args.gpus}")


def run_ours(args):
    import numpy as np
    import torch

    from rpt_b200.scenes import make_renderer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    from rpt_b200 import Context

    ctx = Context(local)
    scene, cam, d = workload(args.workload)
    w, h, spp = d["width"], d["height"], d["num_samples"]
    r = make_renderer(scene, cam, d, seed=0)
    r._ctx = ctx
    r.shard(world, rank)
    batch = spp * world  # weak scaling: every rank traces `spp` samples per pixel
    film = r.begin()

    from rpt_b200 import dist as rd

    def step():
        rd.sample_batch(film, batch)  # trace local shard -> NCCL all-reduce of the f64 batch sums -> commit
        return film.stats()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    launches0 = film.stats()["kernel_launches"]
    waves0 = film.stats()["waves"]
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    dev_ms = ext_ms = con_ms = shd_ms = sdw_ms = 0.0
    rays_e = rays_s = 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        before = film.stats()
        st = step()
        dev_ms += st["device_ms"]  # CUDA events on the library's own stream around the traced batch
        ext_ms += st["extend_ms"]; con_ms += st["connect_ms"]; shd_ms += st["shade_ms"]; sdw_ms += st["shadow_ms"]
        rays_e += st["extend_rays"] - before["extend_rays"]
        rays_s += st["shadow_rays"] - before["shadow_rays"]
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    launches = film.stats()["kernel_launches"] - launches0
    # max over ranks of the per-rank time
    t = torch.tensor([wall, dev_ms / 1e3], device=f"cuda:{local}", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    wall_max, dev_max = float(t[0]), float(t[1])
    step_time = max(wall_max, 1e-12)  # the barrier-bracketed region; the device-event sum is reported beside it (kernel_ms_per_step.device)
    paths = w * h * spp * world * args.steps
    value = paths / step_time / 1e6

    # e2e: the user-facing call from HOST data: lower + upload scene, GPU BVH build, render, image to host
    h2d = sum(v.nbytes + (0 if n is None else n.nbytes) for v, n in r.device_scene().low.meshes)
    e2e_t = 0.0
    n_e2e = max(1, min(args.steps, 3))
    film.close()  # its path-state slab returns to the context, as it does between a user's successive render() calls
    for _ in range(n_e2e):
        barrier()
        t1 = time.perf_counter()
        rr = make_renderer(scene, cam, d, seed=1)
        rr._ctx = ctx
        rr.shard(world, rank)
        rr.device_scene()
        t2 = time.perf_counter()
        f2 = rr.begin()
        t3 = time.perf_counter()
        rd.sample_batch(f2, batch)
        t4 = time.perf_counter()
        img = f2.image()
        barrier()
        t5 = time.perf_counter()
        e2e_t += t5 - t1
        assert img.shape == (h, w, 3)
        f2.close(); rr.device_scene().close()
        if rank == 0:
            print("e2e ms: scene %.1f begin %.1f sample %.1f image %.1f close %.1f" % (
                1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3), 1e3 * (t5 - t4), 1e3 * (time.perf_counter() - t5)), file=sys.stderr)
    te = torch.tensor([e2e_t / n_e2e], device=f"cuda:{local}", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_val = w * h * spp * world / float(te[0]) / 1e6

    if rank == 0:
        peak, which = measured_peak_gbs()
        ext_s = ext_ms / 1e3 / args.steps
        ach = EXTEND_BYTES_PER_RAY * (rays_e / args.steps) / max(ext_s, 1e-12) / 1e9
        cpu = None
        if world == 1 and args.cpu_seconds > 0:
            cpu, _ = cpu_sample(scene, cam, r.params(), args.cpu_seconds)
        line = dict(metric=METRIC, value=value, unit=METRIC, n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=1e3 * step_time / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64",
                    data="synthetic", config=config_of(args, d),
                    e2e=dict(value=e2e_val, unit=METRIC, h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(w * h * 3),
                             note="Renderer -> scene lowering + upload + GPU BVH build + path pool alloc + render + rgb8 image to host"),
                    gpu_launches=int(launches), clocks=clk,
                    roofline=dict(bound="hbm", kernel="k_extend", achieved=ach, peak=peak, unit="GB/s", frac=ach / peak,
                                  traffic=(EXTEND_DRAM_BYTES_PER_RAY[args.workload] * rays_e / max(st["waves"] - waves0, 1)
                                           if args.workload in EXTEND_DRAM_BYTES_PER_RAY else None),
                                  launches=int(st["waves"] - waves0), algorithmic_bytes_per_launch=EXTEND_BYTES_PER_RAY * rays_e / max(st["waves"] - waves0, 1),
                                  peak_source=which,
                                  note="per launch (one wave); closest-hit traversal is issue-bound (62 % of issue slots, 16-23 of 32 lanes), "
                                       "not HBM-bound: the scene stays in L2 and DRAM traffic equals the path-state bytes; see DESIGN.md and profiles/"),
                    rays_per_s=dict(extend=rays_e / max(ext_ms / 1e3, 1e-12),
                                    # one-vertex paths (max_bounces 0): NEE visibility runs inside k_shade_final, no shadow kernel time of its own
                                    shadow=(rays_s / (sdw_ms / 1e3) if sdw_ms > 0.05 * max(shd_ms, 1e-9) else None),
                                    total_step=(rays_e + rays_s) / step_time),
                    kernel_ms_per_step=dict(extend=ext_ms / args.steps, shade=shd_ms / args.steps, shadow=sdw_ms / args.steps, connect=con_ms / args.steps,
                                            device=dev_ms / args.steps, wall=1e3 * wall / args.steps),
                    cpu_baseline=cpu)
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def run_photons(args):
    import numpy as np
    import torch

    from rpt_b200 import Context
    from rpt_b200 import dist as rd
    from rpt_b200.scenes import make_renderer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ctx = Context(local)
    kind = PHOTON_KINDS[args.workload]
    scene, cam, d = workload(args.workload)
    n_total = int(args.photons) * world  # weak scaling
    if args.workload != "photonphoton":
        d = dict(d, watts=(200000.0 / 13650.0) * n_total)  # examples/volumetric_beambeam_lampshade.rs:139-148
    w, h, spp = d["width"], d["height"], d["num_samples"]

    def renderer(seed):
        r = make_renderer(scene, cam, d, seed=seed)
        r._ctx = ctx
        return r.shard(world, rank)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    r = renderer(0)
    r.device_scene()
    launches0 = 0
    for _ in range(args.warmup):
        film, pmap, tm = rd.photon_render(r, n_total, kind)  # full size: the stream-ordered pool reaches its working set here
        launches0 = film.stats()["kernel_launches"]
        film.close(); pmap.close()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    t0 = time.perf_counter()
    acc = dict(emit=0.0, merge=0.0, build=0.0, gather=0.0)
    stats = None
    for _ in range(args.steps):
        film, pmap, tm = rd.photon_render(r, n_total, kind)
        for k in acc:
            acc[k] += tm[k]
        stats = (tm["surface"], tm["volume"], film.stats(), pmap.stats())
        img = film.image()
        film.close(); pmap.close()
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    t = torch.tensor([wall], device=f"cuda:{local}", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    step = float(t[0]) / args.steps
    # e2e: from host scene description to host image, fresh scene every time
    barrier()
    t1 = time.perf_counter()
    r2 = renderer(1)
    film, pmap, _ = rd.photon_render(r2, n_total, kind)
    img = film.image()
    film.close(); pmap.close(); r2.device_scene().close()
    barrier()
    te = torch.tensor([time.perf_counter() - t1], device=f"cuda:{local}", dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    if rank == 0:
        peak, which = measured_peak_gbs()
        ns, nv, fst, mst = stats
        emit_s = acc["emit"] / 1e3 / args.steps
        ach = PHOTON_ROW_BYTES * (ns + nv) / world / max(emit_s, 1e-12) / 1e9
        cpu = None
        if world == 1 and args.cpu_seconds > 0:
            cpu = cpu_photons(scene, cam, r.params(), kind, args.cpu_seconds)
        line = dict(metric="Mphotons/s", value=n_total / step / 1e6, unit="Mphotons/s", n_gpus=world, steps=args.steps, warmup=args.warmup,
                    ms_per_step=1e3 * step, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64", data="synthetic",
                    config=dict(workload=f"{args.workload}: examples/volumetric_{args.workload}_lampshade.rs scene, {n_total} photons, camera pass {w}x{h} x {spp} spp, "
                                         f"gather {d['gather_size']}/{d['gather_size_volume']}",
                                photons=n_total, width=w, height=h, spp=spp, l2_policy="photon rows and maps (GBs) exceed the 126 MB L2",
                                parallelism=f"photon-index-sharded emission + all-gather, sample-sharded gather x{world}"),
                    e2e=dict(value=n_total / float(te[0]) / 1e6, unit="Mphotons/s", h2d_bytes_per_step=4096, d2h_bytes_per_step=int(w * h * 3),
                             note="scene lowering + upload + BVH build + emit + merge + map build + gather + rgb8 image to host"),
                    gpu_launches=int(fst["kernel_launches"] - launches0), clocks=clk,
                    passes_ms={k: v / args.steps for k, v in acc.items()},
                    photons=dict(surface=ns, volume=nv, bvh_items=mst["bvh_items"], gather_samples=w * h * spp * world,
                                 knn_queries=fst["gather_queries"], visibility_rays=fst["gather_visibility"], candidates=fst["gather_candidates"]),
                    roofline=dict(bound="hbm", kernel="k_photon_trace", achieved=ach, peak=peak, unit="GB/s", frac=ach / peak, traffic=None, peak_source=which,
                                  note="104-byte deposit rows per emission second; the walk is issue-bound (12-object scene in L1), see DESIGN.md"),
                    cpu_baseline=cpu)
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


def cpu_photons(scene, cam, params, kind, seconds):
    """Oracle photon_render on a bounded photon count (calibrated to ~`seconds`), all host threads."""
    import orc

    osc = orc.OracleScene(scene)
    n = 20000
    t0 = time.perf_counter()
    osc.photon_render(cam, params, n, kind)
    dt = time.perf_counter() - t0
    n2 = int(min(max(n * seconds / max(dt, 1e-6), n), 2_000_000))
    t0 = time.perf_counter()
    _, _, counts, secs = osc.photon_render(cam, params, n2, kind)
    dt = time.perf_counter() - t0
    return dict(value=n2 / dt / 1e6, unit="Mphotons/s", cores=orc.lib().orc_hardware_threads(), kind="port",
                sample=f"photon_render of {n2} photons ({counts[0]} surface + {counts[1]} volume deposits) + camera pass {params.width}x{params.height}x{params.num_samples}: "
                       f"emit {secs[0]:.2f} s, build {secs[1]:.2f} s, gather {secs[2]:.2f} s")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="fractal_teapots")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="CPU work per cpu_baseline / reference sample")
    ap.add_argument("--width", type=int, default=0, help="override (profiling runs only)")
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--photons", type=float, default=1e7, help="photons per GPU for the photon workloads")
    args = ap.parse_args()
    _OVERRIDES.update(width=args.width, height=args.height, num_samples=args.spp)
    if args.impl == "reference":
        run_reference(args)
    elif args.workload in PHOTON_KINDS:
        run_photons(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

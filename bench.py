#!/usr/bin/env python
"""bench.py — benchmark of the B200 light-transport core on the BASELINE.json configs.

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload NAME] [--scaling strong|weak]

Headline (no --workload): C3, examples/volumetric_pathtrace_lampshade.rs at 1920x1080, 1024 spp — the largest
single-GPU config and one of the two the north-star target names.  A "step" is one pass of the hot path over
one batch: Renderer::sample(spp) over the whole image (src/renderer.rs:158-171) = width*height*spp path-samples,
or one photon_render (src/photon.rs:655-720) of --photons photons for the photon workloads.
Metric: Mpath-samples/s (Mphotons/s for the photon workloads).

At N = 1 the same run also measures the other BASELINE configs (C1 cornell 512^2 x 16 spp mb5, C2 fractal_teapots
1080p x 256 spp at max_bounces 0 and 5, C4 photon-photon and C5 beam-beam at 1e8 photons), each with its own
host-CPU baseline, fewer steps, same sizes: one JSON line each on STDERR and all of them under "configs" in the
ONE stdout line.  --no-extra skips them.

N > 1 (torchrun, one rank per GPU): STRONG scaling of the stated config by default — the batch's spp (photons)
are split over the ranks (rank r traces samples s % N == r / emits its photon-index range), the per-rank f64
batch sums are all-reduced with NCCL, photon records are all-gathered.  `config.spp` is always the TOTAL.

`--impl reference` times the reference's algorithm on the host CPU cores.  rpt is Rust and cannot be built in
this image (no rustc/cargo, dependencies not vendored), so that arm runs the oracle port (oracle/, C++ f64
restatement with rpt's linear object loop, kd-trees, recursion and one task per image row) on a bounded sample
of the same workload.
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
PHOTON_KINDS = {"photonphoton": 0, "beamphoton": 1, "beambeam": 2}
PHOTON_ROW_BYTES = 104
EXAMPLE = {
    "cornell": "examples/cornell.rs", "teapot": "examples/teapot.rs", "fractal_teapots": "examples/fractal_teapots.rs",
    "fractal_teapots_mb5": "examples/fractal_teapots.rs", "lampshade": "examples/volumetric_pathtrace_lampshade.rs",
    "photonphoton": "examples/volumetric_photonphoton_lampshade.rs", "beamphoton": "examples/volumetric_beamphoton_lampshade.rs",
    "beambeam": "examples/volumetric_beambeam_lampshade.rs",
}
# the other BASELINE configs measured beside the headline at N = 1: (workload, steps, warmup)
EXTRA = [("cornell", 10, 3), ("fractal_teapots", 3, 3), ("fractal_teapots_mb5", 3, 3), ("photonphoton", 2, 2), ("beambeam", 2, 2)]

_OVERRIDES = {}


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
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def issue_peak():
    """Measured issue peak of the box's B200 (tools/peaks.cu -> profiles/r2_peaks.json): warp instructions per second
    of the mixed-pipe loop (3.53 of the 4 schedulers' slots per clock and SM)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_peaks.json")) as f:
            p = json.load(f)
        return float(p["issue_mixed"]["warp_inst_per_s"]), "profiles/r2_peaks.json issue_mixed (tools/peaks.cu)"
    except Exception:
        return 148 * 4 * 1.965e9, "fallback: 4 warp instructions per clock and SM at 1965 MHz"


def issue_constants(name):
    """Warp instructions and live lanes per traced vertex of the workload's dominant kernel, from the committed ncu capture of
    this build (profiles/r2_issue.json, written by tools/issue_probe.py from `ncu --metrics smsp__inst_executed.sum,
    smsp__thread_inst_executed.sum`)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r2_issue.json")) as f:
            return json.load(f).get(name)
    except Exception:
        return None


def cpu_sample(scene, cam, params, seconds, threads=0):
    """Time the oracle on every k-th image row of the same workload for ~`seconds`."""
    import orc

    osc = orc.OracleScene(scene)
    of = orc.OracleFilm(osc, cam, params, threads)
    h, w = params.height, params.width
    cores = threads or orc.lib().orc_hardware_threads()
    stride = max(1, h // 64)  # calibrate on 1 spp over ~64 spread rows
    sec, ctr = of.time_rows(0, h, stride, 1)
    rows = len(range(0, h, stride))
    rate = rows * w / max(sec, 1e-9)  # path-samples/s
    # ~`seconds` of CPU work: the workload's spp over as many spread rows as that buys; a workload smaller than the budget
    # (cornell at 16 spp is 0.8 s of CPU) is timed whole, with more spp per pixel (the samples are i.i.d.: same cost each)
    target = seconds * rate
    spp = max(1, params.num_samples)
    want_rows = target / (w * spp)
    if want_rows >= h:
        stride, spp = 1, int(max(spp, target / (h * w)))
    elif want_rows >= 64:
        stride = max(1, int(h / want_rows))
    else:  # keep ~64 rows spread over the image and cut the samples per pixel instead
        stride = max(1, h // 64)
        spp = int(max(1, target / (len(range(0, h, stride)) * w)))
    rows = len(range(0, h, stride))
    sec, ctr = of.time_rows(0, h, stride, spp)
    paths = rows * w * spp
    return dict(value=paths / sec / 1e6, unit=METRIC, cores=cores, kind="port",
                sample=f"{rows} of {h} rows (every {stride}th) x {w} px x {spp} spp = {paths} path-samples in {sec:.2f} s; "
                       f"{(ctr[0] + ctr[1]) / sec / 1e6:.2f} Mrays/s"), sec


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


def config_of(name, d, world, scaling):
    medium = name == "lampshade"
    return dict(workload=f"{name}: {EXAMPLE[name]} scene, {d['width']}x{d['height']}, {d['num_samples']} spp, "
                         + ("Russian roulette 0.8 (medium branch ignores max_bounces)" if medium else f"max_bounces {d['max_bounces']}"),
                width=d["width"], height=d["height"], spp=d["num_samples"], max_bounces=None if medium else d["max_bounces"],
                l2_policy="inputs larger than L2 between steps: every step traces fresh sample indices (new RNG streams, no cached result); the path state "
                          "lives in registers (small scenes) or in a multi-GB pool (BVH scenes); no explicit flush",
                parallelism=f"sample-sharded x{world} ({scaling}: {'total spp fixed' if scaling == 'strong' else 'spp per GPU fixed'})")


def photon_config_of(name, d, n_total, world, scaling):
    return dict(workload=f"{name}: {EXAMPLE[name]} scene, {n_total} photons, camera pass {d['width']}x{d['height']} x {d['num_samples']} spp, "
                         f"gather {d['gather_size']}/{d['gather_size_volume']}",
                photons=n_total, width=d["width"], height=d["height"], spp=d["num_samples"],
                l2_policy="photon rows and maps (GBs) exceed the 126 MB L2; fresh photon seeds every step",
                parallelism=f"photon-index-sharded emission + all-gather, sample-sharded gather x{world} ({scaling})")


def dist_env():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


class Box:
    """Process-wide state of one bench run: device, context, process group."""

    def __init__(self):
        import torch

        self.torch = torch
        self.rank, self.world, self.local = dist_env()
        self.dist = None
        torch.cuda.set_device(self.local)
        if self.world > 1:
            import torch.distributed as dist

            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        from rpt_b200 import Context
        from rpt_b200 import dist as rd

        self.ctx = Context(self.local)
        if self.world > 1 and not os.environ.get("RPTX_TORCH_COLLECTIVES"):
            rd.init_comm(self.ctx)  # the library's own NCCL communicator: the exchanges run inside librptx.so

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, *vals):
        t = self.torch.tensor(list(vals), device=f"cuda:{self.local}", dtype=self.torch.float64)
        if self.dist is not None:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(x) for x in t]

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def run_path(box, name, steps, warmup, scaling, cpu_seconds):
    """One path-tracing workload -> the JSON line (rank 0) or None."""
    from rpt_b200 import dist as rd
    from rpt_b200.scenes import make_renderer

    rank, world = box.rank, box.world
    scene, cam, d = workload(name)
    w, h, spp = d["width"], d["height"], d["num_samples"]
    batch = spp if scaling == "strong" else spp * world  # samples per pixel of one step, all ranks together
    r = make_renderer(scene, cam, d, seed=0)
    r._ctx = box.ctx
    r.shard(world, rank)
    film = r.begin()

    def step():
        rd.sample_batch(film, batch)  # trace the local shard -> NCCL all-reduce of the f64 batch sums -> commit
        return film.stats()

    for _ in range(warmup):
        step()
    st0 = film.stats()
    clocks = ClockSampler(box.local)
    if rank == 0:
        clocks.start()
    box.barrier()
    acc = dict(device=0.0, extend=0.0, shade=0.0, shadow=0.0, connect=0.0)
    t0 = time.perf_counter()
    for _ in range(steps):
        st = step()
        acc["device"] += st["device_ms"]  # CUDA events on the library's own stream around the traced batch
        for k in ("extend", "shade", "shadow", "connect"):
            acc[k] += st[k + "_ms"]
    box.barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    st1 = film.stats()
    wall_max, dev_max = box.max_over_ranks(wall, acc["device"] / 1e3)
    paths = w * h * batch * steps
    value = paths / max(wall_max, 1e-12) / 1e6
    # counters of this rank over the timed steps
    rays_e = st1["extend_rays"] - st0["extend_rays"]
    rays_s = st1["shadow_rays"] - st0["shadow_rays"]
    launches = st1["kernel_launches"] - st0["kernel_launches"]
    mode = {0: "wavefront", 1: "fused", 3: "persistent"}.get(st1["path_mode"], str(st1["path_mode"]))

    # e2e: the user-facing call from HOST data: lower + upload scene, BVH build, path state, render, rgb8 image to host
    h2d = sum(v.nbytes + (0 if n is None else n.nbytes) for v, n in r.device_scene().low.meshes) + 4096
    e2e_t, n_e2e = 0.0, max(1, min(steps, 3))
    film.close()  # its slab returns to the context, as between a user's successive render() calls
    for it in range(n_e2e + 1):  # pass 0 is the warm-up of this path (pool growth after the trim between configs): not timed
        box.barrier()
        t1 = time.perf_counter()
        rr = make_renderer(scene, cam, d, seed=1)
        rr._ctx = box.ctx
        rr.shard(world, rank)
        rr.device_scene()
        t2 = time.perf_counter()
        f2 = rr.begin()
        t3 = time.perf_counter()
        rd.sample_batch(f2, batch)
        t4 = time.perf_counter()
        img = f2.image()
        box.barrier()
        t5 = time.perf_counter()
        e2e_t += (t5 - t1) if it else 0.0
        assert img.shape == (h, w, 3)
        f2.close(); rr.device_scene().close()
        if rank == 0 and it:
            print("e2e ms [%s]: scene %.1f begin %.1f sample %.1f image %.1f" % (name, 1e3 * (t2 - t1), 1e3 * (t3 - t2), 1e3 * (t4 - t3), 1e3 * (t5 - t4)), file=sys.stderr)
    (e2e_max,) = box.max_over_ranks(e2e_t / n_e2e)
    e2e_val = w * h * batch / e2e_max / 1e6
    r.device_scene().close()
    if rank != 0:
        return None

    # roofline of the dominant kernel (by CUDA-event time of THIS run).  Two candidate roofs, both measured on the box:
    #   issue: the SMs' instruction issue rate (tools/peaks.cu -> profiles/r2_peaks.json); the ray kernels keep the scene in
    #          L1/L2 and are bound here.  achieved = warp instructions/s x live lanes / 32, from the vertex count of this run
    #          x the per-vertex instruction counts of the committed ncu capture of this build (profiles/r2_issue.json);
    #   hbm:   algorithmic bytes / time against the measured copy bandwidth; the state kernels of the
    #          wavefront (k_shade, k_connect) are bound here.
    # `bound` is the roof the kernel is closer to.
    vertices = rays_e  # one closest-hit query per path vertex
    ipk, ipk_src = issue_peak()
    hbm_peak, hbm_src = measured_peak_gbs()
    waves = max(st1["waves"] - st0["waves"], 1)
    alg_bytes = {"k_extend": 88, "k_shade": 280, "k_shade_final": 217, "k_shadow": 104, "k_connect": 185, "k_path_persist": 0}
    if mode == "wavefront":
        part = max(("extend", "shade", "shadow", "connect"), key=lambda k: acc[k])
        kname = "k_shade_final" if (part == "shade" and d["max_bounces"] == 0 and name != "lampshade") else "k_" + part
        dom_s = acc[part] / 1e3
        ic = issue_constants(name + ":wavefront")
        ick = (ic or {}).get("kernels", {}).get(kname) if ic else None
        units = rays_s if part == "shadow" else vertices
    else:
        kname, dom_s, units = ("k_path_persist" if mode == "persistent" else "k_path_fused"), acc["device"] / 1e3, vertices
        ic = issue_constants(name)
        ick = ic
    roof = dict(kernel=kname, launches=int(waves), vertices_per_launch=vertices / waves, share_of_step=dom_s / max(acc["device"] / 1e3, 1e-12))
    hbm = dict(achieved=alg_bytes.get(kname, 0) * units / max(dom_s, 1e-12) / 1e9, peak=hbm_peak, unit="GB/s", peak_source=hbm_src,
               algorithmic_bytes_per_vertex=alg_bytes.get(kname, 0))
    hbm["frac"] = hbm["achieved"] / hbm_peak
    issue = dict(peak=ipk / 1e9, unit="Gwarp-inst/s", peak_source=ipk_src)
    if ick:
        winst = ick["warp_inst_per_vertex"] * vertices
        issue.update(achieved=winst * (ick["live_lanes"] / 32.0) / max(dom_s, 1e-12) / 1e9, issue_slots_frac=winst / max(dom_s, 1e-12) / ipk,
                     live_lanes=ick["live_lanes"], warp_inst_per_vertex=ick["warp_inst_per_vertex"], capture=(ic or {}).get("capture"))
        issue["frac"] = issue["achieved"] / issue["peak"]
        traffic = ick.get("dram_bytes_per_vertex")
        roof["traffic"] = traffic * vertices / waves if traffic is not None else ick.get("dram_bytes_per_launch")
    else:
        issue.update(achieved=None, frac=None, note="no ncu capture of this kernel committed (profiles/r2_issue.json)")
        roof["traffic"] = None
    pick = issue if (issue.get("frac") or 0.0) >= hbm["frac"] else hbm
    roof.update(bound="issue" if pick is issue else "hbm", achieved=pick["achieved"], peak=pick["peak"], unit=pick["unit"], frac=pick["frac"],
                peak_source=pick["peak_source"], issue=issue, hbm=hbm)
    cpu = None
    if world == 1 and cpu_seconds > 0:
        cpu, _ = cpu_sample(scene, cam, r.params(), cpu_seconds)
    line = dict(metric=METRIC, value=value, unit=METRIC, n_gpus=world, steps=steps, warmup=warmup, ms_per_step=1e3 * wall_max / steps,
                higher_is_better=True, scaling=scaling, vs_baseline=None, dtype="f64", data="synthetic", config=config_of(name, d, world, scaling),
                e2e=dict(value=e2e_val, unit=METRIC, h2d_bytes_per_step=int(h2d), d2h_bytes_per_step=int(w * h * 3),
                         note="Renderer -> scene lowering + upload + BVH build + path state + render + rgb8 image to host"),
                gpu_launches=int(launches), clocks=clk, path_kernels=mode, roofline=roof,
                rays_per_s=dict(extend=rays_e * world / max(wall_max, 1e-12), shadow_queries=rays_s * world / max(wall_max, 1e-12),
                                vertices_per_path=rays_e / max(st1["path_samples"] - st0["path_samples"], 1)),
                kernel_ms_per_step=dict({k: v / steps for k, v in acc.items()}, wall=1e3 * wall / steps),
                cpu_baseline=cpu, rmse=rmse_probe(box, name))
    if cpu:
        line["x_cpu"] = dict(value=value / cpu["value"], e2e=e2e_val / cpu["value"], note=f"same-run host CPU port, {cpu['cores']} threads")
    return line


def rmse_probe(box, name):
    """Image RMSE against the oracle's converged render (tests/golden/converged_*.npz, made by tests/golden/make_converged.py):
    the same scene at the fixture's reduced size, GPU at the fixture's spp with another seed; relRMSE = sqrt(mean((I-R)^2)) / mean(R)
    on the linear mean image."""
    import numpy as np

    from rpt_b200.scenes import make_renderer

    path = os.path.join(ROOT, "tests", "golden", f"converged_{name}.npz")
    if not os.path.exists(path) or box.world != 1:
        return None
    z = np.load(path)
    ref = z["mean"]
    hh, ww = ref.shape[:2]
    scene, cam, d = workload(name)
    over = dict(width=ww, height=hh, num_samples=int(z["spp"]), seed=12345)
    out = {}
    if name in PHOTON_KINDS:
        n = int(z["photons"])
        if name != "photonphoton":
            over["watts"] = (200000.0 / 13650.0) * n
        r = make_renderer(scene, cam, d, **over)
        r._ctx = box.ctx
        _, mean = r.photon_render(n, PHOTON_KINDS[name], want_mean=True)
        out["photons"] = n
    else:
        r = make_renderer(scene, cam, d, **over)
        r._ctx = box.ctx
        f = r.begin()
        f.sample(int(z["spp"]))
        mean = f.mean()
        f.close()
    r.device_scene().close()
    rel = float(np.sqrt(np.mean((mean - ref) ** 2)) / np.mean(ref))
    out.update(rel_rmse=rel, size=f"{ww}x{hh}", spp=int(z["spp"]), ref_spp=int(z["spp"]), reference="oracle port, tests/golden/" + os.path.basename(path),
               expected_mc=float(z["rel_rmse_mc"]) if "rel_rmse_mc" in z else None, mean_ratio=float(np.mean(mean) / np.mean(ref)))
    return out


def run_photons(box, name, steps, warmup, scaling, cpu_seconds, photons):
    from rpt_b200 import dist as rd
    from rpt_b200.scenes import make_renderer

    rank, world = box.rank, box.world
    kind = PHOTON_KINDS[name]
    scene, cam, d = workload(name)
    n_total = int(photons) if scaling == "strong" else int(photons) * world
    if name != "photonphoton":
        d = dict(d, watts=(200000.0 / 13650.0) * n_total)  # examples/volumetric_beambeam_lampshade.rs:139-148
    w, h, spp = d["width"], d["height"], d["num_samples"]

    def renderer(seed):
        r = make_renderer(scene, cam, d, seed=seed)
        r._ctx = box.ctx
        return r.shard(world, rank)

    r = renderer(0)
    r.device_scene()
    launches0 = 0
    for i in range(warmup):
        r._set("seed", 1000 + i)
        film, pmap, tm = rd.photon_render(r, n_total, kind)  # full size: the stream-ordered pool reaches its working set here
        launches0 = film.stats()["kernel_launches"]
        film.close(); pmap.close()
    clocks = ClockSampler(box.local)
    if rank == 0:
        clocks.start()
    box.barrier()
    t0 = time.perf_counter()
    acc = dict(emit=0.0, merge=0.0, build=0.0, gather=0.0)
    stats = None
    for i in range(steps):
        r._set("seed", i)
        film, pmap, tm = rd.photon_render(r, n_total, kind)
        for k in acc:
            acc[k] += tm[k]
        stats = (tm["surface"], tm["volume"], film.stats(), pmap.stats())
        img = film.image()
        film.close(); pmap.close()
    box.barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    (wall_max,) = box.max_over_ranks(wall)
    step = wall_max / steps
    # e2e: from host scene description to host image, fresh scene every time
    box.barrier()
    t1 = time.perf_counter()
    r2 = renderer(77)
    film, pmap, _ = rd.photon_render(r2, n_total, kind)
    img = film.image()
    film.close(); pmap.close(); r2.device_scene().close()
    box.barrier()
    (e2e_s,) = box.max_over_ranks(time.perf_counter() - t1)
    r.device_scene().close()
    if rank != 0:
        return None
    peak, which = measured_peak_gbs()
    ns, nv, fst, mst = stats
    # dominant pass and its roofline: the camera pass (kNN / beam queries over the map) or the map build move photon
    # records through HBM; algorithmic bytes per SURVEY 8(d)
    dom = max(acc, key=lambda k: acc[k])
    dom_s = acc[dom] / 1e3 / steps
    if dom == "gather":
        k_s, k_v = d["gather_size"], d["gather_size_volume"]
        import math

        node_bytes = 64 * max(1.0, math.log2(max(ns + nv, 16) / 8.0))
        alg = fst["gather_queries"] * (node_bytes + 0.5 * (k_s + k_v) * 56) + fst["gather_candidates"] * 48 + fst["gather_visibility"] * 88
        kernel = "k_photon_gather"
    elif dom == "build":
        alg = (ns + nv) * (104 + 80 + 64 + 4 * 16)  # rows read, SoA written, node, two radix sorts of (key, index)
        kernel = "build_point_map (k_rows_morton, radix sorts, k_rows_gather, build_lbvh_sorted)"
    else:
        alg = PHOTON_ROW_BYTES * (ns + nv) / world
        kernel = "k_photon_trace"
    ach = alg / max(dom_s, 1e-12) / 1e9
    roof = dict(bound="hbm", kernel=kernel, achieved=ach, peak=peak, unit="GB/s", frac=ach / peak, traffic=None, peak_source=which,
                algorithmic_bytes_per_launch=alg, note=f"dominant pass of the step: {dom} ({1e3 * dom_s:.1f} ms of {1e3 * step:.1f})")
    # the walk and the camera pass are issue-bound traversal kernels: with the ncu constants of this build at this size
    # (tools/issue_probe.py <workload> --photons N -> profiles/r2_issue.json "<workload>:photon") report the issue roofline too
    ic = (issue_constants(name + ":photon") or {})
    kc = ic.get("kernels", {}).get({"gather": "k_photon_gather", "emit": "k_photon_trace"}.get(dom, ""))
    if kc and abs(ic.get("photons", 0) - n_total) <= 0.01 * n_total:
        ipeak, isrc = issue_peak()
        units = (w * h * spp if dom == "gather" else n_total) / world
        iach = kc["warp_inst_per_unit"] * units * (kc["live_lanes"] / 32.0) / max(dom_s, 1e-12)
        roof["hbm"] = dict(achieved=ach, peak=peak, unit="GB/s", frac=ach / peak, algorithmic_bytes_per_launch=alg)
        roof["issue"] = dict(achieved=iach / 1e9, peak=ipeak / 1e9, unit="Gwarp-inst/s", frac=iach / ipeak, peak_source=isrc,
                             warp_inst_per_unit=kc["warp_inst_per_unit"], live_lanes=kc["live_lanes"],
                             unit_of_work="gather sample" if dom == "gather" else "emitted photon", capture=ic.get("capture"))
        roof["traffic"] = kc["dram_bytes_per_unit"] * units
        if iach / ipeak > ach / peak:
            roof.update(bound="issue", achieved=iach / 1e9, peak=ipeak / 1e9, unit="Gwarp-inst/s", frac=iach / ipeak, peak_source=isrc)
    cpu = None
    if world == 1 and cpu_seconds > 0:
        cpu = cpu_photons(scene, cam, r.params(), kind, cpu_seconds)
    line = dict(metric="Mphotons/s", value=n_total / step / 1e6, unit="Mphotons/s", n_gpus=world, steps=steps, warmup=warmup,
                ms_per_step=1e3 * step, higher_is_better=True, scaling=scaling, vs_baseline=None, dtype="f64", data="synthetic",
                config=photon_config_of(name, d, n_total, world, scaling),
                e2e=dict(value=n_total / e2e_s / 1e6, unit="Mphotons/s", h2d_bytes_per_step=4096, d2h_bytes_per_step=int(w * h * 3),
                         note="scene lowering + upload + BVH build + emit + merge + map build + gather + rgb8 image to host"),
                gpu_launches=int(fst["kernel_launches"] - launches0), clocks=clk,
                passes_ms={k: v / steps for k, v in acc.items()},
                photons=dict(surface=ns, volume=nv, bvh_items=mst["bvh_items"], gather_samples=w * h * spp,
                             knn_queries=fst["gather_queries"], visibility_rays=fst["gather_visibility"], candidates=fst["gather_candidates"]),
                roofline=roof, cpu_baseline=cpu, rmse=rmse_probe(box, name))
    if cpu:
        line["x_cpu"] = dict(value=line["value"] / cpu["value"], e2e=line["e2e"]["value"] / cpu["value"], note=f"same-run host CPU port, {cpu['cores']} threads")
    return line


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from rpt_b200.scenes import make_renderer

    name = args.workload
    scene, cam, d = workload(name)
    if name in PHOTON_KINDS:
        n_total = int(args.photons) if args.scaling == "strong" else int(args.photons) * args.gpus
        if name != "photonphoton":
            d = dict(d, watts=(200000.0 / 13650.0) * n_total)
        params = make_renderer(scene, cam, d).params()
        kind = PHOTON_KINDS[name]
        t0 = time.perf_counter()
        runs = [cpu_photons(scene, cam, params, kind, args.cpu_seconds) for _ in range(max(args.steps, 1))]
        secs = time.perf_counter() - t0
        last = runs[-1]
        v = sum(x["value"] for x in runs) / len(runs)
        last["value"] = v
        print(json.dumps(dict(impl="reference", metric="Mphotons/s", value=v, unit="Mphotons/s", n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                              ms_per_step=1e3 * secs / max(args.steps, 1), higher_is_better=True, scaling=args.scaling, vs_baseline=None, dtype="f64",
                              data="synthetic", config=photon_config_of(name, d, n_total, args.gpus, args.scaling),
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
                ms_per_step=1e3 * secs / args.steps, higher_is_better=True, scaling=args.scaling, vs_baseline=None, dtype="f64", data="synthetic",
                config=config_of(name, d, args.gpus, args.scaling), cpu_baseline=last,
                e2e=dict(value=v, unit=METRIC, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    print(json.dumps(line))


def run_one(box, name, steps, warmup, args):
    if name in PHOTON_KINDS:
        return run_photons(box, name, steps, warmup, args.scaling, args.cpu_seconds, args.photons)
    return run_path(box, name, steps, warmup, args.scaling, args.cpu_seconds)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="lampshade")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="CPU work per cpu_baseline / reference sample")
    ap.add_argument("--width", type=int, default=0, help="override (profiling runs only)")
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--photons", type=float, default=1e8, help="photons of one photon_render (total under strong scaling, per GPU under weak)")
    ap.add_argument("--no-extra", action="store_true", help="headline workload only")
    args = ap.parse_args()
    _OVERRIDES.update(width=args.width, height=args.height, num_samples=args.spp)
    if args.impl == "reference":
        run_reference(args)
        return
    box = Box()
    line = run_one(box, args.workload, args.steps, args.warmup, args)
    overridden = any(_OVERRIDES.values())
    if box.world == 1 and not args.no_extra and args.workload == "lampshade" and not overridden:
        extras = []
        for name, st, wu in EXTRA:
            try:
                box.ctx.trim()  # each config starts from an empty device: the previous one's cached GBs are not its problem
                x = run_one(box, name, st, wu, args)
            except Exception as e:  # an extra config must not take the headline down
                x = dict(config=dict(workload=name), error=repr(e))
            print(json.dumps(x), file=sys.stderr)
            extras.append(x)
        line["configs"] = extras
    if box.rank == 0:
        print(json.dumps(line))
    box.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Measures, on the GPU box, the per-vertex instruction counts that bench.py's issue roofline uses:

    python tools/issue_probe.py <workload> [bench args...]      (under gpurun; merges into gpurun_out/r2_issue.json)

Runs `bench.py --workload W --steps 1 --warmup 0 --no-extra --cpu-seconds 0 ...` once plainly (for its vertex count) and
once under `ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum`
restricted to the workload's dominant kernel (first launch = the timed step), and writes
  warp_inst_per_vertex = smsp__inst_executed.sum / vertices of that launch
  live_lanes           = smsp__thread_inst_executed.sum / smsp__inst_executed.sum
  dram_bytes_per_launch
Copy the result to profiles/r2_issue.json (committed); bench.py multiplies these constants by the vertex count it
measures live."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out", "r2_issue.json")


PHOTON = ("photonphoton", "beamphoton", "beambeam")


def photon_main(name, extra):
    """Photon workloads: warp instructions, live lanes and DRAM bytes of the emission walk per emitted photon and of the camera
    pass per gather sample, AT the size given (the camera pass's cost per sample grows with the map) -> key `<name>:photon`."""
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", name, "--steps", "1", "--warmup", "0", "--no-extra", "--cpu-seconds", "0"] + extra
    plain = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
    line = json.loads(plain.stdout.strip().splitlines()[-1])
    units = {"k_photon_trace": float(line["config"]["photons"]), "k_photon_gather": float(line["photons"]["gather_samples"])}
    log = os.path.join(ROOT, "gpurun_out", f"issue_{name}.csv")
    metrics = "smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
    subprocess.run(["ncu", "--metrics", metrics, "--clock-control", "none", "-k", "regex:k_photon_trace|k_photon_gather", "-c", "8", "--csv", "--log-file", log] + cmd,
                   capture_output=True, text=True, cwd=ROOT)
    launches = {}  # (kernel, id) -> metrics
    for r in csv.DictReader(l for l in open(log) if l.startswith('"')):
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except (KeyError, ValueError):
            continue
        unit = r.get("Metric Unit", "")
        if r["Metric Name"].startswith("dram__bytes"):
            v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(unit, 1.0)
        if r["Metric Name"].startswith("gpu__time"):
            v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "msecond": 1.0, "ms": 1.0, "second": 1e3}.get(unit, 1e-6)
        base = "k_photon_gather" if "k_photon_gather" in r["Kernel Name"] else "k_photon_trace"
        launches.setdefault((base, r["ID"]), {})[r["Metric Name"]] = v
    kernels = {}
    for base in ("k_photon_trace", "k_photon_gather"):
        cand = [m for (b, _), m in launches.items() if b == base]
        if not cand:
            continue
        m = max(cand, key=lambda x: x.get("gpu__time_duration.sum", 0.0))  # the pilot / warm-up launches are the short ones
        winst, tinst = m["smsp__inst_executed.sum"], m["smsp__thread_inst_executed.sum"]
        kernels[base] = dict(warp_inst_per_unit=winst / units[base], live_lanes=tinst / winst, units=units[base],
                             dram_bytes_per_unit=(m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0)) / units[base],
                             ms=m.get("gpu__time_duration.sum", 0.0))
    rec = dict(kernels=kernels, photons=units["k_photon_trace"],
               capture=f"ncu --metrics {metrics} -k regex:k_photon_trace|k_photon_gather of `bench.py --workload {name} {' '.join(extra)}` (tools/issue_probe.py); "
                       "units: emitted photons (k_photon_trace), gather samples (k_photon_gather*)")
    seed = OUT if os.path.exists(OUT) else os.path.join(ROOT, "profiles", "r2_issue.json")
    data = json.load(open(seed)) if os.path.exists(seed) else {}
    data[name + ":photon"] = rec
    json.dump(data, open(OUT, "w"), indent=1)
    print(name + ":photon", json.dumps(rec))


def main():
    name, extra = sys.argv[1], sys.argv[2:]
    if name in PHOTON:
        return photon_main(name, extra)
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--workload", name, "--steps", "1", "--warmup", "0", "--no-extra", "--cpu-seconds", "0"] + extra
    plain = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT)
    line = json.loads(plain.stdout.strip().splitlines()[-1])
    mode = line["path_kernels"]
    kernel = "k_path_persist" if mode == "persistent" else "k_extend|k_shade|k_shadow|k_connect"
    vertices = line["roofline"]["vertices_per_launch"]
    log = os.path.join(ROOT, "gpurun_out", f"issue_{name}.csv")
    metrics = "smsp__inst_executed.sum,smsp__thread_inst_executed.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum"
    n_waves = 1 if mode == "persistent" else line["roofline"]["launches"]
    # the timed step's kernels come first (warmup 0): one launch (persistent), or <= 4 kernels per wave (wavefront)
    n_launch = 1 if mode == "persistent" else 4 * n_waves
    subprocess.run(["ncu", "--metrics", metrics, "--clock-control", "none", "-k", f"regex:{kernel}", "-c", str(n_launch), "--csv", "--log-file", log] + cmd,
                   capture_output=True, text=True, cwd=ROOT)
    per = {}  # kernel base name -> metric sums
    for r in csv.DictReader(l for l in open(log) if l.startswith('"')):
        try:
            v = float(r["Metric Value"].replace(",", ""))
        except (KeyError, ValueError):
            continue
        unit = r.get("Metric Unit", "")
        if r["Metric Name"].startswith("dram__bytes"):
            v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0}.get(unit, 1.0)
        if r["Metric Name"].startswith("gpu__time"):
            v *= {"ns": 1e-6, "us": 1e-3, "usecond": 1e-3, "nsecond": 1e-6, "msecond": 1.0, "ms": 1.0, "second": 1e3}.get(unit, 1e-6)
        base = r["Kernel Name"].split("<")[0].split("(")[0].replace("void ", "").replace("rptx::", "")
        base = "k_shadow" if base.startswith("k_shadow") else base
        d = per.setdefault(base, {})
        d[r["Metric Name"]] = d.get(r["Metric Name"], 0.0) + v
        if r["Metric Name"] == "gpu__time_duration.sum":
            d["launches"] = d.get("launches", 0) + 1
    total_vertices = vertices * n_waves
    key = name if mode != "wavefront" else name + ":wavefront"
    kernels = {}
    for base, d in per.items():
        winst, tinst = d["smsp__inst_executed.sum"], d["smsp__thread_inst_executed.sum"]
        kv = vertices * max(d.get("launches", 1), 1)  # the capture may span more than one step (fewer than 4 kernels per wave)
        kernels[base] = dict(warp_inst_per_vertex=winst / kv, live_lanes=tinst / winst,
                             dram_bytes_per_vertex=(d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)) / kv,
                             ms=d.get("gpu__time_duration.sum", 0.0), launches=d.get("launches", 0))
    dom = max(kernels, key=lambda k: kernels[k]["ms"])
    rec = dict(kernel=dom, warp_inst_per_vertex=kernels[dom]["warp_inst_per_vertex"], live_lanes=kernels[dom]["live_lanes"],
               dram_bytes_per_launch=kernels[dom]["dram_bytes_per_vertex"] * total_vertices / max(kernels[dom]["launches"], 1),
               dram_bytes_per_vertex=kernels[dom]["dram_bytes_per_vertex"], kernels=kernels,
               vertices=total_vertices, waves=n_waves,
               capture=f"ncu --metrics {metrics} -k regex:{kernel} of `bench.py --workload {name} {' '.join(extra)}` (tools/issue_probe.py)")
    data = {}
    seed = OUT if os.path.exists(OUT) else os.path.join(ROOT, "profiles", "r2_issue.json")  # a fresh box starts from the committed file
    if os.path.exists(seed):
        data = json.load(open(seed))
    data[key] = rec
    json.dump(data, open(OUT, "w"), indent=1)
    print(key, json.dumps(rec))


if __name__ == "__main__":
    main()

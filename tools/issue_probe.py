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


def main():
    name, extra = sys.argv[1], sys.argv[2:]
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

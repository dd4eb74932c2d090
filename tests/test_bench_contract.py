"""CPU: the bench.py contract.  The reference arm (`--impl reference`: the oracle port on the host cores) runs here without a
GPU and must print the ONE JSON line with the contract's keys; the committed records of the GPU arm (profiles/) must carry
the same keys plus `roofline`, `cpu_baseline`, `e2e`, `gpu_launches`, `clocks`, and the constants bench.py derives its
rooflines from (profiles/r2_peaks.json, profiles/r2_issue.json) must hold what it reads."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
             "data", "config", "e2e", "cpu_baseline"}


def _last_json(text):
    lines = [l for l in text.strip().splitlines() if l.startswith("{")]
    assert len(lines) == 1, "exactly one JSON line on stdout"
    return json.loads(lines[0])


def test_reference_arm_prints_the_contract_line():
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-seconds", "0.3"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    d = _last_json(res.stdout)
    assert BASE_KEYS <= set(d), BASE_KEYS - set(d)
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["metric"] == d["unit"] == "Mpath-samples/s" and d["value"] > 0
    assert "lampshade" in d["config"]["workload"] and d["config"]["spp"] == 1024 and d["config"]["width"] == 1920
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "rows" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_runs_on_rank_0_only():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                          "--cpu-seconds", "0.3"], capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert res.returncode == 0 and res.stdout.strip() == ""


def test_committed_gpu_record_has_the_contract_keys():
    d = _last_json(open(os.path.join(ROOT, "profiles", "r2_bench_default_1gpu.json")).read())
    assert BASE_KEYS | {"roofline", "gpu_launches", "clocks"} <= set(d)
    assert d["config"]["workload"].startswith("lampshade") and d["config"]["spp"] == 1024 and d["scaling"] == "strong"
    r = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r) and r["bound"] in ("hbm", "issue")
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and 0 < r["frac"] < 1.2
    assert d["gpu_launches"] > 0 and d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] == 1920 * 1080 * 3
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    names = [c["config"]["workload"].split(":")[0] for c in d["configs"]]
    assert names == ["cornell", "fractal_teapots", "fractal_teapots_mb5", "photonphoton", "beambeam"]
    for c in d["configs"]:
        assert "error" not in c and c["value"] > 0 and "rmse" in c and "roofline" in c and c["cpu_baseline"]["value"] > 0


def test_roofline_constants_hold_what_bench_reads():
    sys.path.insert(0, ROOT)
    import bench

    peak, src = bench.issue_peak()
    assert 5e11 < peak < 2e12 and "r2_peaks.json" in src
    for key in ("lampshade:wavefront", "cornell:wavefront", "fractal_teapots:wavefront", "fractal_teapots_mb5:wavefront"):
        c = bench.issue_constants(key)
        assert c and "k_extend" in c["kernels"] and any(k.startswith("k_shade") for k in c["kernels"])  # k_shade or k_shade_final
        for k in c["kernels"].values():
            assert k["warp_inst_per_vertex"] > 0 and 1.0 <= k["live_lanes"] <= 32.0
    for key in ("photonphoton:photon", "beambeam:photon"):
        c = bench.issue_constants(key)
        assert c and {"k_photon_trace", "k_photon_gather"} <= set(c["kernels"]) and c["photons"] == 1e8

"""CPU: the JSON contract of bench.py — the reference arm run here on a small frame (it needs no GPU: it times the
reference shader compiled as C++, or the oracle port), and the newest recorded B200 line under profiles/ for the keys
the driver and the judge read (value, e2e, roofline, cpu_baseline, clocks, gpu_launches)."""
import glob
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
        "dtype", "data", "config", "e2e"]


def _check_base(line):
    for k in BASE:
        assert k in line, k
    assert isinstance(line["config"], dict) and "workload" in line["config"] and "model" not in line["config"]
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in line["e2e"], k
    assert line["unit"] == "Mrays/s" and line["higher_is_better"] is True and line["dtype"] == "f32"
    assert line["vs_baseline"] is None            # BASELINE.md holds no published number for this metric


def test_reference_arm_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--width", "160", "--height", "120",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    _check_base(line)
    assert line["impl"] == "reference" and line["steps"] == 2 and line["warmup"] == 1
    cb = line["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == line["value"] and cb["sample"]
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["value"] > 0 and line["ms_per_step"] > 0


def test_recorded_b200_line():
    files = glob.glob(os.path.join(ROOT, "profiles", "r*", "bench_v*_n1.json"))
    assert files, "no recorded single-GPU bench line under profiles/"
    newest = max(files, key=lambda f: int(re.search(r"bench_v(\d+)_", f).group(1)))
    line = json.loads(open(newest).read().strip().splitlines()[-1])
    _check_base(line)
    assert line["n_gpus"] == 1 and line["data"] == "synthetic" and line["gpu_launches"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] > 0 and line["e2e"]["d2h_bytes_per_step"] == 16 * 3840 * 2160
    rf = line["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in rf, k
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    if int(re.search(r"bench_v(\d+)_", newest).group(1)) >= 18:
        # round 2: the kernel is issue-bound and the scene cache-resident — the roofline is the issue roofline (a fraction of
        # a physical peak, <= 1 by construction); every memory level is reported as measured traffic / measured peak beside it
        assert rf["bound"] == "issue" and 0 < rf["frac"] <= 1.2 and rf["unit"] == "Gwarp-inst/s"
        mem = rf["memory"]
        assert mem["l2_peak_GBps"] and mem["l1_peak_GBps"] and 0 < mem["l2_frac"] < 1 and 0 < mem["l1_frac"] < 1 and 0 < rf["hbm_frac"] < 1
        assert "over_hbm_peak" in rf["alg_normalised"]
        par = line["parity"]
        assert par["ids_mismatch"] == 0 and par["frac_within_1e-3"] >= 0.999 and par["tie_pixels"] >= 0
        assert "cold_topology_ms" in line["e2e"] and "note" in line["e2e"]
    else:
        assert rf["bound"] in ("hbm", "tensor")
    cb = line["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in cb, k
    ck = line["clocks"]
    assert ck["sm_mhz"] and ck["sm_max_mhz"] and not set(ck["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert line["rays_match_oracle"] is True      # the kernel traced exactly the reference's rays

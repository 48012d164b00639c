"""The committed bench lines (profiles/r01_bench_*_n.json, written by bench.py on a B200) carry every
key the measurement contract names, with consistent values; bench.py exposes the contract's flags."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load(name):
    with open(os.path.join(ROOT, "profiles", name)) as f:
        return json.loads(f.read().strip().splitlines()[-1])


def test_native_line_contract():
    d = load("r01_bench_native_n.json")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["data"] == "synthetic" and d["dtype"] == "f32" and "workload" in d["config"] and "model" not in d["config"]
    # value is whole-job throughput: 8 view iterations per step
    assert abs(d["value"] - 8 * 1e3 / d["ms_per_step"]) <= 1e-3 * d["value"]
    e = d["e2e"]
    assert e["unit"] == d["unit"] and 0 < e["value"] <= d["value"] * 1.02
    assert e["h2d_bytes_per_step"] == 8 * (3 + 1) * 1200 * 680 * 4 and e["d2h_bytes_per_step"] > 0
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-3 and 0 < r["frac"] <= 1.0
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["value"] > 0 and c["sample"]
    assert d["gpu_launches"] > 0
    k = d["clocks"]
    assert k["sm_mhz"] > 0.9 * k["sm_max_mhz"]
    assert not set(k["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}


def test_reference_line_contract():
    n, r = load("r01_bench_native_n.json"), load("r01_bench_reference_n.json")
    assert r["impl"] == "reference"
    for k in ("metric", "unit", "higher_is_better", "n_gpus", "steps", "warmup"):
        assert r[k] == n[k], k
    assert r["config"]["workload"] == n["config"]["workload"]
    assert r["cpu_baseline"]["kind"] == "reference" and r["e2e"]["value"] > 0
    assert n["value"] > 3 * r["value"]            # north_star: >= 3x the reference CUDA rasterizer


def test_bench_flags():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True)
    assert out.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in out.stdout

"""The JSON contract of bench.py (driver-facing): keys of the committed N=1 line, and the reference arm run here on the
host cores (it needs no GPU: it times the reference's CPU implementation of the path)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config", "e2e", "gpu_launches"}


def test_committed_bench_line_carries_the_whole_contract():
    d = json.loads(open(os.path.join(ROOT, "profiles", "r1_bench_n1.json")).readline())
    assert BASE_KEYS | {"roofline", "cpu_baseline", "clocks"} <= set(d)
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert d["metric"] == "KawPow MH/s" and d["unit"] == "MH/s" and d["higher_is_better"] is True and d["scaling"] == "weak"
    assert d["metric"].split()[0].lower() in json.dumps(base).lower()
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["steps"] >= 100 and d["data"] == "synthetic" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"] and "l2" in d["config"]
    assert abs(d["value"] - d["config"]["nonces_per_step_per_gpu"] / d["ms_per_step"] / 1e3) < 0.01 * d["value"]
    e = d["e2e"]
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(e) and e["h2d_bytes_per_step"] > 0 < e["d2h_bytes_per_step"]
    assert e["value"] != d["value"] and 0.9 * d["value"] < e["value"] < 1.05 * d["value"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert abs(r["achieved"] - d["value"] * 1e6 * 16384 / 1e9) < 0.01 * r["achieved"]      # algorithmic bytes, not traffic
    assert r["traffic"] is None or 0.9 < r["traffic"] / (16384 * d["config"]["nonces_per_step_per_gpu"]) < 1.1
    c = d["cpu_baseline"]
    assert {"value", "unit", "cores", "kind", "sample"} <= set(c) and c["kind"] in ("reference", "port") and c["cores"] >= 1
    k = d["clocks"]
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(k)
    assert not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(k["reasons"])
    assert d["gpu_launches"] >= d["steps"]


def test_reference_arm_runs_on_the_host_cores():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    d = json.loads(out.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and BASE_KEYS | {"cpu_baseline"} <= set(d)
    assert d["metric"] == "KawPow MH/s" and d["unit"] == "MH/s" and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] in ("reference", "port")
    assert d["config"]["workload"] == "kawpow_search_epoch448"

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


def test_round2_bench_lines_carry_the_extra_keys():
    """The round-2 lines (profiles/r2_bench_n*.json) keep the base contract and add what the round-1 review asked to see in
    the driver-run line: the epoch build per phase (both exchange routes at N>1, digest check), and the other kernels."""
    for n in (1, 2, 4, 8):
        d = json.loads(open(os.path.join(ROOT, "profiles", f"r2_bench_n{n}.json")).read().strip().splitlines()[-1])
        assert BASE_KEYS | {"roofline", "clocks", "epoch_build", "paths"} <= set(d), n
        assert d["n_gpus"] == n and d["metric"] == "KawPow MH/s" and d["warmup"] >= 3
        assert abs(d["value"] - n * d["config"]["nonces_per_step_per_gpu"] / d["ms_per_step"] / 1e3) < 0.01 * d["value"]
        assert 0.9 * d["value"] < d["e2e"]["value"] < 1.05 * d["value"] and d["e2e"]["h2d_bytes_per_step"] > 0
        assert d["gpu_launches"] >= d["steps"]
        assert not {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(d["clocks"]["reasons"])
        eb = d["epoch_build"]
        assert eb["dag_checksum_equals_reference_digest"] is True and eb["total_s"] < 2.0
        if n == 1:
            assert eb["exchange"] is None and eb["allgather_ms"] is None
            assert d["cpu_baseline"]["kind"] in ("reference", "port")
        else:
            assert eb["exchange"] == "fused" and eb["allgather_ms"] is None  # the exchange is inside the kernel: no separate time
            other = eb["other_route"]
            assert other["exchange"] == "nccl" and other["allgather_ms"] > 0 and other["total_s"] < 2.0
        p = d["paths"]
        assert {"K3_dag_items_kernel", "K6_kawpow_interp_kernel", "K7_ethash_hash_kernel", "nx_miner_search", "equihash_gated"} <= set(p)
        assert 0.5 < p["K3_dag_items_kernel"]["frac_of_gather_peak"] <= 1.0
        assert p["nx_miner_search"]["devices"] == n and p["nx_miner_search"]["mhs_wall"] > 150 * n
        eq = p["equihash_gated"]
        assert "not checked against the reference" in eq["parity"] and all(r["rows_dropped"] == 0 for r in eq["runs"])

"""CPU: the bench JSON contract, checked on the lines committed under profiles/ (produced on a B200 by bench.py)."""
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")


def load(name):
    with open(os.path.join(P, name)) as f:
        return json.loads([l for l in f if l.startswith("{")][0])


def test_default_line_has_every_contract_key():
    d = load("r01_bench_final_default.json")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks"):
        assert k in d, k
    assert d["unit"] == "GLUPS" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["n_gpus"] == 1 and d["warmup"] >= 3
    assert "workload" in d["config"] and "model" not in d["config"] and "l2" in d["config"]
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["traffic"] is None or r["traffic"] > 0
    c = d["cpu_baseline"]
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1 and c["value"] > 0 and c["sample"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 4096 * 4096 * 8 == e["d2h_bytes_per_step"] and 0 < e["value"] < d["value"]
    assert d["gpu_launches"] == d["steps"] * d["config"]["timesteps_per_step"]
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    bad = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert not bad & set(d["clocks"]["reasons"])
    # value is consistent with its own timing: interior updates / time
    lups = (4096 - 2) ** 2 * d["config"]["timesteps_per_step"]
    assert abs(d["value"] - lups / (d["ms_per_step"] * 1e-3) / 1e9) / d["value"] < 1e-6


def test_reference_line():
    d = load("r01_bench_final_reference.json")
    assert d["impl"] == "reference" and d["unit"] == "GLUPS" and d["gpu_launches"] == 0
    assert d["e2e"] == {"value": d["value"], "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] == "port"


@pytest.mark.parametrize("name,n", [("r01_bench_n2_cfg2_final.json", 2), ("r01_bench_n4_cfg2_final.json", 4),
                                    ("r01_bench_n8_cfg2_final.json", 8)])
def test_multi_gpu_lines(name, n):
    d = load(name)
    assert d["n_gpus"] == n and d["scaling"] == "weak" and d["config"]["ny_global"] == 4096 * n
    assert "halo" in d and d["halo"]["exchanges"] > 0
    one = load("r01_bench_n1_cfg2_final.json")["value"]
    assert 0.9 < d["value"] / (n * one) <= 1.02          # weak-scaling efficiency measured in round 1

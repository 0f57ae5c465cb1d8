"""CPU: the bench JSON contract, checked on the lines committed under profiles/ (produced on B200s by bench.py)."""
import json
import os

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
P = os.path.join(ROOT, "profiles")


def load(name):
    with open(os.path.join(P, name)) as f:
        return json.loads([l for l in f if l.startswith("{")][0])


def test_default_line_has_every_contract_key():
    d = load("r02_bench_default_cfg3.json")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks"):
        assert k in d, k
    assert d["unit"] == "GLUPS" and d["higher_is_better"] is True and d["vs_baseline"] is None
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["n_gpus"] == 1 and d["warmup"] >= 3
    c = d["config"]
    assert c["workload"].startswith("cfg3") and c["nx"] == 16384 == c["ny_global"] and c["timesteps_per_step"] == 500
    assert "model" not in c and "l2" in c and c["tblock"] == 1 and c["fmad"] is False
    r = d["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["frac"] >= 0.85                                      # the north-star target on this very configuration
    assert r["traffic"] == r["ncu"]["bytes"] and 0.95 < r["traffic"] / r["algorithmic_bytes_per_launch"] < 1.05
    assert os.path.isfile(os.path.join(ROOT, r["ncu"]["source"]))
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["value"] > 0 and cb["sample"]
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 16384 * 16384 * 8 == e["d2h_bytes_per_step"] and 0 < e["value"] < d["value"]
    assert e["steps"] == d["steps"] and 0 < e["serial"]["value"] < e["value"]
    assert d["gpu_launches"] == d["steps"] * c["timesteps_per_step"]
    assert set(d["clocks"]) >= {"sm_mhz", "sm_max_mhz", "reasons"}
    bad = {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert not bad & set(d["clocks"]["reasons"])
    # value is consistent with its own timing: interior updates / time
    lups = (16384 - 2) ** 2 * c["timesteps_per_step"]
    assert abs(d["value"] - lups / (d["ms_per_step"] * 1e-3) / 1e9) / d["value"] < 1e-6


def test_reference_line_matches_the_gpu_arm_config():
    d = load("r02_bench_reference_cfg3.json")
    assert d["impl"] == "reference" and d["unit"] == "GLUPS" and d["gpu_launches"] == 0
    assert d["e2e"] == {"value": d["value"], "unit": "GLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["kind"] == "port"
    assert d["cpu_baseline"]["cores"] > 1                         # OpenMP threads set explicitly, whatever the launcher exported
    assert d["config"] == load("r02_bench_default_cfg3.json")["config"]


@pytest.mark.parametrize("n", [2, 4, 8])
def test_strong_scaling_lines_carry_parity(n):
    d = load("r02_bench_n%d_cfg4.json" % n)
    assert d["n_gpus"] == n and d["scaling"] == "strong" and d["config"]["ny_global"] == 32768 == d["config"]["nx"]
    assert d["config"]["workload"].startswith("cfg4")
    assert "halo" in d and d["halo"]["exchanges"] > 0
    p = d["parity"]
    assert p["ok"] is True and p["cases"] == 2
    assert all(c["equals_1gpu"] and c["band_equals_cpu_oracle"] for c in p["detail"])
    one = load("r02_bench_n1_cfg4.json")
    assert one["scaling"] == "strong" and one["config"]["workload"].startswith("cfg4")
    assert 0.9 < d["value"] / (n * one["value"]) <= 1.02         # strong-scaling efficiency measured in round 2
    assert d["roofline"]["traffic"] is None                       # no single-GPU ncu figure pasted onto a multi-GPU line


# ------------------------------------------------------------------ bench.py's own choices (no GPU needed)
def _bench_module():
    import importlib.util
    spec = importlib.util.spec_from_file_location("heat_bench", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_workload_choice_follows_the_box(monkeypatch):
    b = _bench_module()
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "0")
    assert b.choose_workload("auto", 1)[0] == "cfg3"                 # single-GPU box: the roofline workload
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "0,1,2,3,4,5,6,7")
    assert b.choose_workload("auto", 1)[0] == "cfg4"                 # N = 1 of a sweep on a multi-GPU box: strong scaling
    for n in (2, 4, 8):
        assert b.choose_workload("auto", n)[0] == "cfg4"
    assert b.choose_workload("cfg2", 8)[0] == "cfg2" and b.choose_workload("cfg5", 1)[0] == "cfg5"
    assert b.WORKLOADS["cfg3"]["nx"] == 16384 and b.WORKLOADS["cfg3"]["timesteps"] == 500
    assert b.WORKLOADS["cfg4"]["nx"] == 32768 and b.WORKLOADS["cfg4"]["scaling"] == "strong"


def test_config_is_a_function_of_the_command_line_only():
    """Both arms build `config` with workload_config from the same arguments, so the driver's same_config holds."""
    import argparse
    b = _bench_module()
    args = argparse.Namespace(dtype="f64", tblock=1, fmad=False, field="hat", kernel="auto", ghost=32, no_overlap=False, halo="auto")
    for name, world in (("cfg3", 1), ("cfg4", 1), ("cfg4", 8), ("cfg2", 4)):
        wl = b.WORKLOADS[name]
        c1 = b.workload_config(args, name, "why", wl, world)
        c2 = b.workload_config(args, name, "why", dict(wl), world)
        assert c1 == c2 and c1["workload"].startswith(name) and "l2" in c1 and "model" not in c1
        assert c1["ny_global"] == (wl["ny"] * world if wl["scaling"] == "weak" else wl["ny"])
        assert ("ghost 32" in c1["parallelism"]) == (world > 1)

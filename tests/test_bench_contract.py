"""The bench.py JSON contract, checked on CPU: the reference arm is run live at a small size, and the committed B200 lines of
the final round-1 run (profiles/r01d_bench_n1.json, r01d_bench_reference_arm.json) are checked for every key the driver reads."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
             "dtype", "data", "config", "e2e", "gpu_launches"}


def _last_json_line(text):
    lines = [l for l in text.strip().splitlines() if l.startswith("{")]
    assert lines, text[-500:]
    return json.loads(lines[-1])


def _check_common(d):
    assert BASE_KEYS <= set(d), sorted(BASE_KEYS - set(d))
    assert d["metric"] == "fused_elementwise_reduce_hbm_gbps" and d["unit"] == "GB/s" and d["higher_is_better"] is True
    assert d["dtype"] == "f32" and d["data"] == "synthetic" and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    assert {"value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"} <= set(d["e2e"])
    assert d["value"] > 0 and d["ms_per_step"] > 0


def test_reference_arm_runs_and_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                        "--ref-size", "256"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-800:]
    d = _last_json_line(r.stdout)
    _check_common(d)
    assert d["impl"] == "reference" and d["gpu_launches"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 == d["e2e"]["d2h_bytes_per_step"]


def test_reference_arm_other_ranks_exit_quietly():
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0", "--ref-size", "256"], capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_committed_b200_lines_carry_every_contract_key():
    p = os.path.join(ROOT, "profiles", "r01d_bench_n1.json")
    if not os.path.exists(p):
        pytest.skip("no committed bench line")
    d = _last_json_line(open(p).read())
    _check_common(d)
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["gpu_launches"] > 0
    rf = d["roofline"]
    assert rf["bound"] == "hbm" and rf["unit"] == "GB/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    assert rf["traffic"] is None or rf["traffic"] > 0
    cb = d["cpu_baseline"]
    assert {"value", "unit", "cores", "kind", "sample"} <= set(cb) and cb["kind"] == "reference" and cb["cores"] == 1
    ck = d["clocks"]
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(ck)
    assert not ({"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(ck["reasons"]))
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0 and d["e2e"]["value"] < d["value"]
    assert "l2" in d["config"]                        # says how the timed region defeats L2 reuse
    ref = _last_json_line(open(os.path.join(ROOT, "profiles", "r01d_bench_reference_arm.json")).read())
    _check_common(ref)
    assert ref["impl"] == "reference" and ref["config"]["workload"] == d["config"]["workload"]

"""The bench JSON line keeps the driver's contract.  Checked on the lines a B200 run of this
round committed under profiles/ (bench.py itself needs a GPU) and on bench.py's argument
parser, which must keep the driver's flags."""
import glob
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _latest(pattern):
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", pattern)))
    if not files:
        pytest.skip("no committed bench line matches %s" % pattern)
    return files[-1]


def test_our_arm_line_has_every_contract_key():
    d = json.load(open(_latest("r02z_bench_n1.json")))
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["higher_is_better"] is True and d["scaling"] in ("weak", "strong")
    assert d["dtype"] == "f64" and d["data"] == "synthetic" and d["vs_baseline"] is None
    assert "workload" in d["config"] and "model" not in d["config"]
    assert d["warmup"] >= 3 and d["gpu_launches"] > 0
    for k in ("sm_mhz", "sm_max_mhz", "reasons"):
        assert k in d["clocks"], k
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    r = d["roofline"]
    for k in ("bound", "achieved", "peak", "unit", "frac", "traffic"):
        assert k in r, k
    assert r["bound"] == "hbm" and r["unit"] == "GB/s"
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    e = d["e2e"]
    for k in ("value", "unit", "h2d_bytes_per_step", "d2h_bytes_per_step"):
        assert k in e, k
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
    c = d["cpu_baseline"]
    for k in ("value", "unit", "cores", "kind", "sample"):
        assert k in c, k
    assert c["kind"] in ("port", "reference") and c["cores"] >= 1
    # value is the whole job: points x steps / time
    pts = 16384 * 16384 * d["n_gpus"]
    assert abs(d["value"] - pts / (d["ms_per_step"] * 1e-3)) <= 1e-6 * d["value"]
    # the two formulations advanced the same model: identical checksums
    assert d["checksums_p_u_v"] == d["one_step"]["checksums_p_u_v"]
    # round 2: frac is a fraction of a roofline again (one pass over the state per launch), the temporal-blocking
    # gain is a separate field; parity ran in the driver's run and found nothing; sustained leg and config 3 present
    assert 0.5 < r["frac"] <= 1.2 and r["effective_bytes_speedup"] == 2.0
    p = d["parity"]
    assert p["checked"] is True and p["mismatches"] == 0 and p["cases"] >= 10
    for leg in ("one_step_sweep", "two_step_sweep"):      # asserted 16384^2 checksums vs the CPU arm
        assert p["full_size"][leg]["max_rel_diff_vs_cpu_long_double"] <= p["full_size"]["tolerance"] == 1e-12
    assert d["sustained"]["seconds"] >= 2.5 and d["sustained"]["value"] <= d["value"] * 1.05
    assert d["config3"]["parity"]["mismatches"] == 0
    assert d["e2e"]["model_steps_per_round_trip"] == 2


def test_both_arms_name_the_same_workload():
    ours = json.load(open(_latest("r02z_bench_n1.json")))
    ref = json.load(open(_latest("r02z_bench_reference_arm.json")))
    assert ours["config"]["workload"] == ref["config"]["workload"]
    assert ours["metric"] == ref["metric"] and ours["unit"] == ref["unit"]
    assert "reference_sample" in ref["config"]
    # the CPU arm's checksums after its steps agree with the long-double sums it prints
    for a, b in zip(ref["checksums_p_u_v"], ref["checksums_p_u_v_long_double"]):
        assert abs(a - b) <= 1e-12 * abs(b)


def test_round3_lines_share_one_config_and_carry_a_summary():
    """round 3: both arms print the same `config` object (arm-specific text lives in `config_notes`) and our
    line carries a compact top-level `summary` (parity, sustained and per-config roofline fractions)"""
    ours = json.load(open(_latest("round3*_bench_n1.json")))
    ref = json.load(open(_latest("round3*_bench_reference_arm.json")))
    if "summary" not in ours:
        pytest.skip("latest committed line predates the summary")
    assert ours["config"] == ref["config"] and set(ours["config"]) == {"workload", "l2"}
    assert "reference_sample" in ref["config_notes"] and "halo" in ours["config_notes"]
    sm = ours["summary"]
    assert sm["parity"]["mismatches"] == 0 and sm["parity"]["checks"] >= 100 and sm["parity"]["full_size_checked"]
    assert sm["sustained_value"] == ours["sustained"]["value"] <= ours["value"] * 1.05
    assert abs(sm["burst_frac"] - ours["roofline"]["frac"]) < 1e-12
    assert sm["config3_mismatches"] == 0 and 0.5 < sm["config3_two_step_frac"] < 1.2
    assert sm["e2e_value"] == ours["e2e"]["value"]
    assert "unpinned" in sm["stencil_parity_note"]


def test_reference_arm_line():
    d = json.load(open(_latest("r02z_bench_reference_arm.json")))
    assert d["impl"] == "reference" and d["gpu_launches"] == 0
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["e2e"]["value"] == d["value"] == d["cpu_baseline"]["value"]
    assert d["cpu_baseline"]["kind"] == "port"


def test_bench_keeps_the_drivers_flags():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--help"], capture_output=True, text=True,
                         timeout=120)
    assert out.returncode == 0
    for flag in ("--gpus", "--steps", "--warmup", "--impl"):
        assert flag in out.stdout, flag

"""CPU checks of bench.py's host side: the reference arm's JSON line (the reference's own ndlqr_Solve on the host
cores, `--impl reference`), the FLOP model of SURVEY 8d and the clock sampler's untimed-continuation helper.
No GPU work: the product arm of bench.py is covered by the `-m gpu` runs."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1                       # ONE JSON line on stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "solves/s" and d["higher_is_better"] is True
    assert d["metric"] == "LQR solves/sec (n=12,m=4,N=256 batched)"
    assert d["value"] > 0 and d["gpu_launches"] == 0
    assert d["e2e"] == {"value": d["value"], "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["value"] == d["value"] and cb["sample"]


def test_flop_model_matches_survey_figures():
    sys.path.insert(0, ROOT)
    import bench
    # structurally non-zero FLOPs per solve
    assert abs(bench.flop_model(12, 4, 256) / 32.2e6 - 1) < 0.01
    assert abs(bench.flop_model(12, 4, 16384) / 3.65e9 - 1) < 0.01
    assert abs(bench.flop_model(64, 16, 1024) / 23.6e9 - 1) < 0.01
    assert abs(bench.flop_model(6, 3, 256) / 4.4e6 - 1) < 0.02


def test_clock_sampler_keep_loaded_runs_the_step_a_fixed_number_of_times():
    sys.path.insert(0, ROOT)
    import bench
    s = bench.ClockSampler(0)            # never started: no nvidia-smi here
    calls = {"step": 0, "sync": 0}
    how = s.keep_loaded(lambda: calls.__setitem__("step", calls["step"] + 1),
                        lambda: calls.__setitem__("sync", calls["sync"] + 1), 37)
    assert calls == {"step": 37, "sync": 1} and "untimed continuation" in how
    s.rows = [["1965", "1965", "Not Active", "Not Active", "Not Active", "Not Active"]] * 3
    assert s.keep_loaded(lambda: 1 / 0, lambda: None, 5) == "during the timed region"
    c = s.summary()
    assert c["sm_mhz"] == 1965.0 and c["reasons"] == [] and c["samples"] == 3

"""bench.py contract on the CPU: the reference arm (`--impl reference`, the CPU port of the reference path timed on the
host cores) prints exactly one JSON line with the keys the driver reads; ranks other than 0 print nothing; the
algorithmic work figures used for the roofline match SURVEY.md section 8d (P_w of cfg A / cfg B)."""
import json
import os
import subprocess
import sys

from conftest import ROOT

sys.path.insert(0, ROOT)
import bench  # noqa: E402


def _run(env_extra=None):
    env = dict(os.environ, **(env_extra or {}))
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                          "--warmup", "1"], capture_output=True, text=True, env=env, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    return [l for l in res.stdout.splitlines() if l.strip()]


def test_reference_arm_prints_one_contract_line():
    lines = _run()
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline", "impl"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == "samples/s"
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("cfgB") and "model" not in d["config"]


def test_reference_arm_runs_on_rank_0_only():
    assert _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"}) == []


def test_algorithmic_work_matches_survey():
    kind, kw, _, _ = bench.CONFIGS["cfgA"]
    assert bench.macs_per_sample(kind, kw) == 237_502_464          # SURVEY.md 8d, probe-verified
    kind, kw, _, _ = bench.CONFIGS["cfgB"]
    assert bench.macs_per_sample(kind, kw) == 83_886_080

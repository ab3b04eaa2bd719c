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


def test_roofline_fraction_follows_from_the_lines_own_fields():
    """`roofline.frac` must be recomputable from the fields the line carries: algorithmic bytes = the hidden
    layers' weights once per direction, duration = the kernel's share of the step that `value` times."""
    class Ctx:
        pk = dict(hbm=6546.2, bf16=1679.4, bf16_sus=1409.5, src="test")
    for prec, wbytes in (("bf16", 2), ("fp32", 4)):
        r = {"B": 256, "precision": prec, "ms": 0.44, "ms_serial": 0.74,
             "kind_ms": {1: [0.02, 1.0, 0.3], 2: [0.02, 1.0, 0.3]}, "kind_launches": {1: [2, 96, 32], 2: [2, 96, 32]}}
        for streams in (1, 2):
            out = bench.hidden_roofline(Ctx(), "cfgB", r, streams)
            kind, kw, C, E, H = bench.dims("cfgB")
            hid_w = bench.macs_per_sample(kind, kw) - bench.tail_macs("cfgB")
            assert out["launches_per_step"] == 96 and out["concurrent_streams"] == streams
            assert abs(out["algorithmic_bytes_per_launch"] * 96 - 2 * hid_w * wbytes) < 1
            t = out["share_of_step"] * out["ms_per_step"] * 1e-3
            assert abs(out["share_of_step"] - 1.0 / 1.32) < 1e-12
            assert out["ms_per_step"] == (0.44 if streams == 2 else 0.74)
            if out["bound"] == "hbm":
                gbs = out["launches_per_step"] * out["algorithmic_bytes_per_launch"] / t / 1e9
                assert abs(out["achieved"] - gbs) < 1e-6 * gbs and abs(out["frac"] - gbs / 6546.2) < 1e-9
            else:                       # parity mode at batch 256: 3 MMA passes, 4 B per weight -> past the ridge
                tf = 2 * 2.0 * hid_w * 256 * 3 / t / 1e12
                assert prec == "fp32" and abs(out["achieved"] - tf) < 1e-6 * tf and abs(out["frac"] - tf / 1679.4) < 1e-9
            assert abs(out["avg_launch_us"] - 1e6 * t * streams / 96) < 1e-9
            step_bytes = 2 * bench.pass_bytes("cfgB", 256, prec)
            assert abs(out["step_frac"] - step_bytes / (out["ms_per_step"] * 1e-3) / 1e9 / 6546.2) < 1e-12
    # minimum HBM bytes per pass = weights once + B * (2C + E + 1) * 4
    assert bench.pass_bytes("cfgA", 64, "bf16") == 237_502_464 * 2 + 64 * (2 * 128 + 128 + 1) * 4

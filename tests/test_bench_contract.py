"""bench.py's driver contract: the reference arm answers with one JSON line and exit 0 wherever it runs; on a GPU the
headline line carries every key the driver reads."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=600):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, timeout=timeout,
                       cwd=ROOT)
    lines = [l for l in r.stdout.strip().splitlines() if l.startswith("{")]
    return r, lines


def test_reference_arm_reports_unavailable_and_exits_zero():
    r, lines = _run(["--impl", "reference", "--gpus", "1", "--steps", "2", "--warmup", "1"], timeout=120)
    assert r.returncode == 0, r.stderr[-400:]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and isinstance(d["unavailable"], str) and d["unavailable"]
    assert "\n" not in d["unavailable"]


@pytest.mark.gpu
def test_headline_line_has_the_contract_keys():
    r, lines = _run(["--steps", "6", "--warmup", "3"])
    assert r.returncode == 0, r.stderr[-800:]
    assert len(lines) == 1, r.stdout[-400:]
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "gpu_launches"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 6 and d["warmup"] == 3 and d["higher_is_better"] is True
    assert d["dtype"] == "bf16" and "synthetic" in d["data"]
    assert d["gpu_launches"] >= 6 * 100  # our kernels, counted from the schedule
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] == 16 * 256 * 256 * 3 + 16 * 256 * 256 and e["d2h_bytes_per_step"] > 0
    assert 0 < e["value"] <= d["value"] * 1.10  # end to end cannot beat the device-only number (6-step timing noise aside)
    assert abs(d["value"] - 16 / (d["ms_per_step"] * 1e-3)) < 1e-6 * d["value"]

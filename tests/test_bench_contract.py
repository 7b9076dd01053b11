"""bench.py prints exactly one JSON line with the keys the driver reads (contract in the task brief)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "cpu_baseline"}


def run_bench(*args, timeout=900):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                       timeout=timeout, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    return json.loads(lines[0])


def test_reference_arm_on_host_cores():
    """`--impl reference` needs no GPU: the oracle port on a bounded sample of the same workload."""
    d = run_bench("--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-rows", "12")
    assert BASE_KEYS <= set(d) and d["impl"] == "reference"
    assert d["unit"] == "pairs/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "rows" in cb["sample"]
    assert "workload" in d["config"] and "model" not in d["config"]


@pytest.mark.gpu
def test_device_arm_line():
    d = run_bench("--steps", "1", "--warmup", "3", "--pairs", "2", "--e2e-pairs", "2", "--cpu-rows", "12", "--no-post")
    assert BASE_KEYS | {"gpu_launches", "clocks", "roofline"} <= set(d)
    assert d["n_gpus"] == 1 and d["value"] > 0 and d["gpu_launches"] > 0 and d["vs_baseline"] is None
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] == 2 * 2 * 375 * 1242 * 3 and e["d2h_bytes_per_step"] == 3 * 2 * 375 * 1242
    r = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(r) and 0 < r["frac"] < 1
    assert abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1

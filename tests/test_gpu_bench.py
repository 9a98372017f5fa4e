"""The bench contract on a small workload: one JSON line with every key the driver reads, and a launch count that
matches a CUDA path (not 0 / 1: a value like that reads as a CPU or library fallback)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_bench_line_contract():
    steps = 3
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", str(steps), "--warmup", "3",
                        "--pairs", "8", "--model", "dinov2_vits14", "--no-cpu-baseline"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in d, k
    assert d["steps"] == steps and d["n_gpus"] == 1 and d["value"] > 0
    # 12 blocks x (2 LayerNorm + 4 GEMM + attention) + front end + matching + control, every step
    assert d["gpu_launches"] >= 80 * steps, d["gpu_launches"]
    assert d["e2e"]["value"] > 0 and d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    assert d["e2e"]["value"] != d["value"]
    rf = d["roofline"]
    assert rf["bound"] in ("hbm", "tensor") and 0 < rf["frac"] < 1.5 and rf["peak"] > 0
    assert "workload" in d["config"]

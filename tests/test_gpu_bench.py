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
    # 12 blocks x (4 GEMM + attention; LayerNorm is folded into the GEMMs) + front end + matching + control, every step
    assert d["gpu_launches"] >= 60 * steps, d["gpu_launches"]
    assert d["e2e"]["value"] > 0 and d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    assert d["e2e"]["value"] != d["value"]
    rf = d["roofline"]
    assert rf["bound"] in ("hbm", "tensor") and 0 < rf["frac"] < 1.5 and rf["peak"] > 0
    assert "workload" in d["config"]
    # round 2: per-step percentiles, parity of the benchmarked batch, and an account of the whole step
    assert d["step_ms_p50"] > 0 and d["step_ms_p95"] >= d["step_ms_p50"]
    assert d["parity_check"]["pairs_checked"] >= 8 and d["parity_check"]["max_twist_rel_err_vs_unbatched"] == 0.0
    acct = d["step_accounting"]
    assert acct["nodes"] >= 60 and acct["kernels_ms"] > 0 and 0.9 < acct["sum_over_step"] < 1.1, acct
    assert d["latency"]["p50_ms"] > 0


@pytest.mark.gpu
def test_bench_match_sweep_contract():
    env = dict(os.environ, VITVS_SWEEP="quick")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--config", "5", "--steps", "3", "--warmup", "3"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][0])
    assert d["metric"] == "match_ibvs_episodes_per_sec" and d["value"] > 0 and d["config"]["baseline_config"] == 5
    pts = [p for p in d["sweep"] if "skipped" not in p]
    assert len(pts) == 4 and all(p["planted_matches_recovered"] >= 0.99 and p["n_selected"] == 24 for p in pts)
    assert all(p["sim_tflops"] > 0 for p in pts)

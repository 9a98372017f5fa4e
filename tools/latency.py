"""Single frame-pair latency probe (the metric's second half): p50 / p95 wall clock of vvs_frame_pairs_to_twist_host for
one pair (pinned host buffers; H2D + one graph replay + D2H + sync) and the device time of the replay alone.
Usage: python tools/latency.py [model ...]   (env: VITVS_PDL, VITVS_GEMM_CTAS, VITVS_LAT_TILES ...)"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vitvs_b200 import Engine, ServoConfig, synth  # noqa: E402
from vitvs_b200.weights import random_state_dict  # noqa: E402

models = sys.argv[1:] or ["dinov2_vits14", "dinov2_vitb14"]
g, c, d = synth.frame_pair(0)
gh, ch, dh = (torch.from_numpy(a[None].copy()).pin_memory() for a in (g, c, d))
tw = torch.empty(1, 6, dtype=torch.float64).pin_memory()
ns = torch.empty(1, dtype=torch.int32).pin_memory()
for m in models:
    cfg = ServoConfig(model_type=m)
    eng = Engine(cfg, max_images=2)
    eng.load_weights(random_state_dict(m, seed=0, max_block=cfg.layer))
    for _ in range(30):
        eng.frame_pairs_to_twist_host(gh, ch, dh, tw, ns)
    lat = []
    for _ in range(300):
        t0 = time.perf_counter()
        eng.frame_pairs_to_twist_host(gh, ch, dh, tw, ns)
        lat.append((time.perf_counter() - t0) * 1e3)
    lat.sort()
    gd, cd, dd = gh.cuda(), ch.cuda(), dh.cuda()
    for _ in range(10):
        eng.frame_pairs_to_twist(gd, cd, dd)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(100):
        eng.frame_pairs_to_twist(gd, cd, dd)
    e1.record()
    torch.cuda.synchronize()
    tr = eng.trace_step(1, replays=3)
    kern = sum(e - b for _, b, e in tr) / 1e3
    span = (tr[-1][2] - tr[0][1]) / 1e3
    print(f"{m} PDL={os.environ.get('VITVS_PDL', '0')} CTAS={os.environ.get('VITVS_GEMM_CTAS', '2')} "
          f"LAT={os.environ.get('VITVS_LAT_TILES', '-')}: host p50 {lat[len(lat) // 2]:.4f} ms p95 {lat[int(len(lat) * .95)]:.4f} "
          f"min {lat[0]:.4f}; device back-to-back {e0.elapsed_time(e1) / 100:.4f} ms/call; traced: {len(tr)} kernels "
          f"{kern:.1f} us busy, span {span:.1f} us", flush=True)
    if os.environ.get("VITVS_LAT_VERBOSE"):
        per, gaps = {}, {}
        for i, (cls, b, e) in enumerate(tr):
            per.setdefault(cls, []).append((e - b) / 1e3)
            if i + 1 < len(tr):
                gaps.setdefault(cls, []).append((tr[i + 1][1] - e) / 1e3)
        for cls, v in sorted(per.items(), key=lambda kv: -sum(kv[1])):
            gl = gaps.get(cls, [0.0])
            print(f"    {cls:12s} n={len(v):3d} busy {sum(v):7.1f} us (mean {sum(v) / len(v):5.1f})  gap after: mean "
                  f"{sum(gl) / len(gl):5.2f} us", flush=True)
    eng.close()

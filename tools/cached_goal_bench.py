"""Cached-goal mode against the pair call (an extra, reported separately from the headline): device time per step of
vvs_frame_pairs_to_twist (goal + current forwarded) and vvs_set_goal once + vvs_current_to_twist (current only).
Usage: python tools/cached_goal_bench.py [model] [pairs]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vitvs_b200 import Engine, ServoConfig, synth  # noqa: E402
from vitvs_b200.weights import random_state_dict  # noqa: E402

model = sys.argv[1] if len(sys.argv) > 1 else "dinov2_vitb14"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
cfg = ServoConfig(model_type=model)
eng = Engine(cfg, max_images=2 * B)
eng.load_weights(random_state_dict(model, seed=0, max_block=cfg.layer))
fr = [synth.frame_pair(i % 8) for i in range(B)]
g, c, d = (torch.from_numpy(np.stack([f[k] for f in fr])).cuda() for k in range(3))


def timed(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


ref = eng.frame_pairs_to_twist(g, c, d)
t_pair = timed(lambda: eng.frame_pairs_to_twist(g, c, d))
eng.set_goal(g)
out = eng.current_to_twist(c, d)
assert torch.equal(out["twist"], ref["twist"])
t_cached = timed(lambda: eng.current_to_twist(c, d))
print(f"{model} {B} pairs: pair call {t_pair:.3f} ms/step ({B / t_pair * 1e3:.0f} pairs/s), cached goal {t_cached:.3f} ms/step "
      f"({B / t_cached * 1e3:.0f} pairs/s), x{t_pair / t_cached:.2f}; twists bit-identical")

"""Asynchronous host-buffer entries (two landing zones): results equal the synchronous call's; pipelined e2e throughput."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vitvs_b200 import Engine, ServoConfig, synth  # noqa: E402
from vitvs_b200.weights import random_state_dict  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
cfg = ServoConfig(model_type="dinov2_vitb14")
eng = Engine(cfg, max_images=2 * B)
eng.load_weights(random_state_dict("dinov2_vitb14", seed=0, max_block=cfg.layer))
sets, raws = [], []
for s in range(3):
    raw = [synth.raw_frame_pair(8 * s + i) for i in range(8)]
    idx = [i % 8 for i in range(B)]
    rz = [(synth.resize_square(g, 308), synth.resize_square(c, 308), d) for g, c, d in raw]
    sets.append(tuple(torch.from_numpy(np.stack([rz[j][k] for j in idx])).pin_memory() for k in range(3)))
    raws.append(tuple(torch.from_numpy(np.stack([raw[j][k] for j in idx])).pin_memory() for k in range(3)))


def outs():
    return (torch.empty(B, 6, dtype=torch.float64).pin_memory(), torch.empty(B, dtype=torch.int32).pin_memory(),
            torch.empty(B, dtype=torch.float32).pin_memory())


ok = True
for name, data, fn in (("pairs", sets, eng.frame_pairs_to_twist_host), ("frames", raws, eng.frames_to_twist_host)):
    ref = []
    for d in data:
        o = outs()
        fn(*d, *o)
        ref.append(tuple(t.clone() for t in o))
    bufs = [outs(), outs()]
    got = []
    N = 9
    for i in range(N):
        fn(*data[i % 3], *bufs[i % 2], wait=False)
        if i >= 1:
            eng.wait(1)
            got.append(tuple(t.clone() for t in bufs[(i - 1) % 2]))
    eng.wait(0)
    got.append(tuple(t.clone() for t in bufs[(N - 1) % 2]))
    same = all(all(torch.equal(a, b) for a, b in zip(got[i], ref[i % 3])) for i in range(N))
    ok &= same
    for mode in ("sync", "async"):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        K = 30
        for i in range(K):
            if mode == "sync":
                fn(*data[i % 3], *bufs[i % 2])
            else:
                fn(*data[i % 3], *bufs[i % 2], wait=False)
                if i >= 1:
                    eng.wait(1)
        eng.wait(0)
        dt = time.perf_counter() - t0
        print(f"{name} {mode}: {K * B / dt:.1f} pairs/s ({1e3 * dt / K:.3f} ms per call)  results equal sync: {same}", flush=True)
print("OK" if ok else "MISMATCH")

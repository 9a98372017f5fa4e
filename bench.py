#!/usr/bin/env python
"""bench.py - frame-pairs -> twist per second of the ViT-VS `dino` servoing hot path on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--pairs B] [--model M]

One "step" = one pass of the whole hot path (DINOv2 forward of goal + current frames, descriptors,
mutual-NN matching, IBVS twist) over one batch of synthetic frame pairs.  At N=1 the workload is
BASELINE.json configs[1]: DINOv2 ViT-B/14, batch of 64 frame pairs, S=308, layer 11, binned descriptors.
N>1 (torchrun): every rank owns its own 64 pairs (weak scaling, no collective on the data path);
barrier + device synchronise on both sides, CUDA-event time, max over ranks.

`--impl reference` times the reference's own CPU path (the oracle port of vitvs_v2.py /
dinov2_extractor.py - the reference is Python and cannot travel; /root/reference is not read) on the
host cores, one frame pair per step.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "frame_pairs_to_twist_per_sec"
UNIT = "frame-pairs/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            d = json.load(fh)
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                 "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def wait_first(self, timeout_s: float = 5.0):
        t0 = time.perf_counter()
        while self.proc is not None and not self.lines and time.perf_counter() - t0 < timeout_s:
            time.sleep(0.02)

    def mark(self):
        """Start of the timed region: samples taken before this belong to the warm-up."""
        self.first = len(self.lines)

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        last = len(self.lines)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        first = getattr(self, "first", 0)
        # the samples of the timed region; a region shorter than the sampling period borrows the one just before it
        window = self.lines[first:last] or self.lines[max(first - 1, 0):last]
        for ln in window:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------- CPU arm
def cpu_reference_setup(model_type: str, size: int):
    import torch
    from oracle import dinov2          # the oracle port: the only CPU implementation, never the product path
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    return dinov2.build(model_type, seed=0), cores


def cpu_reference_step(model, pair, literal: bool):
    from oracle import servo
    goal, cur, depth = pair
    return servo.frame_pair_to_twist(model, goal, cur, depth, literal_similarity=literal)


def cpu_baseline(model_type: str, size: int, budget_s: float = 12.0) -> dict:
    """Bounded sample of the same workload through the oracle port on the host cores."""
    from vitvs_b200 import synth
    model, cores = cpu_reference_setup(model_type, size)
    pairs = [synth.frame_pair(i, size) for i in range(2)]
    t0 = time.perf_counter()
    n = 0
    while n < 2 or time.perf_counter() - t0 < budget_s:          # a bounded sample: ~budget_s seconds of CPU work
        cpu_reference_step(model, pairs[n % len(pairs)], literal=True)
        n += 1
    dt = time.perf_counter() - t0
    t1 = time.perf_counter()
    cpu_reference_step(model, pairs[0], literal=False)
    dv = time.perf_counter() - t1
    return {"value": n / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} frame pair(s) of the same workload ({model_type}, S={size}, binned) through oracle/servo.py "
                      f"with the reference's per-token CosineSimilarity loop (vitvs_v2.py:49-56), torch fp32, "
                      f"{cores} threads, {dt:.1f} s; with that loop vectorised into one GEMM: {1.0 / dv:.3f} pairs/s"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    from vitvs_b200 import synth
    size = 308
    literal = os.environ.get("VITVS_REF_VECTORISED", "0") != "1"
    model, cores = cpu_reference_setup(args.model, size)
    pairs = [synth.frame_pair(i, size) for i in range(2)]
    for i in range(args.warmup):
        cpu_reference_step(model, pairs[i % 2], literal)
    t0 = time.perf_counter()
    for i in range(args.steps):
        cpu_reference_step(model, pairs[i % 2], literal)
    dt = time.perf_counter() - t0
    val = args.steps / dt
    sample = (f"1 frame pair per step ({args.model}, S={size}, binned descriptors, layer 11), oracle port of the "
              f"reference torch path, {'literal per-token similarity loop' if literal else 'vectorised similarity'}, "
              f"fp32, {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args.model, args.pairs), "pairs_per_step": 1, "image_size": size,
                   "layer": 11, "binning": True},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ----------------------------------------------------------------------------------------- GPU arm
def workload_name(model: str, pairs: int) -> str:
    nice = {"dinov2_vits14": "ViT-S/14", "dinov2_vitb14": "ViT-B/14", "dinov2_vitl14": "ViT-L/14",
            "dinov2_vitg14": "ViT-g/14"}[model]
    return f"DINOv2 {nice} batch of {pairs} frame pairs per GPU (S=308, layer 11, 3x3-binned descriptors, K=24)"


def ncu_traffic(cls):
    """DRAM bytes per launch of a kernel class from the committed ncu capture (ViT-B, 64 pairs only), else None."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1f_traffic.json")) as fh:
            return json.load(fh)["bytes_per_launch"].get(cls)
    except Exception:
        return None


def flops_per_pair(dim, tokens, layers, bin_parts):
    n = tokens + 1
    vit = 2.0 * 588 * dim * tokens + layers * (24.0 * n * dim * dim + 4.0 * n * n * dim)
    return 2 * vit + 2.0 * tokens * tokens * bin_parts * dim


def gemm_class_flops(cls, M, D, H, Mp, tokens, pairs, dprime):
    return {"patch_gemm": 2.0 * Mp * 588 * D, "qkv_gemm": 2.0 * M * D * 3 * D, "proj_gemm": 2.0 * M * D * D,
            "ffn1_gemm": 2.0 * M * D * H, "ffn2_gemm": 2.0 * M * D * H,
            "sim_gemm": 2.0 * pairs * tokens * tokens * dprime}.get(cls)


def run_ours(args):
    import torch
    import torch.distributed as dist
    from vitvs_b200 import ARCHS, Engine, ServoConfig, synth
    from vitvs_b200.weights import random_state_dict

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    B = args.pairs
    cfg = ServoConfig(model_type=args.model)
    dim, heads, depth, hidden, swiglu = ARCHS[args.model]
    layers = min(cfg.layer, depth - 1) + 1
    eng = Engine(cfg, max_images=2 * B)
    eng.load_weights(random_state_dict(args.model, seed=0, max_block=cfg.layer))
    S, T = cfg.dino_input_size, eng.tokens

    # synthetic inputs: 8 distinct pairs tiled to the batch, NSETS rotated input sets (rank-offset seeds)
    NSETS = 4
    base = rank * 1000
    import numpy as np
    sets_host, sets_dev, raw_host = [], [], []
    NRAW = 2                                   # un-resized 640x480 camera-frame sets for the e2e_frames leg
    for s in range(NSETS):
        raw = [synth.raw_frame_pair(base + 8 * s + i) for i in range(8)]
        prs = [(synth.resize_square(gr, S), synth.resize_square(cr, S), dr) for gr, cr, dr in raw]
        idx = [i % 8 for i in range(B)]
        g = torch.from_numpy(np.stack([prs[j][0] for j in idx]))
        c = torch.from_numpy(np.stack([prs[j][1] for j in idx]))
        d = torch.from_numpy(np.stack([prs[j][2] for j in idx]))
        sets_host.append((g.pin_memory(), c.pin_memory(), d.pin_memory()))
        sets_dev.append((g.to(dev), c.to(dev), d.to(dev)))
        if s < NRAW:
            raw_host.append((torch.from_numpy(np.stack([raw[j][0] for j in idx])).pin_memory(),
                             torch.from_numpy(np.stack([raw[j][1] for j in idx])).pin_memory(), sets_host[-1][2]))
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    # ---- device-resident throughput (value)
    sampler = ClockSampler(local)          # running since before the warm-up: nvidia-smi takes a while to come up
    if rank == 0:
        sampler.start()
        sampler.wait_first()
    for i in range(args.warmup):
        eng.frame_pairs_to_twist(*sets_dev[i % NSETS])
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    sampler.mark()
    e0.record(stream)
    for i in range(args.steps):
        out = eng.frame_pairs_to_twist(*sets_dev[i % NSETS])
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = eng.last_launch_count() * args.steps
    n_sel = out["n_selected"].float().mean().item()

    # ---- end to end through the host-buffer C-ABI call (H2D + path + D2H + sync each step)
    tw = torch.empty(B, 6, dtype=torch.float64).pin_memory()
    ns = torch.empty(B, dtype=torch.int32).pin_memory()
    msim = torch.empty(B, dtype=torch.float32).pin_memory()
    for i in range(min(args.warmup, 3)):
        eng.frame_pairs_to_twist_host(*sets_host[i % NSETS], tw, ns, msim)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record(stream)
    for i in range(args.steps):
        eng.frame_pairs_to_twist_host(*sets_host[i % NSETS], tw, ns, msim)
    e3.record(stream)
    barrier()
    ms_e2e = e2.elapsed_time(e3)

    # ---- the same from the un-resized 640x480 camera frames: PIL-exact bicubic resize on the GPU
    for i in range(min(args.warmup, 3)):
        eng.frames_to_twist_host(*raw_host[i % NRAW], tw, ns, msim)
    barrier()
    e4, e5 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e4.record(stream)
    for i in range(args.steps):
        eng.frames_to_twist_host(*raw_host[i % NRAW], tw, ns, msim)
    e5.record(stream)
    barrier()
    ms_e2e_frames = e4.elapsed_time(e5)
    launches_frames = eng.last_launch_count()

    # ---- p50 single frame-pair latency (the metric's second half): one pair per call through the same
    # host-buffer C-ABI call (H2D + device path as one CUDA-graph replay + D2H + sync), wall clock per call
    latency = None
    if rank == 0:
        eng1 = Engine(cfg, max_images=2)
        eng1.load_weights(random_state_dict(args.model, seed=0, max_block=cfg.layer))
        g1, c1, d1 = (t[:1].clone().pin_memory() for t in sets_host[0])
        tw1 = torch.empty(1, 6, dtype=torch.float64).pin_memory()
        ns1 = torch.empty(1, dtype=torch.int32).pin_memory()
        for _ in range(20):
            eng1.frame_pairs_to_twist_host(g1, c1, d1, tw1, ns1)
        lat = []
        for _ in range(200):
            t0 = time.perf_counter()
            eng1.frame_pairs_to_twist_host(g1, c1, d1, tw1, ns1)
            lat.append((time.perf_counter() - t0) * 1e3)
        lat.sort()
        latency = {"p50_ms": round(lat[len(lat) // 2], 4), "p95_ms": round(lat[int(len(lat) * 0.95)], 4),
                   "min_ms": round(lat[0], 4), "calls": len(lat), "warmup": 20, "pairs_per_call": 1,
                   "timing": "host wall clock around vvs_frame_pairs_to_twist_host (pinned buffers; H2D, device "
                             "path replayed as one CUDA graph, D2H, stream sync)"}
        eng1.close()

    if world > 1:
        t = torch.tensor([ms, ms_e2e, ms_e2e_frames], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e, ms_e2e_frames = t.tolist()

    # ---- per-kernel-class device time, live, on the launching stream (2 extra profiled steps)
    prof = None
    if rank == 0:
        eng.profile(True)
        PS = 2
        raw_dev = tuple(t.to(dev) for t in raw_host[0])
        for i in range(PS):
            eng.frames_to_twist(*raw_dev)          # resize (both frames of every pair) + the pre-resized path
        torch.cuda.synchronize(dev)
        del raw_dev
        prof = eng.profile_read()
        eng.profile(False)
        prof = {k: (v[0] / PS, v[1] // PS) for k, v in prof.items()}

    if rank == 0:
        pk = peaks()
        total_pairs = world * B * args.steps
        value = total_pairs / (ms / 1e3)
        e2e_value = total_pairs / (ms_e2e / 1e3)
        M, Mp = 2 * B * (T + 1), 2 * B * T
        parts = 9 if cfg.use_feature_binning else 1
        sim_k = dim if eng.gram_stencil else parts * dim     # executed K of the similarity product
        step_ms_prof = sum(v[0] for v in prof.values())
        kernels = {}
        for k, (tot_ms, cnt) in sorted(prof.items(), key=lambda kv: -kv[1][0]):
            ent = {"ms_per_step": round(tot_ms, 4), "launches_per_step": cnt, "share": round(tot_ms / step_ms_prof, 4)}
            fl = gemm_class_flops(k, M, dim, hidden, Mp, T, B, sim_k)
            if k == "attention":
                fl = 4.0 * (T + 1) * (T + 1) * dim * 2 * B
            if fl:
                ent["tflops"] = round(fl / (tot_ms / cnt * 1e-3) / 1e12, 1)
                ent["frac_of_sustained_peak"] = round(ent["tflops"] / pk["tf_sustained"], 3)
            # HBM-bound classes: ALGORITHMIC bytes per launch (SURVEY 8d / DESIGN.md 3) against the measured copy bandwidth
            S_ = cfg.dino_input_size
            by = {"layernorm": M * dim * (4 + 2) + 2 * dim * 4,
                  "proj_gemm": M * dim * (2 + 4 + 4),                                   # h read, x read-modify-write
                  "im2col": 2 * B * (S_ * S_ * 3 + (T + 1) * 640 * 2),
                  "descriptors": 2 * B * T * (dim * 4 + parts * dim * 2 + 4),           # tokens read once, binned rows written
                  "resize": B * (cfg.u_max * cfg.v_max * 3 + S_ * S_ * 3),              # per launch pair (H + V) of one frame set
                  }.get(k)
            if by:
                n_units = cnt if k != "resize" else max(cnt // 2, 1)
                ent["gbps"] = round(by / (tot_ms / n_units * 1e-3) / 1e9, 1)
                ent["frac_of_hbm_peak"] = round(ent["gbps"] / pk["hbm"], 3)
            kernels[k] = ent
        dom = next(k for k in kernels if k.endswith("_gemm"))
        dom_fl = gemm_class_flops(dom, M, dim, hidden, Mp, T, B, sim_k)
        dom_ms = prof[dom][0] / prof[dom][1]
        achieved = dom_fl / (dom_ms * 1e-3) / 1e12
        fl_pair = flops_per_pair(dim, T, layers, 1 if eng.gram_stencil else parts)   # executed flops
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(args.model, B), "pairs_per_gpu_per_step": B, "image_size": S,
                       "layer": cfg.layer, "binning": bool(cfg.use_feature_binning), "num_pairs": cfg.num_pairs,
                       "weights": "random-init (seed 0)",
                       "l2": f"{NSETS} rotated input sets; per-step activation traffic (~GBs) exceeds the 126 MB L2",
                       "mean_selected_pairs": n_sel,
                       "matching": ("Gram-stencil: 9-point stencil over the [T,D]x[D,T] token Gram matrix (executed "
                                    "2*T^2*D flops; the 9D-wide binned descriptors are never written)"
                                    if eng.gram_stencil else "direct [T,9D]x[9D,T] product on materialised descriptors"),
                       "algorithmic_gflop_per_pair": round(fl_pair / 1e9, 2),
                       "whole_path_tflops": round(value / world * fl_pair / 1e12, 1),
                       "whole_path_frac_of_sustained_peak": round(value / world * fl_pair / 1e12 / pk["tf_sustained"], 3)},
            "e2e": {"value": e2e_value, "unit": UNIT,
                    "h2d_bytes_per_step": B * (2 * S * S * 3 + cfg.u_max * cfg.v_max * 2),
                    "d2h_bytes_per_step": B * (6 * 8 + 4 + 4), "ms_per_step": ms_e2e / args.steps},
            "e2e_frames": {"value": total_pairs / (ms_e2e_frames / 1e3), "unit": UNIT,
                           "h2d_bytes_per_step": B * (2 * cfg.u_max * cfg.v_max * 3 + cfg.u_max * cfg.v_max * 2),
                           "d2h_bytes_per_step": B * (6 * 8 + 4 + 4), "ms_per_step": ms_e2e_frames / args.steps,
                           "gpu_launches_per_step": launches_frames,
                           "what": "vvs_frames_to_twist_host: un-resized 640x480 RGB frames in pinned host memory; the "
                                   "PIL-exact bicubic resize (vitvs_v2.py:474-475) runs on the GPU"},
            "latency": latency,
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": f"gemm_tc_kernel ({dom})", "achieved": round(achieved, 1),
                         "peak": pk["tf_sustained"], "unit": "TFLOP/s", "frac": round(achieved / pk["tf_sustained"], 3),
                         "traffic": ncu_traffic(dom) if (args.model == "dinov2_vitb14" and B == 64) else None,
                         "traffic_unit": "DRAM bytes per launch (profiles/r1f_traffic.json, one ncu --set full capture)",
                         "peak_source": pk["source"] + ", sustained figure (kernel timed inside a long step)",
                         "avg_launch_ms": round(dom_ms, 4), "flops_per_launch": dom_fl},
            "kernels": kernels,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(args.model, S)
        print(json.dumps(line), flush=True)
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=64, help="frame pairs per GPU per step (BASELINE configs[1]: 64)")
    ap.add_argument("--model", default="dinov2_vitb14")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

"""CPU-side checks: the C-ABI library loads and exports every symbol include/vitvs_b200.h declares (no
compute without a GPU), the host logic (config mirror, pos-embed interpolation, weights, sharding incl. a
world_size-2 gloo run) and the loud failure when no GPU is present."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from vitvs_b200 import _lib
    from vitvs_b200.build import build_library
    build_library()
    hdr = open(os.path.join(ROOT, "include", "vitvs_b200.h")).read()
    declared = set(re.findall(r"VVS_API\s+(?:const\s+char\*|int)\s+(vvs_\w+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = _lib.load()
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    for name in declared:
        assert getattr(lib, name) is not None
    # struct layouts the ctypes mirror assumes
    import ctypes as C
    # see vvs_config: ... + patch_stride, padded to the 8-byte alignment of the doubles
    assert C.sizeof(_lib.Config) == (7 * 4 + 2 * 4 + 4 + 6 * 8 + 3 * 4 + 4 + 2 * 4 + 4) + 4 + 8
    assert C.sizeof(_lib.MatchOut) == 13 * 8 and C.sizeof(_lib.PairOut) == 16 * 8


def test_last_error_and_null_handle_paths_do_not_need_a_gpu():
    from vitvs_b200 import _lib
    lib = _lib.load()
    assert lib.vvs_destroy(None) == 0
    assert lib.vvs_num_sms(None) == 0
    assert lib.vvs_weights_ready(None) != 0
    import ctypes as C
    h = C.c_void_p()
    rc = lib.vvs_create(None, C.byref(h))
    assert rc == -1 and b"null" in lib.vvs_last_error(None)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_engine_fails_loudly_without_gpu():
    from vitvs_b200 import Engine, VvsError
    with pytest.raises(VvsError, match="no CPU fallback"):
        Engine()


def test_config_mirrors_reference_yaml(tmp_path):
    from vitvs_b200 import ServoConfig
    y = tmp_path / "config.yaml"
    y.write_text("u_max: 640\nv_max: 480\nlambda_: 0.03\nf_x: 502.3016357421875\nf_y: 502.3016357421875\n"
                 "num_pairs: 24\nimage_path: 'goalrgb.jpg'\nmax_velocity: 1.0\ndino_input_size: 308\n"
                 "use_feature_binning: true\nnum_samples: 500\nema_alpha: 0.8\nmin_iterations: 300\n")
    c = ServoConfig.from_yaml(str(y))
    assert (c.u_max, c.v_max, c.num_pairs, c.dino_input_size, c.use_feature_binning) == (640, 480, 24, 308, True)
    assert c.c_x == 320 and c.c_y == 240 and c.lambda_ == 0.03 and c.ema_alpha == 0.8
    assert c.layer == 11 and c.facet == "token" and c.same_image_threshold == 0.99      # hard-coded in the reference
    assert ServoConfig() == c


def test_pos_embed_interpolation_matches_oracle():
    from oracle import dinov2
    from vitvs_b200.engine import interpolate_pos_embed
    m = dinov2.build("dinov2_vits14", seed=0)
    for grid in (22, 7, 37):
        S = grid * 14
        x = torch.zeros(1, 1 + grid * grid, 384)
        with torch.no_grad():
            ref = m.interpolate_pos_encoding(x, S, S)[0]
        mine = interpolate_pos_embed(m.pos_embed, grid)
        assert torch.equal(ref, mine)


def test_random_state_dict_matches_architecture():
    from oracle import dinov2
    from vitvs_b200.weights import random_state_dict
    for name in ("dinov2_vits14", "dinov2_vitg14"):
        sd = random_state_dict(name, seed=0, max_block=1)
        arch = dinov2.ARCHS[name]
        ref = dinov2.DinoV2(dinov2.Arch(arch.name, arch.dim, arch.heads, 2, arch.swiglu)).state_dict()
        for k, v in sd.items():
            assert k in ref and tuple(ref[k].shape) == tuple(v.shape), k
        assert {k for k in ref if k.startswith("blocks.")} == {k for k in sd if k.startswith("blocks.")}


def test_synthetic_inputs_are_deterministic_and_well_formed():
    from vitvs_b200 import synth
    g, c, d = synth.frame_pair(0)
    g2, c2, d2 = synth.frame_pair(0)
    assert np.array_equal(g, g2) and np.array_equal(c, c2) and np.array_equal(d, d2)
    assert g.shape == (308, 308, 3) and g.dtype == np.uint8 and d.shape == (480, 640) and d.dtype == np.uint16
    assert 0.01 < (d == 0).mean() < 0.03 and 400 < np.median(d[d > 0]) < 900
    assert not np.array_equal(g, c)
    a, b, perm = synth.descriptor_sets(64, 16)
    assert a.shape == b.shape == (64, 16) and sorted(perm.tolist()) == list(range(64))


def test_shard_bounds_partition():
    from vitvs_b200.sharding import shard_bounds
    for n in (0, 1, 7, 64, 256, 513):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard_bounds(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [h - l for l, h in blocks]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_bounds(4, 2, 2)


_GLOO_WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from vitvs_b200.sharding import run_sharded, shard_bounds
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:" + sys.argv[2], rank=int(sys.argv[3]), world_size=2)
n = 7                                                    # ragged: blocks of 4 and 3
goal = torch.arange(n, dtype=torch.float64).reshape(n, 1)
def step(g, c, d):                                       # stands in for Engine.frame_pairs_to_twist on this rank's block
    return torch.cat([g * (k + 1) for k in range(6)], dim=1)
out = run_sharded(step, goal, goal, goal)
ref = step(goal, goal, goal)
assert torch.equal(out, ref), (out, ref)                 # same values whatever the sharding, unit order kept
lo, hi = shard_bounds(n, dist.get_rank(), 2)
assert (hi - lo) == (4 if dist.get_rank() == 0 else 3)
dist.barrier()
dist.destroy_process_group()
print("ok", dist.get_rank() if False else sys.argv[3])
"""


def test_sharded_run_world_size_2_gloo(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = str(s.getsockname()[1])
    s.close()
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, port, str(r)], stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True) for r in range(2)]
    outs = [p.communicate(timeout=120)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
        assert "ok" in o

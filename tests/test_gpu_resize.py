"""the GPU resize of the camera frames is bit-identical to the oracle's restatement of Pillow's
8-bit bicubic (itself pinned against Pillow in tests/test_oracle_resize.py), and the raw-frame entry points give
bit-identical twists to the pre-resized ones.  Integer work => exact equality, no tolerance."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import dinov2, pil_resize  # noqa: E402  (the checker)
from vitvs_b200 import synth  # noqa: E402


@pytest.fixture(scope="module")
def vits():
    return dinov2.build("dinov2_vits14", seed=0)


@pytest.fixture(scope="module")
def eng(vits):
    from vitvs_b200 import Engine, ServoConfig
    e = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=40)
    e.load_weights(vits)
    yield e
    e.close()


def _frames():
    rng = np.random.default_rng(5)
    g = synth.goal_frame(1)
    edges = np.zeros((480, 640, 3), np.uint8)          # overshoot / clipping after each pass
    edges[::7, :, 0] = 255
    edges[:, ::5, 1] = 255
    edges[100:300, 200:500, 2] = 255
    return np.stack([g, synth.current_frame(1, g), rng.integers(0, 256, (480, 640, 3), dtype=np.uint8), edges,
                     np.full((480, 640, 3), 255, np.uint8), np.zeros((480, 640, 3), np.uint8)])


def test_resize_bit_exact_vs_oracle(eng):
    frames = _frames()
    out = eng.resize_frames(torch.from_numpy(frames).cuda()).cpu().numpy()
    assert out.shape == (len(frames), 308, 308, 3) and out.dtype == np.uint8
    for i, f in enumerate(frames):
        assert np.array_equal(out[i], pil_resize.resize_bicubic_u8(f, 308, 308)), f"frame {i}"
    assert np.array_equal(out[0], synth.resize_square(frames[0], 308))        # == Pillow itself


@pytest.mark.parametrize("w,h,size", [(320, 240, 224), (200, 150, 224), (640, 480, 518), (322, 480, 322), (224, 224, 224),
                                      (53, 37, 28)])
def test_resize_other_geometries(vits, w, h, size):
    """up-scaling, mixed up/down, one pass skipped, both skipped, ragged rows (row bytes not a multiple of 4 or 16)."""
    from vitvs_b200 import Engine, ServoConfig
    e = Engine(ServoConfig(model_type="dinov2_vits14", u_max=w, v_max=h, dino_input_size=size), max_images=3)
    rng = np.random.default_rng(w + size)
    frames = rng.integers(0, 256, (3, h, w, 3), dtype=np.uint8)
    out = e.resize_frames(torch.from_numpy(frames).cuda()).cpu().numpy()
    for i in range(3):
        assert np.array_equal(out[i], pil_resize.resize_bicubic_u8(frames[i], size, size))
    e.close()


def test_frames_entry_equals_preresized_entry(eng):
    n = 3
    graw, craw, depth = synth.raw_frame_batch(n)
    g = np.stack([synth.resize_square(f, 308) for f in graw])
    c = np.stack([synth.resize_square(f, 308) for f in craw])
    d = torch.from_numpy(depth).cuda()
    a = eng.frame_pairs_to_twist(torch.from_numpy(g).cuda(), torch.from_numpy(c).cuda(), d, want_match=True)
    b = eng.frames_to_twist(torch.from_numpy(graw).cuda(), torch.from_numpy(craw).cuda(), d, want_match=True)
    for k in ("twist", "n_selected", "mean_sim_sel", "nn1", "nn2", "sim1", "uv", "uv_star", "sel_tok"):
        assert torch.equal(a[k], b[k]), k
    assert int(b["n_selected"].min()) == 24


@pytest.mark.parametrize("n", [1, 5, 17, 20])
def test_host_entries_pipelined_equal_device_entry(eng, n):
    """the sub-batched host pipeline (n >= 16 => two sub-batches, uneven for n = 17) returns what one device call does"""
    graw, craw, depth = synth.raw_frame_batch(n, distinct=5)
    ref = eng.frames_to_twist(torch.from_numpy(graw).cuda(), torch.from_numpy(craw).cuda(),
                              torch.from_numpy(depth).cuda())
    tw_ref, ns_ref, ms_ref = ref["twist"].cpu(), ref["n_selected"].cpu(), ref["mean_sim_sel"].cpu()
    gh, ch, dh = (torch.from_numpy(x).pin_memory() for x in (graw, craw, depth))
    for rep in range(3):               # eager, graph capture, graph replay
        tw = torch.zeros(n, 6, dtype=torch.float64).pin_memory()
        ns = torch.zeros(n, dtype=torch.int32).pin_memory()
        ms = torch.zeros(n, dtype=torch.float32).pin_memory()
        eng.frames_to_twist_host(gh, ch, dh, tw, ns, ms)
        assert torch.equal(tw, tw_ref) and torch.equal(ns, ns_ref) and torch.equal(ms, ms_ref), rep
        assert eng.last_launch_count() > 65
    # pre-resized host entry through the same pipeline
    g = torch.from_numpy(np.stack([synth.resize_square(f, 308) for f in graw])).pin_memory()
    c = torch.from_numpy(np.stack([synth.resize_square(f, 308) for f in craw])).pin_memory()
    for rep in range(2):
        tw = torch.zeros(n, 6, dtype=torch.float64).pin_memory()
        eng.frame_pairs_to_twist_host(g, c, dh, tw)
        assert torch.equal(tw, tw_ref), rep


def test_resize_argument_errors(eng):
    from vitvs_b200 import VvsError
    with pytest.raises(AssertionError):
        eng.resize_frames(torch.zeros(1, 308, 308, 3, dtype=torch.uint8).cuda())
    with pytest.raises(VvsError):
        eng.resize_frames(torch.zeros(41, 480, 640, 3, dtype=torch.uint8).cuda())     # > max_images

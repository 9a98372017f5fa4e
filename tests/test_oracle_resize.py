"""Pins oracle/pil_resize.py (restatement of Pillow's 8-bit antialiased bicubic resize)
against Pillow itself - the third-party library the reference calls at vitvs_v2.py:474-475 - bit for bit."""
import numpy as np
import pytest
from PIL import Image

from oracle import pil_resize
from vitvs_b200 import synth


def _pil(a, w, h):
    return np.asarray(Image.fromarray(a).resize((w, h)))


@pytest.mark.parametrize("h,w,ow,oh", [
    (480, 640, 308, 308),      # the path's geometry (config.yaml dino_input_size)
    (480, 640, 518, 518),      # DINOv2 native side: down-scale in x, up-scale in y
    (480, 640, 224, 224),
    (37, 53, 14, 28),          # ragged small sizes
    (100, 100, 224, 224),      # pure up-scaling (support not widened)
    (480, 640, 640, 480),      # identity: both passes skipped
    (480, 640, 308, 480),      # horizontal pass only
    (9, 7, 3, 2),
])
def test_restated_resize_equals_pillow(h, w, ow, oh):
    rng = np.random.default_rng(h * 1000 + ow)
    a = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    assert np.array_equal(pil_resize.resize_bicubic_u8(a, ow, oh), _pil(a, ow, oh))


def test_clipping_cases_equal_pillow():
    # hard 0/255 edges make the cubic lobes overshoot: exercises the clip after EACH pass
    a = np.zeros((480, 640, 3), np.uint8)
    a[::7, :, 0] = 255
    a[:, ::5, 1] = 255
    a[100:300, 200:500, 2] = 255
    assert np.array_equal(pil_resize.resize_bicubic_u8(a, 308, 308), _pil(a, 308, 308))


def test_synthetic_frames_equal_pillow_and_synth():
    g = synth.goal_frame(0)
    c = synth.current_frame(0, g)
    for f in (g, c):
        r = pil_resize.resize_bicubic_u8(f, 308, 308)
        assert np.array_equal(r, _pil(f, 308, 308))
        assert np.array_equal(r, synth.resize_square(f, 308))


def test_coefficient_tables():
    ks, bounds, kk = pil_resize.precompute_coeffs(640, 308)
    assert ks == 11 and bounds.shape == (308, 2) and kk.shape == (308, 11)
    assert (bounds[:, 0] >= 0).all() and (bounds[:, 0] + bounds[:, 1] <= 640).all()
    # normalised taps: every row sums to 2^22 up to the rounding of its taps
    assert np.abs(kk.sum(1) - (1 << 22)).max() <= 6
    assert pil_resize.precompute_coeffs(480, 308)[0] == 9
    assert pil_resize.precompute_coeffs(100, 224)[0] == 5

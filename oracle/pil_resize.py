"""Oracle: CPU restatement of Pillow's 8-bit bicubic ``Image.resize``.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference calls ``PIL.Image.resize((S, S))`` (vitvs_v2.py:474-475) with Pillow's default
resampling filter.  Pillow is a third-party dependency that is not vendored in /root/reference;
the version in this image is **Pillow 12.2.0** (``PIL.__version__``), whose default filter for
``Image.resize`` on RGB images is BICUBIC with antialiasing (support scaled by the down-scale
factor) and no ``reducing_gap``.  This file restates the published algorithm of
``src/libImaging/Resample.c`` (``precompute_coeffs``, ``normalize_coeffs_8bpc``,
``ImagingResampleHorizontal_8bpc`` / ``ImagingResampleVertical_8bpc``): two separable passes,
horizontal first, coefficients normalised in double then rounded to 22-bit fixed point, a
rounding bias of 2^21, arithmetic shift and clip to [0, 255] after EACH pass.

PARITY PINNED: tests/test_oracle_resize.py checks the restatement bit for bit against Pillow
itself (which is importable on both the build container and the GPU box) on the path's
geometry 640x480 -> 308x308 and on ragged / up-scaling / identity geometries.
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2          # Resample.c: 8-bit pixels, 2 bits of head-room
BICUBIC_SUPPORT = 2.0


def bicubic_filter(x: float) -> float:
    """Resample.c bicubic_filter: Keys cubic convolution kernel with a = -0.5."""
    a = -0.5
    if x < 0.0:
        x = -x
    if x < 1.0:
        return ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    if x < 2.0:
        return (((x - 5) * x + 8) * x - 4) * a
    return 0.0


def precompute_coeffs(in_size: int, out_size: int):
    """Resample.c precompute_coeffs + normalize_coeffs_8bpc for the full box [0, in_size).
    -> ksize, bounds[out_size, 2] (first tap, tap count), kk[out_size, ksize] int32."""
    scale = float(in_size) / out_size
    filterscale = max(scale, 1.0)
    support = BICUBIC_SUPPORT * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), np.int32)
    kk = np.zeros((out_size, ksize), np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = (xx + 0.5) * scale
        xmin = int(center - support + 0.5)          # C (int) cast of a value >= -support+0.5: trunc
        if xmin < 0:
            xmin = 0
        xmax = int(center + support + 0.5)
        if xmax > in_size:
            xmax = in_size
        xmax -= xmin
        w = [bicubic_filter((x + xmin - center + 0.5) * ss) for x in range(xmax)]
        ww = 0.0
        for v in w:
            ww += v
        for x in range(xmax):
            k = w[x] / ww if ww != 0.0 else w[x]
            # normalize_coeffs_8bpc: round half away from zero through a truncating cast
            kk[xx, x] = int(-0.5 + k * (1 << PRECISION_BITS)) if k < 0 else int(0.5 + k * (1 << PRECISION_BITS))
        bounds[xx] = (xmin, xmax)
    return ksize, bounds, kk


def _clip8(acc: np.ndarray) -> np.ndarray:
    """Resample.c clip8: arithmetic shift by PRECISION_BITS, clamp to [0, 255]."""
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def _pass(img: np.ndarray, out_size: int, axis: int) -> np.ndarray:
    in_size = img.shape[axis]
    if in_size == out_size:                          # ImagingResample skips a pass that changes nothing
        return img
    _, bounds, kk = precompute_coeffs(in_size, out_size)
    src = np.moveaxis(img, axis, 0).astype(np.int64)
    out = np.empty((out_size,) + src.shape[1:], np.uint8)
    for xx in range(out_size):
        xmin, n = bounds[xx]
        acc = np.full(src.shape[1:], 1 << (PRECISION_BITS - 1), np.int64)
        for x in range(n):
            acc += src[xmin + x] * int(kk[xx, x])
        out[xx] = _clip8(acc)
    return np.moveaxis(out, 0, axis)


def resize_bicubic_u8(frame_u8: np.ndarray, out_w: int, out_h: int) -> np.ndarray:
    """``np.asarray(Image.fromarray(frame_u8).resize((out_w, out_h)))`` for an [H, W, C] uint8 image:
    horizontal pass first (ImagingResample), then vertical, uint8 in between."""
    assert frame_u8.dtype == np.uint8 and frame_u8.ndim == 3
    tmp = _pass(frame_u8, out_w, axis=1)
    return np.ascontiguousarray(_pass(tmp, out_h, axis=0))

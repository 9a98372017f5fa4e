"""vvs_resize_frames_static: the camera-frame resize for callers that cannot allocate device memory.

Who that is: a servo node sharing its GPU with a process-wide memory pool that already owns the device (a cudaMalloc
from this library would fail, or not, depending on what ran before), and a control / camera-callback thread with a
deadline (cudaMalloc and cudaFree can wait for every stream on the device).  For them the library carries its tap
tables and the intermediate image of the two-pass resize in a fixed device array of 8.25 MiB instead of in a handle.

What is checked, through the C ABI (ctypes): the output equals that of the handle-based vvs_resize_frames byte for
byte (integer arithmetic - the resize tests use no tolerance and neither does this one) on two calls in a row that
share the array, with a change of geometry and back in between (tables rewritten in place), at the largest batch the
array takes; that one more frame is refused with a message naming the limit; and that a warmed-up call leaves the
device's free memory where it was.  The file sorts last on purpose: pytest -x stops at the first failure, and a failure
here must not hide the other tests.  It makes its own engine, so it passes alone in a fresh process as well as after
the rest of the suite."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from vitvs_b200 import _lib  # noqa: E402


def _static_resize(lib, frames: torch.Tensor, size: int) -> torch.Tensor:
    """The caller owns input and output (torch allocates them here, outside the call under test)."""
    n, h, w, _ = frames.shape
    out = torch.empty(n, size, size, 3, dtype=torch.uint8, device=frames.device)
    stream = C.c_void_p(torch.cuda.current_stream(frames.device).cuda_stream)
    rc = lib.vvs_resize_frames_static(frames.data_ptr(), n, h, w, size, out.data_ptr(), stream)
    if rc != 0:
        raise _lib.VvsError(f"vvs_resize_frames_static failed ({rc}): {lib.vvs_last_error(None).decode()}")
    return out


def _frames(seed: int, n: int, h: int, w: int) -> torch.Tensor:
    rng = np.random.default_rng(seed)
    f = rng.integers(0, 256, (n, h, w, 3), dtype=np.uint8)
    f[0, ::7, :, 0] = 255          # hard edges: overshoot and clipping after each pass
    f[0, :, ::5, 1] = 0
    return torch.from_numpy(f).cuda()


def test_static_resize_equals_handle_resize_on_consecutive_calls():
    from vitvs_b200 import Engine, ServoConfig
    lib = _lib.load()
    torch.cuda.set_device(0)
    eng = Engine(ServoConfig(model_type="dinov2_vits14"), max_images=18)                      # 640x480 -> 308
    small = Engine(ServoConfig(model_type="dinov2_vits14", u_max=320, v_max=240, dino_input_size=224), max_images=4)
    try:
        a, b = _frames(11, 6, 480, 640), _frames(12, 18, 480, 640)       # 18 frames = the whole 8 MiB intermediate
        want_a, want_b = eng.resize_frames(a), eng.resize_frames(b)
        # two calls in a row on one stream: the second reuses the array the first one filled
        got_a = _static_resize(lib, a, 308)
        got_b = _static_resize(lib, b, 308)
        assert torch.equal(got_a, want_a)
        assert torch.equal(got_b, want_b)
        # another geometry rewrites the tables in place (up-scaling 224 <- 240 rows is a different filter length) ...
        c = _frames(13, 4, 240, 320)
        assert torch.equal(_static_resize(lib, c, 224), small.resize_frames(c))
        # ... and the first geometry still gives the first answer afterwards
        assert torch.equal(_static_resize(lib, a, 308), want_a)
        # only one pass needed (no intermediate) and no pass needed (plain copy)
        d = _frames(14, 2, 308, 640)
        one_pass = Engine(ServoConfig(model_type="dinov2_vits14", u_max=640, v_max=308), max_images=2)
        assert torch.equal(_static_resize(lib, d, 308), one_pass.resize_frames(d))
        one_pass.close()
        e = _frames(15, 2, 308, 308)
        assert torch.equal(_static_resize(lib, e, 308), e)

        # warmed up, same geometry as the previous two-pass call but one: the call takes no device memory
        _static_resize(lib, a, 308)
        torch.cuda.synchronize()
        out = torch.empty_like(want_b)
        free_before, _ = torch.cuda.mem_get_info()
        rc = lib.vvs_resize_frames_static(b.data_ptr(), 18, 480, 640, 308, out.data_ptr(),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert rc == 0 and torch.equal(out, want_b)
        assert free_after == free_before
    finally:
        eng.close()
        small.close()


def test_static_resize_refuses_what_does_not_fit():
    lib = _lib.load()
    torch.cuda.set_device(0)
    too_many = torch.zeros(19, 480, 640, 3, dtype=torch.uint8, device="cuda")       # 19 * 480 * 308 * 3 > 8 MiB
    with pytest.raises(_lib.VvsError, match="exceed the built-in scratch of 8388608 bytes"):
        _static_resize(lib, too_many, 308)
    with pytest.raises(_lib.VvsError, match="exceeds the built-in limit of 1024"):
        _static_resize(lib, too_many[:1], 1036)
    out = torch.empty(1, 308, 308, 3, dtype=torch.uint8, device="cuda")
    assert lib.vvs_resize_frames_static(None, 1, 480, 640, 308, out.data_ptr(), None) == -1
    assert b"null argument" in lib.vvs_last_error(None)
    # a refused call leaves the array usable
    f = _frames(16, 1, 480, 640)
    from oracle import pil_resize
    assert np.array_equal(_static_resize(lib, f, 308)[0].cpu().numpy(),
                          pil_resize.resize_bicubic_u8(f[0].cpu().numpy(), 308, 308))

"""ctypes binding of libvitvs_b200.so (include/vitvs_b200.h).  Loads the in-tree library and fails
loudly when it is missing or cannot be used - there is no Python/CPU fallback for any kernel."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VITVS_LIB", os.path.join(HERE, "libvitvs_b200.so"))   # override: experiment builds only

c_i32, c_i64, c_f32, c_f64, c_vp = C.c_int32, C.c_int64, C.c_float, C.c_double, C.c_void_p

F32, BF16 = 0, 1
IMAGE_U8_HWC, IMAGE_F32_NCHW = 0, 1
EPI_BIAS_BF16, EPI_GELU_BF16, EPI_RESID_F32, EPI_PATCH_F32, EPI_SWIGLU_BF16, EPI_SIM, EPI_STORE_F32 = range(7)
# folded-LayerNorm variants (csrc/vvs_internal.h)
EPI_RESID_LN_F32, EPI_PATCH_LN_F32, EPI_BIAS_LN_BF16, EPI_GELU_LN_BF16, EPI_SWIGLU_LN_BF16 = range(7, 12)


class VvsError(RuntimeError):
    pass


class Config(C.Structure):
    _fields_ = [
        ("dim", c_i32), ("heads", c_i32), ("depth", c_i32), ("ffn_hidden", c_i32), ("swiglu", c_i32),
        ("image_size", c_i32), ("max_images", c_i32),
        ("frame_width", c_i32), ("frame_height", c_i32),
        ("f_x", c_f64), ("f_y", c_f64), ("c_x", c_f64), ("c_y", c_f64),
        ("lambda_", c_f64), ("ema_alpha", c_f64),
        ("num_pairs", c_i32), ("layer", c_i32), ("use_feature_binning", c_i32),
        ("same_image_threshold", c_f32),
        ("max_match_tokens", c_i32), ("max_match_dim", c_i32), ("patch_stride", c_i32),
        ("damping_mu", c_f64),
    ]


class MatchOut(C.Structure):
    _fields_ = [(n, c_vp) for n in (
        "nn1", "nn2", "sim1", "candidate", "mean_sim1", "n_candidates", "n_selected", "flags",
        "sel_tok", "sel_sim", "uv_star", "uv", "sel_order")]


class PairOut(C.Structure):
    _fields_ = [("twist", c_vp), ("n_selected", c_vp), ("mean_sim_sel", c_vp), ("match", MatchOut)]


# every symbol include/vitvs_b200.h declares: (restype, argtypes)
SYMBOLS = {
    "vvs_create": (c_i32, [C.POINTER(Config), C.POINTER(c_vp)]),
    "vvs_destroy": (c_i32, [c_vp]),
    "vvs_last_error": (C.c_char_p, [c_vp]),
    "vvs_num_sms": (c_i32, [c_vp]),
    "vvs_build_id": (C.c_char_p, []),
    "vvs_transform_to_real_world": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "vvs_interaction_matrix": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_vp, c_vp]),
    "vvs_set_weight": (c_i32, [c_vp, C.c_char_p, c_vp, c_i32, C.POINTER(c_i64), c_i32, c_vp]),
    "vvs_weights_ready": (c_i32, [c_vp]),
    "vvs_forward_tokens": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp]),
    "vvs_get_tokens": (c_i32, [c_vp, c_vp, c_i32, c_vp]),
    "vvs_descriptors_hierarchy": (c_i32, [c_vp, c_i32, c_i32, c_vp, c_vp]),
    "vvs_forward_facet": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_i32, c_vp]),
    "vvs_facet_descriptors": (c_i32, [c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "vvs_descriptors": (c_i32, [c_vp, c_i32, c_i32, c_vp, c_vp]),
    "vvs_match": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_i32, C.POINTER(MatchOut), c_vp]),
    "vvs_similarity": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "vvs_ibvs": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "vvs_ibvs_terms": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "vvs_frame_pairs_to_twist": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, C.POINTER(PairOut), c_vp]),
    "vvs_frame_pairs_to_twist_host": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "vvs_frame_pairs_to_twist_host_async": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "vvs_frames_to_twist_host_async": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "vvs_wait": (c_i32, [c_vp, c_i32]),
    "vvs_resize_frames": (c_i32, [c_vp, c_vp, c_i32, c_vp, c_vp]),
    "vvs_resize_frames_static": (c_i32, [c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp]),
    "vvs_frames_to_twist": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, C.POINTER(PairOut), c_vp]),
    "vvs_frames_to_twist_host": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp]),
    "vvs_set_goal": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_vp]),
    "vvs_current_to_twist": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, C.POINTER(PairOut), c_vp]),
    "vvs_render_plane": (c_i32, [c_vp, c_vp, c_i32, c_i32, c_f64, c_f64, c_vp, c_i32, c_vp, c_vp, c_vp]),
    "vvs_last_launch_count": (c_i32, [c_vp]),
    "vvs_uses_gram_stencil": (c_i32, [c_vp]),
    "vvs_set_num_pairs": (c_i32, [c_vp, c_i32]),
    "vvs_profile_enable": (c_i32, [c_vp, c_i32]),
    "vvs_trace_step": (c_i32, [c_vp, c_i32, c_i32, C.POINTER(c_i32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), c_i32,
                               C.POINTER(c_i32), c_vp]),
    "vvs_profile_read": (c_i32, [c_vp, C.POINTER(c_f64), C.POINTER(c_i32), c_i32]),
    "vvs_dbg_gemm": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32,
                             c_f32, c_vp, c_i32, c_vp]),
    "vvs_dbg_gemm_ln": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_i32, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32,
                                c_f32, c_vp, c_i32, c_vp, c_vp, c_vp, c_vp, c_f32, c_vp]),
    "vvs_dbg_attention": (c_i32, [c_vp, c_vp, c_vp, c_i32, c_i32, c_i32, c_vp]),
    "vvs_dbg_layernorm": (c_i32, [c_vp, c_vp, c_vp, c_vp, c_vp, c_i32, c_i32, c_vp]),
}

KERNEL_CLASSES = ["other", "im2col", "patch_gemm", "layernorm", "qkv_gemm", "attention", "proj_gemm", "ffn1_gemm",
                  "ffn2_gemm", "descriptors", "sim_gemm", "select", "ibvs", "resize", "stencil"]

_lib = None


def load() -> C.CDLL:
    """dlopen the in-tree library and bind every declared symbol.  Raises VvsError when absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise VvsError(
            f"{LIB_PATH} is missing - build it with `python -m vitvs_b200.build` (needs nvcc). "
            "vitvs_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)       # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(handle, rc: int, what: str):
    if rc != 0:
        msg = load().vvs_last_error(handle)
        raise VvsError(f"{what} failed ({rc}): {msg.decode() if msg else '?'}")

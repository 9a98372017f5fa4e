"""Oracle helper: import the reference's OWN hot-path modules from /root/reference.

TEST INFRASTRUCTURE ONLY.  Works only in the build container (the GPU box has no
/root/reference): callers must check ``available()`` first.  Nothing is copied -
the modules are executed where they lie, with the ROS / matplotlib / timm imports
at the top of vitvs_v2.py:2-42 and dinov2_extractor.py:1-12 stubbed by MagicMock
(SURVEY §8c [probed]).  Used by oracle/make_golden.py and tests/test_oracle_vs_reference.py.
"""
from __future__ import annotations

import importlib
import os
import sys
import types
from unittest.mock import MagicMock

REF_ROOT = os.environ.get("VITVS_REFERENCE_ROOT", "/root/reference")
_SRC = os.path.join(REF_ROOT, "catkin_ws", "ibvs", "src")
_STUBS = [
    "rospy", "gazebo_msgs", "gazebo_msgs.msg", "gazebo_msgs.srv", "geometry_msgs",
    "geometry_msgs.msg", "sensor_msgs", "sensor_msgs.msg", "cv_bridge", "timm",
    "matplotlib", "matplotlib.pyplot", "matplotlib.patches", "tf2_ros", "tf_conversions",
]
_cache: dict = {}


def available() -> bool:
    return os.path.isfile(os.path.join(_SRC, "vitvs_v2.py"))


def _load():
    if _cache:
        return _cache
    if not available():
        raise RuntimeError(f"reference not present under {REF_ROOT}")
    sys.dont_write_bytecode = True          # /root/reference is read-only
    for name in _STUBS:
        sys.modules.setdefault(name, MagicMock(name=name))
    sys.path.insert(0, _SRC)
    try:
        _cache["extractor"] = importlib.import_module("dinov2_extractor")
        _cache["vitvs"] = importlib.import_module("vitvs_v2")
    finally:
        sys.path.remove(_SRC)
    # dino_patch/attention.py imports standalone (xFormers absent => SDPA branch)
    spec = importlib.util.spec_from_file_location(
        "_ref_dino_patch_attention", os.path.join(REF_ROOT, "dino_patch", "attention.py"))
    mod = importlib.util.module_from_spec(spec)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        spec.loader.exec_module(mod)
    _cache["attention"] = mod
    return _cache


def vitvs():
    """The reference module catkin_ws/ibvs/src/vitvs_v2.py."""
    return _load()["vitvs"]


def extractor():
    """The reference module catkin_ws/ibvs/src/dinov2_extractor.py."""
    return _load()["extractor"]


def attention():
    """The reference module dino_patch/attention.py."""
    return _load()["attention"]


def block():
    """The reference module dino_patch/block.py (Block, NestedTensorBlock), executed where it lies.

    block.py:17-20 imports four siblings; only ``attention.py`` is vendored in the reference.  The other three
    (``drop_path``, ``layer_scale``, ``mlp``) belong to the un-vendored upstream package (github facebookresearch/dinov2,
    dinov2/layers/, unpinned) and are supplied here as minimal stand-ins that state the published definitions: DropPath is
    the identity outside training, LayerScale multiplies by a learned per-channel ``gamma``, Mlp is fc1 -> act -> fc2.
    What is pinned by importing block.py itself is everything the reference's file decides: which sub-modules a block
    has, LayerNorm placement, the residual order and the inference branch (block.py:90-95,112-114)."""
    if "block" in _cache:
        return _cache["block"]
    _load()
    import torch
    from torch import nn
    pkg_name = "_ref_dino_patch"
    pkg_dir = os.path.join(REF_ROOT, "dino_patch")
    pkg = types.ModuleType(pkg_name)
    pkg.__path__ = [pkg_dir]                    # `.attention` resolves to the reference's own file
    sys.modules[pkg_name] = pkg

    class DropPath(nn.Module):
        def __init__(self, drop_prob=None):
            super().__init__()
            self.drop_prob = drop_prob

        def forward(self, x):
            assert not self.training, "stand-in: inference only"
            return x

    class LayerScale(nn.Module):
        def __init__(self, dim, init_values=1e-5, inplace=False):
            super().__init__()
            self.gamma = nn.Parameter(init_values * torch.ones(dim))

        def forward(self, x):
            return x * self.gamma

    class Mlp(nn.Module):
        def __init__(self, in_features, hidden_features=None, out_features=None, act_layer=nn.GELU, drop=0.0, bias=True):
            super().__init__()
            self.fc1 = nn.Linear(in_features, hidden_features or in_features, bias=bias)
            self.act = act_layer()
            self.fc2 = nn.Linear(hidden_features or in_features, out_features or in_features, bias=bias)

        def forward(self, x):
            return self.fc2(self.act(self.fc1(x)))

    for sub, attrs in (("drop_path", {"DropPath": DropPath}), ("layer_scale", {"LayerScale": LayerScale}),
                       ("mlp", {"Mlp": Mlp})):
        m = types.ModuleType(f"{pkg_name}.{sub}")
        m.__dict__.update(attrs)
        sys.modules[m.__name__] = m
    spec = importlib.util.spec_from_file_location(f"{pkg_name}.block", os.path.join(pkg_dir, "block.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[spec.name] = mod
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        spec.loader.exec_module(mod)
    _cache["block"] = mod
    return mod


def make_controller(cfg: dict, depth_u16=None):
    """A reference ``Controller`` without ROS: ``Controller.__new__`` + the attributes
    ``load_parameters`` (vitvs_v2.py:272-323) and ``initialize_ema`` (:325-327) would set."""
    C = vitvs().Controller
    c = C.__new__(C)
    c.u_max, c.v_max = cfg["u_max"], cfg["v_max"]
    c.f_x, c.f_y = cfg["f_x"], cfg["f_y"]
    c.c_x, c.c_y = c.u_max / 2, c.v_max / 2
    c.lambda_ = cfg["lambda_"]
    c.num_pairs = cfg["num_pairs"]
    c.dino_input_size = cfg["dino_input_size"]
    c.use_feature_binning = cfg["use_feature_binning"]
    c.ema_alpha = cfg.get("ema_alpha", 0.1)
    c.max_velocity = cfg.get("max_velocity", 1.0)
    c.latest_image_depth = depth_u16
    c.initialize_ema()
    return c


def make_extractor(model, model_type: str):
    """The reference's ViTExtractor wrapped around an injected backbone
    (dinov2_extractor.py:25-39: ``model=`` seam)."""
    return extractor().ViTExtractor(model_type=model_type, stride=14, model=model, device="cpu")

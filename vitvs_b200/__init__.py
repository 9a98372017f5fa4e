"""vitvs_b200 - B200-native (sm_100a) implementation of the ViT-VS `--method dino` servoing hot path:
DINOv2 ViT forward -> patch descriptors -> mutual-NN matching -> IBVS twist.  Hand-written CUDA
(tcgen05 / TMEM / TMA) behind a C ABI (include/vitvs_b200.h); this package is the thin host mirror of the
reference's plugin surface.  No CPU fallback."""
from .config import ARCHS, ServoConfig  # noqa: F401


def __getattr__(name):   # lazy: importing the package must not need torch.cuda / the .so (build step, CPU tests)
    if name in ("Engine", "interpolate_pos_embed"):
        from . import engine
        return getattr(engine, name)
    if name in ("B200ViTExtractor", "find_correspondences_batch", "DinoServoMixin", "DinoServo",
                "set_default_engine"):
        from . import plugin
        return getattr(plugin, name)
    if name == "VvsError":
        from ._lib import VvsError
        return VvsError
    raise AttributeError(name)

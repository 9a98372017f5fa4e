"""Configuration mirror of the reference's YAML (catkin_ws/ibvs/config/config.yaml, read by
Controller.load_parameters, vitvs_v2.py:272-323) plus the constants the reference hard-codes, promoted
to fields with identical defaults (SURVEY §5 "Config / flags")."""
from __future__ import annotations

from dataclasses import dataclass, fields

ARCHS = {  # name -> (dim, heads, depth, ffn_hidden, swiglu)
    "dinov2_vits14": (384, 6, 12, 1536, False),
    "dinov2_vitb14": (768, 12, 12, 3072, False),
    "dinov2_vitl14": (1024, 16, 24, 4096, False),
    "dinov2_vitg14": (1536, 24, 40, 4096, True),
}


@dataclass
class ServoConfig:
    # config.yaml keys consumed by the hot path
    u_max: int = 640
    v_max: int = 480
    f_x: float = 502.3016357421875
    f_y: float = 502.3016357421875
    lambda_: float = 0.03
    num_pairs: int = 24
    dino_input_size: int = 308
    use_feature_binning: bool = True
    ema_alpha: float = 0.8
    max_velocity: float = 1.0
    max_velocity_vector_history: int = 200
    # hard-coded in the reference
    model_type: str = "dinov2_vits14"        # vitvs_v2.py:250
    stride: int = 14                          # vitvs_v2.py:250
    layer: int = 11                           # vitvs_v2.py:484
    facet: str = "token"                      # vitvs_v2.py:485
    distance_threshold: float = 1.0           # vitvs_v2.py:72
    same_image_threshold: float = 0.99        # vitvs_v2.py:84
    # not in the reference: damping of the control law, v = -lambda (L^T L + mu I)^-1 L^T e; 0 = the reference's pinv(L)
    damping_mu: float = 0.0

    @property
    def c_x(self) -> float:                   # vitvs_v2.py:282
        return self.u_max / 2

    @property
    def c_y(self) -> float:                   # vitvs_v2.py:283
        return self.v_max / 2

    @classmethod
    def from_yaml(cls, path: str, **overrides) -> "ServoConfig":
        import yaml
        with open(path, "r") as fh:
            raw = yaml.safe_load(fh)
        known = {f.name for f in fields(cls)}
        kw = {k: v for k, v in raw.items() if k in known}
        kw.update(overrides)
        return cls(**kw)

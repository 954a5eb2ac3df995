from .tiny_ecg import TinyECG, Tiny1D, FlatParams, param_offsets, param_shapes, num_params  # noqa: F401


def build_model(name: str, num_classes: int = 2):
    """Model registry: ``tiny_ecg`` (reference model) or ``resnet1d_wide`` / ``resnet1d_small``."""
    if name in ("tiny_ecg", "tinyecg", "TinyECG"):
        return TinyECG(num_classes=num_classes)
    from . import resnet1d_wide as _r
    if name == "resnet1d_wide":
        return _r.resnet1d_wide_25m(num_classes)
    if name == "resnet1d_small":
        return _r.resnet1d_small(num_classes)
    raise ValueError(f"unknown model {name!r}")

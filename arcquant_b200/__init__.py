"""arcquant_b200 -- B200-native (sm_100a) NVFP4 + Augmented-Residual-Channel hot path.

``arcquant_b200.agemm`` mirrors the reference's ``agemm`` extension module; ``import agemm`` (the
top-level shim at the repo root) gives reference callers the same module name.
"""
from . import agemm  # noqa: F401

__version__ = "0.1.0"

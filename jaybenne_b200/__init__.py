"""jaybenne_b200 — B200-native implementation of jaybenne's IMC/DDMC hot path.

CUDA kernels (sm_100a) behind a C ABI (include/jaybenne_b200.h, built into
jaybenne_b200/libjaybenne_b200.so), plus a thin host mirror of the reference's package
interface (package.py) and driver application (mcblock.py). There is no CPU fallback.
"""
from . import capi  # noqa: F401
from .capi import JbError, device_api, load_library  # noqa: F401
from .mcblock import Mcblock, Mesh, ParameterInput, uniform_deck_text  # noqa: F401
from .package import Jaybenne, default_config, radiation_step_hosted  # noqa: F401

__all__ = ["capi", "JbError", "device_api", "load_library", "Mcblock", "Mesh", "ParameterInput",
           "uniform_deck_text", "Jaybenne", "default_config", "radiation_step_hosted"]

"""The drop-in boundary: libjaybenne_b200.so loads without a GPU, exports every symbol
include/jaybenne_b200.h declares, and its compute entry points refuse to run without a CUDA
device (there is no CPU fallback). The deck / mesh host helpers are plain C++ and run here.
"""
import ctypes as C
import os
import re

import pytest

from jaybenne_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "jaybenne_b200.h")


def header_functions():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(jb_[a-z0-9_]+)\s*\(", text)))


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    names = header_functions()
    assert len(names) >= 50
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_python_binding_covers_the_header():
    assert set(capi.declared_symbols()) == set(header_functions())


def test_abi_version_and_struct_sizes():
    api = capi.device_api()
    assert api.abi_version() == 1
    # the ctypes mirrors must match the C layout the header promises
    assert C.sizeof(capi.jb_block) == 8 + 24 + 24 + 24 + 8
    assert C.sizeof(capi.jb_event) == 96
    assert int(api.exchange_record_bytes()) == 96
    assert C.sizeof(capi.jb_cycle_stats) == 11 * 8 + 8 + 7 * 8


def test_config_default_matches_reference_defaults():
    """<jaybenne> defaults, src/jaybenne/jaybenne.cpp:163-223."""
    cfg = capi.jb_config()
    capi.device_api().config_default(C.byref(cfg))
    assert cfg.dt == pytest.approx(1.7976931348623157e308)
    assert cfg.min_swarm_occupancy == 0.0
    assert cfg.unique_rank_seeds == 1 and cfg.seed == 123
    assert cfg.max_transport_iterations == 10000
    assert cfg.use_ddmc == 0 and cfg.tau_ddmc == 5.0
    assert cfg.source_strategy == capi.JB_STRATEGY_UNIFORM
    assert cfg.do_emission == 1 and cfg.do_feedback == 1
    assert cfg.rng_mode == capi.JB_RNG_PHILOX and cfg.gray_mode == capi.JB_GRAY_KIRCHHOFF


@pytest.mark.skipif(has_gpu(), reason="checks the no-device behaviour")
def test_compute_fails_loudly_without_a_device():
    api = capi.device_api()
    cfg = capi.jb_config()
    api.config_default(C.byref(cfg))
    cfg.num_particles = 10
    ctx = capi.ctx_p()
    assert api.create(C.byref(cfg), C.byref(ctx)) == capi.JB_ERR_NO_DEVICE
    assert not ctx.value


def test_create_validates_like_jaybenne_initialize():
    api = capi.device_api()
    cfg = capi.jb_config()
    api.config_default(C.byref(cfg))
    cfg.num_particles = 10
    cfg.min_swarm_occupancy = 1.0           # jaybenne.cpp:171-172: must be in [0, 1)
    ctx = capi.ctx_p()
    assert api.create(C.byref(cfg), C.byref(ctx)) == capi.JB_ERR_INVALID
    cfg.min_swarm_occupancy = 0.0
    cfg.rng_mode = capi.JB_RNG_XORSHIFT_POOL  # oracle-only generator
    assert api.create(C.byref(cfg), C.byref(ctx)) == capi.JB_ERR_UNSUPPORTED


def test_package_has_no_oracle_dependency():
    """Nothing under jaybenne_b200/ may import, link or name the oracle."""
    pkg = os.path.join(ROOT, "jaybenne_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath or "__pycache__" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", "Makefile")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                for needle in ("libjb_oracle", "oracle_binding", "jb_oracle.cpp", "oracle/", "import oracle"):
                    assert needle not in text, (needle, os.path.join(dirpath, f))
    # and the shared library itself carries no oracle entry point
    lib = capi.load_library()
    assert not hasattr(lib, "jbo_create") and not hasattr(lib, "jbo_radiation_step")

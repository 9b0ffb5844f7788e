"""TEST INFRASTRUCTURE: loads the CPU oracle (oracle/build, oracle/_ref) and exposes it
through the same `Api` shape as the CUDA library so the package's host code can drive it.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this."""
import ctypes as C
import os
import subprocess

from jaybenne_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
FLAVOURS = {
    "portable": os.path.join(ORACLE_DIR, "build", "libjb_oracle.so"),      # restated + jb_math.h (GPU twin)
    "std": os.path.join(ORACLE_DIR, "build", "libjb_oracle_std.so"),       # restated + libm
    "ref": os.path.join(ORACLE_DIR, "_ref", "libjb_oracle_ref.so"),        # reference headers + libm
    "fast": os.path.join(ORACLE_DIR, "build", "libjb_oracle_fast.so"),     # -O3 -march=native timing build
}
_cache = {}

PROBES = {
    "flavour": (C.c_char_p, []),
    "max_threads": (C.c_int32, []),
    "set_num_threads": (None, [C.c_int32]),
    "config_default": (None, [C.POINTER(capi.jb_config)]),
    "probe_absorption": (C.c_double, [capi.ctx_p, C.c_double, C.c_double, C.c_double]),
    "probe_scattering": (C.c_double, [capi.ctx_p, C.c_double, C.c_double, C.c_double]),
    "probe_emissivity": (C.c_double, [capi.ctx_p, C.c_double, C.c_double]),
    "probe_temperature": (C.c_double, [capi.ctx_p, C.c_double, C.c_double]),
    "probe_specific_heat": (C.c_double, [capi.ctx_p, C.c_double, C.c_double]),
    "probe_eps": (C.c_double, [C.c_int]),
    "probe_philox": (None, [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int32,
                            C.POINTER(C.c_double)]),
    "probe_philox_raw": (None, [C.POINTER(C.c_uint32), C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "probe_xorshift": (None, [C.c_uint64, C.c_int32, C.POINTER(C.c_double)]),
    "probe_imc_step": (C.c_int32, [C.c_uint64, C.c_uint32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                   C.c_double, C.c_double, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                   C.POINTER(C.c_double), C.POINTER(C.c_int32)]),
    "probe_ddmc_step": (C.c_int32, [C.c_uint64, C.c_uint32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double,
                                    C.c_double, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double),
                                    C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int32),
                                    C.POINTER(C.c_int32)]),
    "probe_planck": (C.c_double, [C.c_uint64, C.c_uint32, C.c_double, C.c_double]),
    "math": (C.c_double, [C.c_int32, C.c_double]),
    "math_array": (None, [C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_double), C.c_int64]),
}


def build():
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR])
    if os.path.isdir("/root/reference/src/jaybenne"):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref"])


def available(flavour):
    return os.path.exists(FLAVOURS[flavour])


def oracle_api(flavour="portable"):
    if flavour not in _cache:
        path = FLAVOURS[flavour]
        if not os.path.exists(path):
            build()
        lib = C.CDLL(path)
        _cache[flavour] = capi.Api(lib, "jbo_", (capi.COMPUTE_API, PROBES))
    return _cache[flavour]

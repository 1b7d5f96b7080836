import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def emul():
    """CPU emulation of the LSB_HD kernel bodies (tests/emul/emul.cpp) — test infrastructure."""
    import ctypes as C
    so = os.path.join(ROOT, "tests", "_build", "libemul.so")
    src = os.path.join(ROOT, "tests", "emul", "emul.cpp")
    hdrs = [os.path.join(ROOT, "lattisense_b200", "csrc", h) for h in os.listdir(os.path.join(ROOT, "lattisense_b200", "csrc")) if h.endswith(".h")]
    newest = max(os.path.getmtime(p) for p in [src] + hdrs)
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        os.makedirs(os.path.dirname(so), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off",
                               "-I/usr/local/cuda/include", "-o", so, src])
    lib = C.CDLL(so)
    u64p = C.POINTER(C.c_uint64)
    i32p = C.POINTER(C.c_int)
    lib.emul_ntt.argtypes = [C.c_int, C.c_uint64, u64p, C.c_int, C.c_int]
    lib.emul_ntt_cluster.argtypes = [C.c_int, C.c_uint64, u64p, C.c_int, C.c_int]
    lib.emul_ntt_persist.argtypes = [C.c_int, C.c_uint64, u64p, C.c_int, C.c_int]
    lib.emul_basis_extend.argtypes = [C.c_int, u64p, C.c_int, i32p, C.c_int, i32p, C.c_int, C.c_int, u64p, u64p]
    lib.emul_bex_mixed_targets.argtypes = [u64p, C.c_int, i32p, C.c_int, i32p, C.c_int]
    lib.emul_reduce_wide.restype = C.c_uint64
    lib.emul_reduce_wide.argtypes = [C.c_uint64] * 3
    for n in ("emul_mulmod", "emul_barrett64", "emul_mul_shoup"):
        getattr(lib, n).restype = C.c_uint64
    lib.emul_mulmod.argtypes = [C.c_uint64] * 3
    lib.emul_barrett64.argtypes = [C.c_uint64] * 2
    lib.emul_mul_shoup.argtypes = [C.c_uint64] * 3
    return lib

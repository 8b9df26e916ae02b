"""CPU check of vector_b200/csrc/npquant.h: the C++ statement of numpy's float32 percentile index arithmetic that
b200vec_normalize_db runs on the host must agree with hostmath.percentile_ranks / lerp (themselves pinned to
np.percentile by tests/test_hostmath.py) — ranks and gamma bit for bit, including above 2^24 elements."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from vector_b200 import hostmath as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("npquant") / "npquant.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", out,
                    os.path.join(ROOT, "tests", "native", "npquant_host.cpp")], check=True)
    lib = C.CDLL(out)
    lib.b2host_pct_ranks.argtypes = [C.c_longlong, C.c_double, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong),
                                     C.POINTER(C.c_float)]
    lib.b2host_lerp.restype = C.c_float
    lib.b2host_lerp.argtypes = [C.c_float, C.c_float, C.c_float]
    return lib


def test_ranks_and_gamma_match_hostmath(lib):
    rng = np.random.default_rng(7)
    counts = [1, 2, 3, 10, 999, 1000, 10_234_880, 16_777_216, 16_777_217, 20_000_001, 64_611_328, 111_998_976,
              2_000_000_000] + [int(x) for x in rng.integers(1, 300_000_000, 200)]
    for n in counts:
        for q in (0, 5, 5.0, 10.0, 25, 50, 95.0, 99.9, 100):
            lo, hi, g = C.c_longlong(), C.c_longlong(), C.c_float()
            lib.b2host_pct_ranks(n, float(q), C.byref(lo), C.byref(hi), C.byref(g))
            want = H.percentile_ranks(n, q, np.float32)
            assert (lo.value, hi.value) == (want[0], want[1]), (n, q)
            assert np.float32(g.value) == want[2] and want[2].dtype == np.float32, (n, q)


def test_lerp_matches_hostmath(lib):
    rng = np.random.default_rng(8)
    a = rng.standard_normal(2000).astype(np.float32) * 50
    b = a + np.abs(rng.standard_normal(2000).astype(np.float32))
    t = rng.uniform(0, 1, 2000).astype(np.float32)
    t[:4] = [0.0, 0.5, 1.0, np.float32(0.49999997)]
    for x, y, w in zip(a, b, t):
        assert np.float32(lib.b2host_lerp(x, y, w)) == H.lerp(x, y, w)

"""b200vec_power_stats_static — the allocation-free variant of b200vec_power_stats (include/b200vec.h).

No counterpart in the reference: the entry point exists for callers that may not call cudaMalloc.  Its scratch is a
file-scope device array inside libb200vec.so, shared by every call and keeping the last call's contents, so the
checks here are made on calls IN A ROW (a larger input first, then smaller ones whose stale partials would show),
against the existing entry point through the same C-ABI.  Collected last on purpose (pytest -x).
"""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu

MAX_SAMPLES = 1 << 30          # kStaticMaxSamples in vector_b200/csrc/b200vec.cu
SUM_RTOL = 1e-12               # both sums are fp64; they differ only in the order of the additions
SIZES = [5_000_003, 1, 65_536, 65_537, 1_000_003, 300, 5_000_003]


def _lib():
    from vector_b200 import _lib as L
    return L.load()


def _signal(n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(n, 2, generator=g, device="cuda", dtype=torch.float32) * 0.37
    return torch.view_as_complex(x).contiguous()


def _existing(lib, ctx, x, pre_scale):
    st = (C.c_double * 2)()
    rc = lib.b200vec_power_stats(ctx, x.data_ptr(), x.numel(), C.c_float(pre_scale), st, None)
    assert rc == 0, lib.b200vec_last_error().decode()
    return float(st[0]), float(st[1])


def _static(lib, x, pre_scale, out, stream=None):
    rc = lib.b200vec_power_stats_static(0, x.data_ptr() if x.numel() else None, x.numel(), C.c_float(pre_scale),
                                        out.data_ptr(), stream)
    assert rc == 0, lib.b200vec_last_error().decode()


@pytest.fixture(scope="module")
def ctx():
    lib = _lib()
    h = C.c_void_p()
    assert lib.b200vec_create(0, C.byref(h)) == 0, lib.b200vec_last_error().decode()
    yield h
    lib.b200vec_destroy(h)


def test_static_power_stats_matches_existing_entry_on_calls_in_a_row(ctx):
    lib = _lib()
    out = torch.full((2,), -1.0, device="cuda", dtype=torch.float64)
    inputs = [(_signal(n, 100 + i), 1.0 if i % 2 == 0 else 0.731) for i, n in enumerate(SIZES)]
    want = [_existing(lib, ctx, x, ps) for x, ps in inputs]
    torch.cuda.synchronize()
    for rep in range(2):                                   # the whole series twice: every call follows another one
        for (x, ps), (mx, sm) in zip(inputs, want):
            out.fill_(-1.0)
            _static(lib, x, ps, out)
            got = out.cpu().numpy()                        # synchronises
            assert got[0] == mx, (x.numel(), rep)          # max |y|^2: bit for bit
            assert abs(got[1] - sm) <= SUM_RTOL * abs(sm), (x.numel(), rep, got[1], sm)


def test_static_power_stats_same_input_twice_is_reproducible():
    lib = _lib()
    x = _signal(3_333_333, 7)
    out = torch.zeros(2, device="cuda", dtype=torch.float64)
    res = []
    for _ in range(2):
        out.fill_(-1.0)
        _static(lib, x, 1.0, out)
        res.append(out.cpu().numpy().copy())
    assert res[0].tobytes() == res[1].tobytes()            # fixed summation order: identical bits
    # and against numpy on the host (float64 accumulation of float32 |y|^2): the same 1e-7 the mat2wv tests use
    xh = x.cpu().numpy()
    a = np.abs(xh).astype(np.float32)
    assert abs(res[0][1] - float(np.sum((a * a).astype(np.float64)))) <= 1e-6 * res[0][1]


def test_static_power_stats_empty_and_limit():
    lib = _lib()
    out = torch.full((2,), 5.0, device="cuda", dtype=torch.float64)
    rc = lib.b200vec_power_stats_static(0, None, 0, C.c_float(1.0), out.data_ptr(), None)
    assert rc == 0
    assert out.cpu().tolist() == [0.0, 0.0]
    # beyond the static scratch: refused before anything is launched (the pointer is never dereferenced)
    out.fill_(5.0)
    rc = lib.b200vec_power_stats_static(0, out.data_ptr(), MAX_SAMPLES + 1, C.c_float(1.0), out.data_ptr(), None)
    assert rc != 0
    msg = lib.b200vec_last_error().decode()
    assert "2^30" in msg and "static scratch" in msg
    assert out.cpu().tolist() == [5.0, 5.0]
    assert lib.b200vec_power_stats_static(0, out.data_ptr(), 16, C.c_float(1.0), None, None) != 0
    assert lib.b200vec_power_stats_static(99, out.data_ptr(), 16, C.c_float(1.0), out.data_ptr(), None) != 0


def test_static_power_stats_allocates_nothing():
    lib = _lib()
    x = _signal(2_000_000, 11)
    out = torch.zeros(2, device="cuda", dtype=torch.float64)
    _static(lib, x, 1.0, out)                              # first call on the device: set-up
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(20):
        _static(lib, x, 1.0, out)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0

"""CPU check of the max-pass planner (vector_b200/csrc/maxspans.cuh, host code of libb200vec): the spans it
hands to the kernels must cover every sample of the ranges that can hold max|v| — samples under two or more
instances, or under an instance of a packet without a free-standing instance — and nothing outside the ranges;
with them, max over spans U packet maxima equals the max of the whole vector (unified_gui.py:1778-1779)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from vector_b200 import hostmath as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def planner(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("maxspans") / "maxspans.so")
    src = os.path.join(ROOT, "tests", "native", "maxspans_host.cpp")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", out, src], check=True)
    lib = C.CDLL(out)
    lib.b2host_max_spans.restype = C.c_longlong
    lib.b2host_max_spans.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_longlong]
    lib.b2host_max_eval_ranges.argtypes = [C.c_void_p, C.c_int, C.c_longlong, C.c_void_p]
    return lib


def _spans(lib, pk, ranges):
    pk = np.ascontiguousarray(pk, np.int64).reshape(-1, 5)
    rg = np.ascontiguousarray(ranges, np.int64).reshape(-1)
    cap = 1 << 20
    out = np.empty((cap, 2), np.int64)
    n = lib.b2host_max_spans(pk.ctypes.data, len(pk), rg.ctypes.data, len(rg) // 2, out.ctypes.data, cap)
    assert n <= cap
    return out[:n]


def _scenario(rng, total):
    np_ = int(rng.integers(1, 6))
    rows, lengths = [], []
    for _ in range(np_):
        length = int(rng.integers(20, 900))
        period = int(rng.integers(max(8, length // 3), 3000))
        start = int(rng.integers(0, 2000))
        rows.append((start, period, H.instance_count(start, period, length, total), 0))
        lengths.append(length)
    return rows, lengths


@pytest.mark.parametrize("seed", range(40))
def test_spans_cover_what_can_hold_the_max(planner, seed):
    rng = np.random.default_rng(seed)
    total = int(rng.integers(3000, 20000))
    rows, lengths = _scenario(rng, total)
    flags = H.free_instance_flags(rows, lengths, total)
    if seed % 3 == 0:
        flags = [False] * len(rows)
    cov = np.zeros(total, np.int32)
    unf = np.zeros(total, np.int32)
    vec = np.zeros(total, np.float64)
    pmax = []
    for (s, p, k, _), n, fl in zip(rows, lengths, flags):
        y = rng.standard_normal(n)
        pmax.append(np.abs(y).max() if (fl and k > 0) else 0.0)
        for i in range(k):
            cov[s + i * p:s + i * p + n] += 1
            vec[s + i * p:s + i * p + n] += y
            if not fl:
                unf[s + i * p:s + i * p + n] += 1
    ranges = H.max_eval_ranges(rows, lengths, total)
    if seed % 4 == 1:                                              # an arbitrary sub-range pair
        a, b = sorted(int(x) for x in rng.integers(0, total, 2))
        ranges = [(0, a), (b, total)]
    pk = [[s, p, k, n, int(fl)] for (s, p, k, _), n, fl in zip(rows, lengths, flags)]
    sp = _spans(planner, pk, ranges)
    in_range = np.zeros(total, bool)
    for lo, hi in ranges:
        in_range[lo:hi] = True
    covered = np.zeros(total, bool)
    for s0, n in sp:
        assert 1 <= n <= 512 and s0 >= 0 and s0 + n <= total
        assert in_range[s0:s0 + n].all()                           # never outside the caller's ranges
        assert not covered[s0:s0 + n].any()                        # disjoint
        covered[s0:s0 + n] = True
    must = in_range & ((cov >= 2) | (unf >= 1))
    assert covered[must].all()
    # the planner's promise: max over spans U flagged packet maxima == max over the ranges
    m_spans = np.abs(vec[covered]).max() if covered.any() else 0.0
    m_ref = np.abs(vec[in_range]).max() if in_range.any() else 0.0
    if ranges == H.max_eval_ranges(rows, lengths, total):
        assert max([m_spans] + pmax) == np.abs(vec).max()
    else:
        assert m_spans <= m_ref and max([m_spans] + pmax) >= m_ref


def test_baseline_config_needs_few_spans(planner):
    """cfg2: 1 238 of the 16 298 half-frames of the max ranges lie under two instances."""
    total, n = 56_000_000, 56_100
    rows = [(0, 5600000, 10, 0), (559900, 2800000, 20, 0), (1119900, 1120000, 49, 0), (1679900, 560000, 97, 0),
            (2239900, 280000, 192, 0), (2799900, 112000, 475, 0)]
    ranges = H.max_eval_ranges(rows, [n] * 6, total)
    sp = _spans(planner, [[s, p, k, n, 1] for s, p, k, _ in rows], ranges)
    assert 0 < len(sp) < 2000
    sp0 = _spans(planner, [[s, p, k, n, 0] for s, p, k, _ in rows], ranges)     # no flags: everything under an instance
    assert len(sp0) > 5 * len(sp)


@pytest.mark.parametrize("seed", range(30))
def test_max_eval_ranges_matches_hostmath(planner, seed):
    rng = np.random.default_rng(1000 + seed)
    total = int(rng.integers(3000, 60000))
    rows, lengths = _scenario(rng, total)
    if seed % 2:
        rows = [(s, p - p % 50 + 50, k, 0) for s, p, k, _ in rows]     # shared factors: a common period often exists
        rows = [(s, p, H.instance_count(s, p, n, total), 0) for (s, p, _, _), n in zip(rows, lengths)]
    pk = np.ascontiguousarray([[s, p, k, n, 0] for (s, p, k, _), n in zip(rows, lengths)], np.int64)
    out = np.zeros(4, np.int64)
    nr = planner.b2host_max_eval_ranges(pk.ctypes.data, len(pk), total, out.ctypes.data)
    got = [(int(out[2 * r]), int(out[2 * r + 1])) for r in range(nr)]
    assert got == H.max_eval_ranges(rows, lengths, total)

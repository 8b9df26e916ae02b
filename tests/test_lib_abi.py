"""The C-ABI library builds, loads and exports every symbol include/b200vec.h declares;
without a GPU it refuses to create a context (no CPU fallback).  CPU only, no compute."""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    from vector_b200 import _lib
    return _lib.load()


def test_every_declared_symbol_is_exported(lib):
    from vector_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "b200vec.h")).read()
    declared = set(re.findall(r"\b(b200vec_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    nm = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    for sym in declared:
        assert re.search(rf"\bT {sym}\b", nm), sym
        assert getattr(lib, sym) is not None


def test_struct_layout_matches_header():
    from vector_b200._lib import BatchPkt, PktDesc
    assert C.sizeof(BatchPkt) == 40 and BatchPkt.freq_hz.offset == 32 and BatchPkt.src_slot.offset == 24
    assert C.sizeof(PktDesc) == 32
    assert (PktDesc.start.offset, PktDesc.period.offset, PktDesc.count.offset, PktDesc.slot.offset) == (0, 8, 16, 24)


def test_sass_is_sm100a_only(lib):
    from vector_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    ctx = C.c_void_p()
    assert lib.b200vec_create(0, C.byref(ctx)) != 0
    assert b"no CUDA device" in lib.b200vec_last_error()
    import vector_b200 as vb
    import numpy as np
    from vector_b200._lib import B200VecError
    with pytest.raises(B200VecError, match="no CPU fallback"):
        vb.apply_frequency_shift(np.ones(8, np.complex64), 1e6, 56e6)
    with pytest.raises(B200VecError):
        vb.create_spectrogram(np.ones(8000, np.complex64), 56e6)
    assert vb.apply_frequency_shift(np.ones(8), 0, 56e6) is not None     # identity branch needs no GPU


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "vector_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), f
                assert "ref_port" not in src and "/root/reference" not in src, f


def test_fused_kernel_code_size_below_the_cliff():
    """profiles/README.md, "Code-size cliff": every build whose fused wf1024_kernel instantiation exceeded
    64 KB of SASS ran the unchanged step 40 us slower.  Keep it below, with headroom."""
    import re, shutil, subprocess
    from vector_b200._lib import LIB_PATH
    if shutil.which("cuobjdump") is None or not os.path.exists(LIB_PATH):
        pytest.skip("cuobjdump or the built library is not available")
    sass = subprocess.run(["cuobjdump", "-sass", LIB_PATH], capture_output=True, text=True, timeout=300).stdout
    inside, last = False, None
    for line in sass.splitlines():
        if "Function :" in line:
            if inside:
                break
            inside = "wf1024_kernelILi4ELi0ELb1ELb0" in line
        elif inside:
            m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/", line)
            if m:
                last = int(m.group(1), 16)
    assert last is not None, "fused wf1024_kernel<4,0,true,false> not found in the library"
    assert last + 16 <= 62 * 1024, f"fused kernel is {last + 16} bytes of SASS (cliff at 65536)"

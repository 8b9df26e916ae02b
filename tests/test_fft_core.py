"""CPU check of the register FFT core the nfft-1024 kernels run (vector_b200/csrc/fft32.cuh).

The header compiles as plain C++; tests/native/fft_core_host.cpp walks the very same code lane
by lane (fmaf = FFMA, same operation order), so the arithmetic the GPU executes is checked here
without a GPU: against scipy's float32 spectrogram (the reference's own call,
vector_analyzer/spectrogram_analysis.py:40-49) and against an fp64 truth, with the parity
rule of SURVEY.md §8c.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
import scipy.signal

from oracle import ref_port as O
from tests import parity as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def core(tmp_path_factory):
    out = str(tmp_path_factory.mktemp("fftcore") / "fftcore.so")
    src = os.path.join(ROOT, "tests", "native", "fft_core_host.cpp")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", out, src],
                   check=True)
    lib = C.CDLL(out)
    lib.b2host_stft1024.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_float, C.c_void_p]
    return lib


def _host_stft(lib, x, hop, window, scale):
    x = np.ascontiguousarray(x, dtype=np.complex64)
    nf = (x.size - 1024) // hop + 1
    w32 = np.ascontiguousarray(window, dtype=np.float32)
    psd = np.empty((nf, 1024), np.float32)
    lib.b2host_stft1024(x.ctypes.data, hop, nf, w32.ctypes.data, np.float32(np.sqrt(np.float32(scale))),
                        psd.ctypes.data)
    return np.fft.fftshift(psd.T, axes=0)


def test_core_matches_scipy_on_a_built_vector(core):
    fs = 56e6
    cfgs = O._rows_to_configs([("tone", 2, 1.0, 0.2), ("chirp", -3, 1.2, 0.6), ("bpsk", 1, 1.5, 0.0)])
    vec = O.build_vector(O.packets_for(cfgs), cfgs, 280_000, True)[0]
    win = scipy.signal.get_window("hann", 1024)
    scale = 1.0 / (fs * (win * win).sum())
    S = _host_stft(core, vec, 512, win, scale)
    _, _, Sr = O.full_rate_spectrogram(vec, fs)
    _, _, St = O.stft_psd_f64(vec, fs, "hann", 1024, 512, 1024, "density")
    assert S.shape == Sr.shape
    assert P.spec_db_error(S, Sr) <= P.SPEC_DB
    ours, refs = P.band_errors(S, St), P.band_errors(Sr, St)
    for band, (emax, e99) in ours.items():
        if band in refs:
            assert e99 <= 2 * refs[band][1] + 1e-4, (band, e99, refs[band])
    zero_cols = ~np.any(Sr, axis=0)
    assert zero_cols.any() and not np.any(S[:, zero_cols])


@pytest.mark.parametrize("hop,win", [(512, "hann"), (256, "blackmanharris"), (34, "boxcar")])
def test_core_against_fp64_dft(core, hop, win):
    rng = np.random.default_rng(hop)
    x = (rng.standard_normal(1024 + 7 * hop) + 1j * rng.standard_normal(1024 + 7 * hop)).astype(np.complex64)
    w = scipy.signal.get_window(win, 1024)
    S = _host_stft(core, x, hop, w, 1.0)
    w32 = w.astype(np.float32).astype(np.float64)
    for f in range(S.shape[1]):
        X = np.fft.fftshift(np.fft.fft(x[f * hop: f * hop + 1024].astype(np.complex128) * w32))
        want = np.abs(X) ** 2
        assert np.max(np.abs(S[:, f] - want)) <= 3e-6 * want.max()


def test_core_single_bin_and_impulse(core):
    n = np.arange(1024)
    ones = np.ones(1024)
    for k in (0, 1, 31, 32, 33, 511, 512, 1023):
        x = np.exp(2j * np.pi * k * n / 1024).astype(np.complex64)
        S = np.fft.fftshift(_host_stft(core, x, 512, ones, 1.0)[:, 0])      # undo the shift: natural order
        assert np.argmax(S) == k and abs(S[k] - 1024.0 ** 2) <= 1e-5 * 1024.0 ** 2
        assert np.max(np.delete(S, k)) <= 1e-6 * 1024.0 ** 2
    for pos in (0, 1, 32, 517, 1023):
        x = np.zeros(1024, np.complex64)
        x[pos] = 2.0 - 1.0j
        S = _host_stft(core, x, 512, ones, 1.0)[:, 0]
        assert np.allclose(S, 5.0, rtol=1e-6)

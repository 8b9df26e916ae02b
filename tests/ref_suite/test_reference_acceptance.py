"""The reference's own acceptance tests for the vector-build path, restated over vector_b200's names
("must pass unchanged with the new engine installed under the utils names").

Every case names the reference test it restates (file:line under the reference root).  Each runs twice:
  impl = "oracle"  the CPU restatement — shows the restated inequality is the reference's (CPU, always runs);
  impl = "engine"  vector_b200 (GPU, C-ABI) — the acceptance run proper; where both produce arrays the engine
                   must also agree with the oracle to the path's tolerances.
Plot-only tests (tests/test_utils.py:86-160) need matplotlib: unavailable on this box.
"""
import os

import numpy as np
import pytest
import scipy.io as sio

from oracle import ref_port as O
from tests import parity as P

torch = pytest.importorskip("torch")


class _Oracle:
    apply_frequency_shift = staticmethod(O.apply_frequency_shift)
    create_spectrogram = staticmethod(O.create_spectrogram)
    normalize_spectrogram = staticmethod(O.normalize_spectrogram)
    save_vector = staticmethod(O.save_vector)
    save_vector_wv = staticmethod(O.save_vector_wv)
    load_packet_info = staticmethod(O.load_packet_info)

    @staticmethod
    def load_packet(path):
        return O.load_packet_info(path)[0]

    @staticmethod
    def build_vector(configs, total, normalize, packets):
        v, mk, counts, _ = O.build_vector(packets, configs, total, normalize)
        return v, mk, counts


@pytest.fixture(params=["oracle", pytest.param("engine", marks=pytest.mark.gpu)])
def impl(request):
    if request.param == "oracle":
        return _Oracle
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import vector_b200 as vb
    return vb


def _burst(sr, duration, tones, noise=0.0, seed=0):
    """hanning-windowed sum of complex tones: the test signal of tests/test_spectrogram_cleanliness.py."""
    t = np.linspace(0, duration, int(sr * duration), endpoint=False)
    x = np.hanning(len(t)) * sum(a * np.exp(2j * np.pi * f * t) for a, f in tones)
    if noise:
        rng = np.random.default_rng(seed)
        x = x + noise * (rng.standard_normal(len(t)) + 1j * rng.standard_normal(len(t)))
    return t, x


def _db(impl, x, sr):
    f, ts, S = impl.create_spectrogram(x, sr, time_resolution_us=5)
    db, vmin, vmax = impl.normalize_spectrogram(S)
    return f, ts, S, db, vmin, vmax


def _agree(impl, x, sr):
    """engine vs oracle on the same input in the path's dtype (complex64 — what load_packet_info and the vector
    build hand to create_spectrogram): spectrogram 0.01 dB on the >= -80 dBc bins, colour limits 0.01 dB."""
    if impl is _Oracle:
        return
    x = x.astype(np.complex64)
    f, ts, S, db, vmin, vmax = _db(impl, x, sr)
    fr, tr, Sr, dbr, vminr, vmaxr = _db(_Oracle, x, sr)
    assert S.shape == Sr.shape and S.dtype == Sr.dtype and np.array_equal(f, fr) and np.array_equal(ts, tr)
    assert P.spec_db_error(S, Sr) <= P.SPEC_DB
    assert abs(vmin - vminr) <= 0.01 and abs(vmax - vmaxr) <= 0.01


# ---------------------------------------------------------------- tests/test_spectrogram_cleanliness.py:25-242
def test_dynamic_range_limits(impl):                                     # :25-50
    _, x = _burst(1_000_000, 0.0005, [(1.0, 100_000)])
    *_, vmin, vmax = _db(impl, x, 1_000_000)
    assert 20 <= vmax - vmin <= 35
    _agree(impl, x, 1_000_000)


def test_spectrogram_shape_adequacy(impl):                               # :52-76
    _, x = _burst(1_000_000, 0.001, [(1.0, 150_000)])
    _, _, S = impl.create_spectrogram(x, 1_000_000, time_resolution_us=5)
    assert S.shape[1] >= 50 and S.shape[0] >= 200
    _agree(impl, x, 1_000_000)


def test_tone_burst_clarity(impl):                                       # :78-113
    _, x = _burst(1_000_000, 0.0005, [(1.0, 100_000)])
    f, _, _, db, _, _ = _db(impl, x, 1_000_000)
    k = int(np.argmin(np.abs(f - 100_000)))
    others = np.concatenate([db[:k - 10, :], db[k + 10:, :]])
    assert np.mean(db[k, :]) - np.mean(others) >= 15


def test_multi_tone_separation(impl):                                    # :115-161
    _, x = _burst(1_000_000, 0.0005, [(0.7, 80_000), (0.5, 120_000), (0.3, 160_000)])
    f, _, _, db, _, _ = _db(impl, x, 1_000_000)
    idx = [int(np.argmin(np.abs(f - fr))) for fr in (80_000, 120_000, 160_000)]
    background = np.mean(db[0:idx[0] - 20, :])
    snr = [np.mean(db[i, :]) - background for i in idx]
    assert snr[0] >= 10 and snr[1] >= 10 and snr[2] >= 8
    _agree(impl, x, 1_000_000)


def test_frequency_sweep_clarity(impl):                                  # :163-199
    sr, dur = 1_000_000, 0.0008
    t = np.linspace(0, dur, int(sr * dur), endpoint=False)
    x = np.hanning(len(t)) * np.exp(2j * np.pi * (70_000 * t + (180_000 - 70_000) * t ** 2 / (2 * dur)))
    _, _, S, db, _, _ = _db(impl, x, sr)
    cols = range(0, S.shape[1], max(1, S.shape[1] // min(10, S.shape[1])))
    gains = [db[:, i].max() - np.mean(db[:, i]) for i in cols]
    assert np.mean(gains) >= 12
    _agree(impl, x, sr)


def test_noise_suppression(impl):                                        # :201-242
    _, x = _burst(1_000_000, 0.0005, [(1.0, 110_000)], noise=0.1, seed=42)
    f, _, _, db, vmin, vmax = _db(impl, x, 1_000_000)
    assert vmax - vmin <= 35
    k = int(np.argmin(np.abs(f - 110_000)))
    rest = np.concatenate([np.arange(0, max(0, k - 20)), np.arange(min(len(f), k + 20), len(f))])
    assert np.mean(db[k, :]) - np.mean(db[rest, :]) >= 12
    _agree(impl, x, 1_000_000)


# ---------------------------------------------------------------- tests/test_utils.py:43-83,163-170
def _sample_packet(duration, sr, frequency):                             # utils.generate_sample_packet, utils.py:679-686
    t = np.linspace(0, duration, int(sr * duration), endpoint=False)
    return np.exp(2j * np.pi * frequency * t)


def test_apply_frequency_shift_moves_the_peak(impl):                     # tests/test_utils.py:43-60
    sr, base, shift = 8000, 1000, 500
    sig = _sample_packet(0.1, sr, base)

    def peak(x):
        sp = np.fft.fftshift(np.fft.fft(x))
        return np.fft.fftshift(np.fft.fftfreq(len(x), d=1 / sr))[np.argmax(np.abs(sp))]
    for s in (shift, -shift):
        out = impl.apply_frequency_shift(sig, s, sr)
        assert out.dtype == np.complex64 and abs(peak(out) - (base + s)) < 1
        if impl is not _Oracle:
            assert P.rel_l2(out, O.apply_frequency_shift(sig, s, sr)) <= 2e-7


def test_spectrogram_reflects_shift(impl):                               # tests/test_utils.py:63-76
    sr = 8000
    shifted = impl.apply_frequency_shift(_sample_packet(0.1, sr, 1000), 500, sr)
    f, _, S = impl.create_spectrogram(shifted, sr)
    assert abs(f[np.unravel_index(np.argmax(np.abs(S)), S.shape)[0]] - 1500) < 2


def test_create_spectrogram_preserves_rate(impl):                        # tests/test_utils.py:79-83
    sr = 10_000_000
    f, _, _ = impl.create_spectrogram(np.zeros(1_500_000, dtype=np.float32), sr)
    assert np.isclose(max(abs(f)), sr / 2, rtol=0.01)


def test_save_vector_wv_header(impl, tmp_path):                          # tests/test_utils.py:163-170
    out = tmp_path / "out.wv"
    impl.save_vector_wv(np.array([1 + 1j, -1 - 1j], dtype=np.complex64), str(out), sample_rate=1e6)
    assert out.exists() and b"TYPE: SMU-WV" in out.read_bytes()[:20]


# ---------------------------------------------------------------- test_unified_gui.py:131-317
def _tone56(freq=5e6, duration=0.001):
    fs = 56e6
    t = np.linspace(0, duration, int(fs * duration), endpoint=False)
    return np.exp(2j * np.pi * freq * t).astype(np.complex64)


def test_spectrogram_creation_shapes(impl):                              # test_unified_gui.py:134-150
    f, ts, S = impl.create_spectrogram(_tone56(), 56e6)
    assert len(f) > 0 and len(ts) > 0 and S.shape == (len(f), len(ts))


def test_frequency_shift_changes_signal(impl):                           # test_unified_gui.py:152-172
    sig = _tone56()
    out = impl.apply_frequency_shift(sig, 2e6, 56e6)
    assert not np.allclose(sig, out) and out.dtype == np.complex64


def test_packet_loading_round_trip(impl, tmp_path):                      # test_unified_gui.py:174-191
    rng = np.random.default_rng(0)
    pkt = (rng.random(1000) + 1j * rng.random(1000)).astype(np.complex64)
    sio.savemat(str(tmp_path / "p.mat"), {"Y": pkt})
    got = impl.load_packet(str(tmp_path / "p.mat"))
    np.testing.assert_array_almost_equal(pkt, got)
    assert got.dtype == np.complex64


def test_complete_packet_to_vector_workflow(impl, tmp_path):             # test_unified_gui.py:195-317 (TestFullWorkflow)
    """Three packets cut from a three-tone signal, shifts +2/-3/+1 MHz, the reference test's "period_ms" / 1000
    unit slip kept (so only the first packet's instances fit the 0.2 ms vector — it passes on the reference the
    same way): length, dtype, |v| <= 1, energy ratio, MAT round trip; and the build equals the inline loop."""
    fs = 56e6
    rng = np.random.default_rng(5)
    t = np.linspace(0, 0.0001, int(fs * 0.0001), endpoint=False)
    sig = (np.exp(2j * np.pi * 5e6 * t) + 0.7 * np.exp(2j * np.pi * 10e6 * t) + 0.5 * np.exp(2j * np.pi * 15e6 * t))
    sig = (sig + 0.1 * (rng.random(len(t)) + 1j * rng.random(len(t)))).astype(np.complex64)
    packets = [(sig[i * 1000: i * 1000 + n], 0) for i, n in enumerate((500, 750, 1000))]
    total = int(0.2 / 1000.0 * fs)
    spec = [(2e6, 50, 0), (-3e6, 75, 10), (1e6, 60, 25)]                # (shift, "period_ms", "start_time_ms")
    configs = [dict(file=f"test_packet_{i+1}.mat", freq_shift=s, period=p / 1000.0, start_time=st / 1000.0)
               for i, (s, p, st) in enumerate(spec)]
    # the inline loop of the reference test, with ITS helper for the shift
    want = np.zeros(total, dtype=np.complex64)
    inst = 0
    for (y, _), (s, p, st) in zip(packets, spec):
        y = O.apply_frequency_shift(y, s, fs)
        per, pos = int(p / 1000.0 * fs), int(st / 1000.0 * fs)
        while pos + len(y) <= total:
            want[pos:pos + len(y)] += y
            inst += 1
            pos += per
    assert inst > 0
    want = want / np.max(np.abs(want))
    vec, _, counts = impl.build_vector(configs, total, True, packets)
    assert sum(counts) == inst
    assert len(vec) == total and vec.dtype == np.complex64 and np.max(np.abs(vec)) <= 1.0 + 1e-6
    assert np.sum(np.abs(vec) ** 2) / len(vec) > 0.01
    assert P.rel_l2(vec, want) <= P.VEC_REL_L2
    path = str(tmp_path / "test_vector_complete.mat")
    impl.save_vector(vec, path)
    np.testing.assert_array_almost_equal(vec, sio.loadmat(path, squeeze_me=True)["Y"])


# ---------------------------------------------------------------- test_complete_workflow.py:91-285
def test_vector_creation_2_5ms(impl, tmp_path):                          # test_complete_workflow.py:91-172
    """2.5 ms from the tone and chirp fixtures at 0.2 ms / 1.0 ms and 0.6 ms / 1.2 ms.  The reference script
    places by `int(current_time * fs)` with no pre-buffer subtraction; pre_samples = 0 packets make
    generate_vector's arithmetic the same placement."""
    fs = 56e6
    total = int(2.5 / 1000.0 * fs)
    bank = O.synth_bank(n=2)
    packets = [(bank[0][0], 0), (bank[1][0], 0)]
    want = np.zeros(total, dtype=np.complex64)
    n_inst = []
    for (y, _), (start, period) in zip(packets, ((0.2e-3, 1.0e-3), (0.6e-3, 1.2e-3))):
        tcur, k = start, 0
        while tcur + len(y) / fs <= 2.5e-3:
            s0 = int(tcur * fs)
            if s0 + len(y) <= total:
                want[s0:s0 + len(y)] += y
                k += 1
            tcur += period
        n_inst.append(k)
    configs = [dict(file="packet_1_5MHz.mat", freq_shift=0.0, period=1.0e-3, start_time=0.2e-3),
               dict(file="packet_2_chirp.mat", freq_shift=0.0, period=1.2e-3, start_time=0.6e-3)]
    vec, markers, counts = impl.build_vector(configs, total, False, packets)
    assert counts == n_inst == [2, 1] and len(markers) == 3
    assert np.array_equal(vec, want)                                    # float32 adds only: bit-exact
    path = str(tmp_path / "vector_2_5ms.mat")
    impl.save_vector(vec, path)
    y, pre = impl.load_packet_info(path)
    assert np.array_equal(y, want) and pre == 0 and len(y) / fs * 1000 == 2.5


def test_microsecond_timing_precision(impl):                             # test_complete_workflow.py:174-241
    """10 us packets of ones every 100 us over 5 ms: the spacing is exact in samples (5600)."""
    fs = 56e6
    total = int(0.005 * fs)
    pkt = (np.ones(int(10e-6 * fs), dtype=np.complex64), 0)
    cfg = [dict(file="ones.mat", freq_shift=0.0, period=100e-6, start_time=0.0)]
    vec, markers, counts = impl.build_vector(cfg, total, False, [pkt])
    times_us = np.array([m[0] for m in markers]) * 1e6
    assert counts[0] == len(times_us) >= 49
    assert abs(np.mean(np.diff(times_us)) - 100) < 1.0
    starts = np.flatnonzero(np.diff(np.concatenate([[0], np.abs(vec) > 0]).astype(np.int8)) == 1)
    assert np.all(np.diff(starts) == 5600) and starts[0] == 0


def test_spectrogram_of_chirp_packet(impl):                              # test_complete_workflow.py:243-285
    pkt = O.synth_bank(n=2)[1][0]
    f, t, S = impl.create_spectrogram(pkt, 56e6, time_resolution_us=1)
    assert len(f) > 1 and len(t) > 1 and S.shape == (len(f), len(t))
    assert (t[1] - t[0]) * 1e6 <= 1.01 and 50 < (f[-1] - f[0]) / 1e6 < 56.1
    if impl is not _Oracle:
        _, _, Sr = O.create_spectrogram(pkt, 56e6, time_resolution_us=1)
        assert P.spec_db_error(S, Sr) <= P.SPEC_DB

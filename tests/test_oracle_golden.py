"""Pin the CPU oracle (oracle/ref_port.py) to outputs of the REAL reference.

tests/golden/* were written by tests/golden/make_golden.py running ramiyako/vector's own
generate_vector / utils functions headless.  Here the port recomputes the same cases from
the same seeded packets and must agree: integers/markers/plans exactly, sample data
bit-for-bit (SHA-1) when the numpy/scipy versions match the generator's, and in any case
on the stored sample windows to float32 round-off.
"""
import numpy as np
import pytest
import scipy

from oracle import ref_port as O
from tests.golden import cases


def _same_env(meta):
    return meta["numpy"] == np.__version__ and meta["scipy"] == scipy.__version__


def _check_vec(name, vec, meta, arrays, exact):
    g = meta["builds"][name]
    assert vec.dtype == np.complex64 and vec.shape[0] == g["n"]
    starts = arrays[f"{name}__starts"]
    wins = arrays[f"{name}__windows"]
    for s, w in zip(starts, wins):
        got = vec[s:s + w.shape[0]]
        if exact:
            assert np.array_equal(got, w), f"{name}: window @{s} differs"
        else:
            np.testing.assert_allclose(got, w, rtol=2e-6, atol=2e-7)
    assert int(np.count_nonzero(vec)) == g["nonzero"]
    if exact:
        assert cases.sha1(vec) == g["sha1"]
        if f"{name}__full" in arrays.files:
            assert np.array_equal(vec, arrays[f"{name}__full"])
    assert abs(float(np.abs(vec).sum(dtype=np.float64)) - g["sum_abs"]) <= 1e-6 * max(1.0, g["sum_abs"])


def test_bank_matches_generator(golden_meta):
    bank = dict(zip(O.PACKET_KINDS, O.synth_bank(n=5)))
    for k, (y, pre) in bank.items():
        assert pre == 100 and y.dtype == np.complex64 and len(y) == 56100
        if _same_env(golden_meta):
            assert cases.sha1(y) == golden_meta["bank_sha1"][k]


@pytest.mark.parametrize("name", ["wf_2p5ms_plain", "wf_2p5ms_shift_norm", "overlap_5ms",
                                  "empty_0p5ms", "cfg1_100ms"])
def test_build_matches_reference(name, golden_meta, golden_arrays):
    case = cases.build_cases()[name]
    pk = O.packets_for(case["configs"])
    n = O.total_samples_unified(case["length_ms"])
    vec, markers, counts, max_abs = O.build_vector(pk, case["configs"], n, case["normalize"])
    g = golden_meta["builds"][name]
    assert [m[0] for m in markers] == g["marker_times"]
    assert [m[2] for m in markers] == g["marker_names"]
    _check_vec(name, vec, golden_meta, golden_arrays, _same_env(golden_meta))
    title, details = O.validate_packet_timing(markers, case["configs"])
    assert title == g["title"]
    if g["details"]:
        for d, e in zip(details, g["details"]):
            assert d["packet_name"] == e["packet_name"] and d["instances"] == e["instances"]
            assert d["explanations"] == e["explanations"]
            assert d["start_time_error_ms"] == pytest.approx(e["start_time_error_ms"], abs=1e-12)


@pytest.mark.slow
def test_cfg2_matches_reference(golden_meta, golden_arrays):
    name = "cfg2_1000ms"
    case = cases.build_cases()[name]
    pk = O.packets_for(case["configs"])
    n = O.total_samples_unified(case["length_ms"])
    assert n == 56_000_000
    vec, markers, counts, max_abs = O.build_vector(pk, case["configs"], n, True)
    assert counts == [10, 20, 49, 97, 192, 475]
    assert [m[0] for m in markers] == golden_meta["builds"][name]["marker_times"]
    _check_vec(name, vec, golden_meta, golden_arrays, _same_env(golden_meta))
    assert O.validate_packet_timing(markers, case["configs"])[0] == golden_meta["builds"][name]["title"]


def test_known_answers_cfg1(golden_meta):
    """instance counts 99/83/66 and first start index 11 100."""
    cfgs = O.cfg1_configs()
    n = O.total_samples_unified("100")
    assert n == 5_600_000
    got = [O.placement(c, 100, 56100, n) for c in cfgs]
    assert [g[2] for g in got] == [99, 83, 66]
    assert got[0][0] == 11_100


def test_spectrogram_of_build_matches_reference(golden_meta, golden_arrays):
    for name in ("wf_2p5ms_shift_norm", "cfg1_100ms"):
        case = cases.build_cases()[name]
        pk = O.packets_for(case["configs"])
        n = O.total_samples_unified(case["length_ms"])
        vec = O.build_vector(pk, case["configs"], n, case["normalize"])[0]
        f, t, S = O.create_spectrogram(vec, 56e6, time_resolution_us=1)
        g = golden_meta["builds"][name]["spectrogram"]
        assert list(S.shape) == g["shape"] and str(S.dtype) == g["dtype"]
        assert (float(f[0]), float(f[-1]), float(t[0]), float(t[-1])) == (g["f0"], g["f1"], g["t0"], g["t1"])
        sub = S[::16, ::max(1, S.shape[1] // 64)]
        ref = golden_arrays[f"{name}__spec_sub"]
        db, vmin, vmax = O.normalize_spectrogram(S, 5, 95, 25)
        if _same_env(golden_meta):
            assert cases.sha1(S) == g["sha1"]
            assert cases.sha1(f) == g["freqs_sha1"] and cases.sha1(t) == g["times_sha1"]
            assert cases.sha1(db) == g["db_sha1"]
            assert (float(vmin), float(vmax)) == (g["vmin"], g["vmax"])
        else:
            np.testing.assert_allclose(sub, ref, rtol=1e-4, atol=1e-12 * g["peak"])
            assert abs(float(vmin) - g["vmin"]) < 1e-3 and abs(float(vmax) - g["vmax"]) < 1e-3


def test_plans_match_reference(golden_meta):
    for key, g in golden_meta["plans"].items():
        if "@" in key:
            continue
        n, tr = (int(v) for v in key.split(":"))
        p = O.spectrogram_plan(n, 56e6, time_resolution_us=tr)
        got = dict(n_dec=p.n_dec, fs=p.fs, window=p.window, nperseg=p.nperseg,
                   noverlap=p.noverlap, nfft=p.nfft, scaling=p.scaling)
        assert got == g, (key, got, g)
    g = golden_meta["plans"]["2000000000:1@200e6"]
    p = O.spectrogram_plan(20_000_000, 200e6, time_resolution_us=1)
    assert (p.n_dec, p.fs, p.window, p.nperseg, p.noverlap, p.nfft) == (
        g["n_dec"], g["fs"], g["window"], g["nperseg"], g["noverlap"], g["nfft"])


def test_shift_and_resample_match_reference(golden_meta, golden_arrays):
    y, _ = O.synth_bank(n=1)[0]
    sh = O.apply_frequency_shift(y, 2e6, 56e6)
    exact = _same_env(golden_meta)
    if exact:
        assert np.array_equal(sh, golden_arrays["shift_tone_p2MHz"])
        assert cases.sha1(O.apply_frequency_shift(y, -12.3456e6, 56e6)) == \
            golden_meta["misc"]["shift_tone_m12p3456MHz_sha1"]
    else:
        np.testing.assert_allclose(sh, golden_arrays["shift_tone_p2MHz"], rtol=1e-6, atol=1e-7)
    assert O.apply_frequency_shift(y, 0, 56e6) is y            # utils.py:122-123
    assert O.resample_signal(y, 56e6, 56e6) is y               # utils.py:109-110
    for (n_in, sr_in, sr_out) in ((2020, 20e6, 56e6), (4020, 40e6, 56e6), (56100, 56e6, 200e6),
                                  (5000, 56e6, 20e6), (4096, 56e6, 28e6)):
        key = f"resample_{n_in}_{int(sr_in/1e6)}to{int(sr_out/1e6)}"
        x = O.synth_packet("chirp", fs=sr_in, body_s=(n_in - 20) / sr_in, pre=20,
                           seed=300 + n_in).astype(np.complex64)[:n_in]
        r = O.resample_signal(x, sr_in, sr_out)
        g = golden_meta["misc"][key]
        assert len(r) == g["n_out"] and r.dtype == np.complex64
        if exact:
            assert cases.sha1(x) == g["in_sha1"] and cases.sha1(r) == g["sha1"]
        np.testing.assert_allclose(r[:1024], golden_arrays[key + "__head"], rtol=1e-5, atol=1e-6)
        # the fp64 restatement agrees with scipy's single-precision path to fp32 round-off
        r64 = O.resample_fourier_f64(x, len(r))
        err = np.linalg.norm(r64 - r) / np.linalg.norm(r64)
        assert err < 2e-6, (key, err)


def test_main_flavour_with_resample_matches_reference(golden_meta, golden_arrays):
    g = golden_meta["builds"]["main_resample"]
    pk, cfgs = [], []
    for i, c in enumerate(g["configs"]):
        kind = c["file"].split("_")[1].split(".")[0]
        y = O.synth_packet(kind, fs=c["sample_rate"], body_s=1e-4, pre=20, seed=201 + i)
        pk.append((y.astype(np.complex64), 20))
        cfgs.append(c)
    n = O.total_samples_main("0.004")
    vec = O.build_vector(pk, cfgs, n, True, resample=True)[0]
    _check_vec("main_resample", vec, golden_meta, golden_arrays, _same_env(golden_meta))


def test_wv_bytes_match_reference(golden_meta):
    case = cases.build_cases()["wf_2p5ms_shift_norm"]
    pk = O.packets_for(case["configs"])
    n = O.total_samples_unified(case["length_ms"])
    vec = O.build_vector(pk, case["configs"], n, True)[0]
    hdr_ref = golden_meta["misc"]["wv_header_wf_2p5ms_shift_norm"]
    date = hdr_ref.split("{DATE: ")[1].split("}")[0]
    assert O.wv_header(n, 56e6, date=date).decode() == hdr_ref
    assert hdr_ref.startswith("{TYPE: SMU-WV,0}")                  # tests/test_utils.py:163-170
    if _same_env(golden_meta):
        payload = O.wv_quantize(vec)
        assert cases.sha1(payload.view(np.uint8)) == golden_meta["misc"]["wv_payload_sha1"]


def test_zero_signal_fallback_shape(golden_meta):
    f, t, S = O.create_spectrogram(np.zeros(20000, np.complex64), 56e6)
    assert list(S.shape) == golden_meta["misc"]["spec_zeros_shape"]
    assert float(t[0]) == golden_meta["misc"]["spec_zeros_t0"]
    with pytest.raises(ValueError, match="Signal is empty"):
        O.create_spectrogram(np.zeros(0, np.complex64), 56e6)


def test_normalisation_is_reciprocal_multiply():
    """numpy divides complex64 by a real scalar as v * (1/m) (Smith's algorithm with a zero
    imaginary divisor), NOT as a per-component IEEE division — the device scale pass must do
    the same to be bit-exact (unified_gui.py:1781)."""
    rng = np.random.default_rng(3)
    v = (rng.standard_normal(100_000) + 1j * rng.standard_normal(100_000)).astype(np.complex64)
    m = np.max(np.abs(v))
    assert m.dtype == np.float32
    assert np.array_equal(v / m, v * (np.float32(1) / m))

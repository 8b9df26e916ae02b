"""Host logic of the product (vector_b200/hostmath.py) against the oracle and the golden
fixtures of the real reference.  CPU only."""
import numpy as np
import pytest
import scipy.signal

from oracle import ref_port as O
from tests.golden import cases
from vector_b200 import hostmath as H


def test_placement_integers_match_oracle():
    rng = np.random.default_rng(0)
    for _ in range(300):
        cfg = dict(period=round(rng.uniform(0.01, 10), 3) / 1000.0, start_time=round(rng.uniform(0, 5), 4) / 1000.0,
                   freq_shift=0.0)
        pre = int(rng.integers(0, 300))
        L = int(rng.integers(1, 80000))
        total = int(rng.integers(0, 3_000_000))
        start, period, count = O.placement(cfg, pre, L, total)
        assert H.period_samples(cfg["period"]) == period
        assert H.start_offset(cfg["start_time"], pre) == start
        assert H.instance_count(start, period, L, total) == count
    assert H.total_samples(float("1000") / 1000.0) == 56_000_000 == O.total_samples_unified("1000")
    assert H.total_samples(float("2.5") / 1000.0) == 140_000
    with pytest.raises(ValueError):
        H.instance_count(0, 0, 10, 100)


def test_cfg_known_answers():
    cfgs = O.cfg2_configs()
    n = 56_000_000
    got = [H.instance_count(H.start_offset(c["start_time"], 100), H.period_samples(c["period"]), 56100, n)
           for c in cfgs]
    assert got == [10, 20, 49, 97, 192, 475]


def test_plans_match_reference_golden(golden_meta):
    for key, g in golden_meta["plans"].items():
        if "@" in key:
            n, sr, tr = 20_000_000, 200e6, 1
        else:
            n, tr = (int(v) for v in key.split(":"))
            sr = 56e6
        p = H.plan_spectrogram(n, sr, time_resolution_us=tr)
        assert (p.n_dec, p.fs, p.window, p.nperseg, p.noverlap, p.nfft, p.scaling) == (
            g["n_dec"], g["fs"], g["window"], g["nperseg"], g["noverlap"], g["nfft"], g["scaling"]), key
        o = O.spectrogram_plan(n, sr, time_resolution_us=tr)
        assert (p.factor, p.heavy, p.n_frames) == (o.factor, o.heavy, o.n_frames)


def test_plan_sweep_matches_oracle():
    rng = np.random.default_rng(5)
    for n in list(rng.integers(1, 6_000_000, 200)) + [1, 2, 9, 10, 11, 5_000_000, 5_000_001]:
        for tr in (1, 7, 50, None):
            for ad in (True, False):
                for ms in (2_000_000, 1_000_000, 1000):
                    try:
                        b = O.spectrogram_plan(int(n), 56e6, ms, tr, ad)
                    except TypeError:            # the reference itself raises for None + adaptive/heavy
                        with pytest.raises(TypeError):
                            H.plan_spectrogram(int(n), 56e6, ms, tr, ad)
                        continue
                    a = H.plan_spectrogram(int(n), 56e6, ms, tr, ad)
                    assert (a.factor, a.n_dec, a.fs, a.window, a.nperseg, a.noverlap, a.nfft) == (
                        b.factor, b.n_dec, b.fs, b.window, b.nperseg, b.noverlap, b.nfft)
    with pytest.raises(ValueError, match="Signal is empty"):
        H.plan_spectrogram(0, 56e6)


def test_times_freqs_scale_match_scipy():
    rng = np.random.default_rng(2)
    for n, win, nps, nov, nfft, sc in [(140000, "blackmanharris", 512, 456, 1024, "spectrum"),
                                        (933334, "hann", 934, 0, 1024, "spectrum"),
                                        (30000, "hann", 1024, 512, 1024, "density"),
                                        (700, "boxcar", 32, 16, 64, "spectrum"),
                                        (20000, "hann", 64, 16, 128, "spectrum")]:
        x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
        fs = 56e6 / 3
        f, t, S = scipy.signal.spectrogram(x, fs=fs, window=win, nperseg=nps, noverlap=nov, nfft=nfft,
                                           return_onesided=False, detrend=False, scaling=sc)
        p = H.SpectrogramPlan(n, 1, n, fs, False, win, nps, nov, nfft, sc)
        assert np.array_equal(p.times(), t) and np.array_equal(p.freqs(), f) and p.n_frames == S.shape[1]
        taps, scale = H.window_and_scale(win, nps, fs, sc)
        assert taps.dtype == np.float32 and np.array_equal(taps, scipy.signal.get_window(win, nps).astype(np.float32))
        # reproduce one frame with numpy: |fft(frame*win)|^2*scale == scipy's column
        X = np.fft.fft((x[:nps] * taps).astype(np.complex64), n=nfft)
        col = (X.real.astype(np.float32) ** 2 + X.imag.astype(np.float32) ** 2) * scale
        np.testing.assert_allclose(col, S[:, 0], rtol=2e-4, atol=1e-6 * S[:, 0].max())


def test_resolve_plan_mirrors_scipy_argument_checks():
    p = H.SpectrogramPlan(10, 1, 10, 1.0, False, "boxcar", 32, 16, 64)
    with pytest.raises(ValueError, match="noverlap must be less than nperseg"):
        H.resolve_plan(p)                                    # nperseg clamped to 10 <= noverlap 16
    with pytest.raises(ValueError):
        scipy.signal.spectrogram(np.zeros(10, np.complex64), window="boxcar", nperseg=32, noverlap=16, nfft=64,
                                 return_onesided=False, detrend=False)
    p = H.SpectrogramPlan(20, 1, 20, 1.0, False, "boxcar", 32, 16, 64)
    assert H.resolve_plan(p).nperseg == 20
    with pytest.raises(ValueError, match="positive"):
        H.resolve_plan(H.SpectrogramPlan(3, 1, 3, 1.0, False, "hann", 0, 0, 128))


def test_percentile_index_arithmetic_is_numpys():
    rng = np.random.default_rng(1)
    for n in (1, 2, 3, 10, 999, 1000, 10_234_880, 20_000_001):
        a = rng.standard_normal(n).astype(np.float32)
        srt = np.sort(a)
        for q in (5, 5.0, 10.0, 95.0, 0, 100, 50):
            lo, hi, g = H.percentile_ranks(n, q, np.float32)
            got = H.lerp(srt[lo], srt[hi], g)
            ref = np.percentile(a, q)
            assert got == ref and got.dtype == ref.dtype, (n, q)


def test_colour_limits_match_oracle():
    rng = np.random.default_rng(3)
    for spread in (0.01, 1.0, 30.0, 300.0):
        S = (rng.standard_normal((64, 50)) ** 2 * 10 ** (rng.uniform(-spread / 10, 0, (64, 50)))).astype(np.float32)
        db, vmin, vmax = O.normalize_spectrogram(S, 5, 95, 25)
        lo = np.percentile(db, 5)
        hi = np.percentile(db, 95)
        assert H.colour_limits(lo, hi, db.min(), db.max(), 25) == (vmin, vmax)
    db = np.full((4, 4), -30.0, np.float32)
    assert H.colour_limits(np.float32(-30), np.float32(-30), db.min(), db.max(), 25) == (-30, -5)   # flat input: vmax = vmin + range (utils.py:379-382)


def test_validation_matches_oracle_and_reference(golden_meta):
    for name, case in cases.build_cases().items():
        if name == "cfg2_1000ms":
            continue
        pk = O.packets_for(case["configs"])
        n = O.total_samples_unified(case["length_ms"])
        markers = []
        for (y, pre), c in zip(pk, case["configs"]):
            start, period, count = O.placement(c, pre, len(y), n)
            base = c["file"].rsplit(".", 1)[0]
            markers += [((start + k * period + pre) / 56e6, c["freq_shift"], base, "x", "C0") for k in range(count)]
        title, details = H.validate_packet_timing(markers, case["configs"])
        g = golden_meta["builds"][name]
        assert title == g["title"] == O.validate_packet_timing(markers, case["configs"])[0]
        for d, e in zip(details, g["details"] or []):
            assert d["explanations"] == e["explanations"] and d["instances"] == e["instances"]
    assert H.validate_packet_timing([], [None]) == ("No packets to validate", [])


def test_compute_freq_ranges():
    assert H.compute_freq_ranges([]) is None and H.compute_freq_ranges([0, 0]) is None
    assert H.compute_freq_ranges([2e6, -3e6, 0]) == [(-4e6, 3e6)] == O.compute_freq_ranges([2e6, -3e6, 0])
    assert H.compute_freq_ranges([True, "a"]) is None


def test_slab_and_frame_partition_cover_exactly_once():
    for total in (1, 15, 16, 17, 5_600_000, 56_000_001, 2_000_000_000):
        for world in (1, 2, 3, 4, 8):
            cuts = [H.slab_bounds(total, world, r) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            for (a, b), (c, d) in zip(cuts, cuts[1:]):
                assert b == c and a <= b
            assert all(lo % 16 == 0 for lo, _ in cuts)
            for stride, nps, hop in ((1, 1024, 512), (56, 512, 100), (2000, 512, 10)):
                n_dec = (total + stride - 1) // stride
                nf = 0 if n_dec < nps else (n_dec - nps) // hop + 1
                fr = [H.frame_range_for_slab(lo, hi, total, stride, nps, hop, nf) for lo, hi in cuts]
                assert fr[0][0] == 0 and fr[-1][1] == nf
                for (a, b, h1), (c, d, h2) in zip(fr, fr[1:]):
                    assert b == c
                for (lo, hi), (a, b, halo) in zip(cuts, fr):
                    if b > a:
                        assert lo <= a * hop * stride < hi or a == b
                        assert ((b - 1) * hop + nps - 1) * stride < halo <= total


def test_product_synth_matches_oracle_synth():
    """The product's scenario generator (vector_b200/synth.py) and the oracle's are separate
    code; they must produce the same packets and tables."""
    from vector_b200 import synth
    for i, k in enumerate(synth.KINDS):
        assert np.array_equal(synth.packet(k, seed=101 + i), O.synth_packet(k, seed=101 + i))
    strip = lambda cs: [{k: v for k, v in c.items()} for c in cs]
    assert strip(synth.configs(synth.CFG1_ROWS)) == strip(O.cfg1_configs())
    assert strip(synth.configs(synth.CFG2_ROWS)) == strip(O.cfg2_configs())
    cfgs, pk = synth.scenario(synth.CFG2_ROWS)
    for (a, pa), (b, pb) in zip(pk, O.packets_for(O.cfg2_configs())):
        assert pa == pb and np.array_equal(a, b)


# ------------------------------------------------------------------------------ max-abs ranges
def _brute_vector(rows, packets, total):
    """The reference's placement loop (unified_gui.py:1757-1769) in float32, packet then instance order."""
    v = np.zeros(total, np.complex64)
    for (start, period, count, _), y in zip(rows, packets):
        for k in range(count):
            pos = start + k * period
            v[pos:pos + len(y)] += y
    return v


@pytest.mark.parametrize("seed", range(12))
def test_max_eval_ranges_give_the_exact_max(seed):
    """max|v| over hostmath.max_eval_ranges == max|v| over the whole vector, bit for bit, and the
    claimed period really repeats; random trains incl. self-overlapping ones (L > P)."""
    rng = np.random.default_rng(seed)
    total = int(rng.integers(20_000, 60_000))
    npk = int(rng.integers(1, 5))
    rows, packets = [], []
    for p in range(npk):
        period = int(rng.choice([60, 100, 150, 200, 300, 400, 600]))
        length = int(rng.integers(10, 2 * period if seed % 3 == 0 else period))
        start = int(rng.integers(0, 3000))
        y = (rng.standard_normal(length) + 1j * rng.standard_normal(length)).astype(np.complex64)
        count = H.instance_count(start, period, length, total)
        if seed == 5 and p == 0:
            count = 1 if count else 0                         # a packet placed once: must fall back to the whole vector
        rows.append((start, period, count, p)); packets.append(y)
    v = _brute_vector(rows, packets, total)
    ranges = H.max_eval_ranges(rows, [len(y) for y in packets], total)
    assert ranges and ranges[0][0] == 0 and all(lo < hi for lo, hi in ranges)
    assert all(a[1] <= b[0] for a, b in zip(ranges, ranges[1:]))
    sub = np.concatenate([np.abs(v[lo:hi]) for lo, hi in ranges])
    assert sub.max() == np.abs(v).max()
    if seed == 5:
        assert ranges == [(0, total)]
    for world in (2, 3):
        parts = [H.split_ranges(ranges, world, r) for r in range(world)]
        assert sum(hi - lo for part in parts for lo, hi in part) == sum(hi - lo for lo, hi in ranges)
        assert max(np.abs(v[lo:hi]).max() for part in parts for lo, hi in part) == np.abs(v).max()


def test_max_eval_ranges_cfg2_shape():
    """BASELINE cfg2: periods 100/50/20/10/5/2 ms -> one 100 ms period + the 49 ms head + the 100-sample tail."""
    rows = []
    for per, st in [(100, 0), (50, 10), (20, 20), (10, 30), (5, 40), (2, 50)]:
        p = H.period_samples(per / 1000, 56e6); s = H.start_offset(st / 1000, 100, 56e6)
        rows.append((s, p, H.instance_count(s, p, 56100, 56_000_000), 0))
    assert H.max_eval_ranges(rows, [56100] * 6, 56_000_000) == [(0, 8_344_000), (55_999_900, 56_000_000)]


# ------------------------------------------------------------------------------ MAT v5 fast writer
@pytest.mark.parametrize("n", [0, 1, 2, 3, 5, 6, 1000, 56_101])
def test_save_vector_matches_savemat_bytes(tmp_path, n):
    """api.save_vector writes savemat's exact element bytes (utils.py:659-670): everything after the
    128-byte text header (which carries the time of day) is identical, and load_packet_info reads the
    file back like the reference's own files."""
    import io
    import scipy.io as sio
    from vector_b200 import api
    rng = np.random.default_rng(n)
    v = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    path = tmp_path / "v.mat"
    api.save_vector(v, str(path))
    got = path.read_bytes()
    ref = io.BytesIO()
    sio.savemat(ref, {"Y": v, "pre_samples": 0})
    ref = ref.getvalue()
    assert len(got) == len(ref) and got[128:] == ref[128:]
    assert got[116:128] == ref[116:128] and got[:48] == ref[:48]        # version / endian, "MATLAB 5.0 MAT-file Platform: ..."
    back = sio.loadmat(str(path))
    assert back["Y"].dtype == np.complex64 and np.array_equal(back["Y"].ravel(), v) and int(back["pre_samples"].item()) == 0
    if n >= 2:                       # (a 1-sample file squeezes to a Python scalar in the reference's loader too)
        y, pre = api.load_packet_info(str(path))
        assert np.array_equal(y, v) and pre == 0


def test_save_vector_flattens_and_casts(tmp_path):
    from vector_b200 import api
    import scipy.io as sio
    v = (np.arange(12) - 3j * np.arange(12)).reshape(3, 4).astype(np.complex128)
    api.save_vector(v, str(tmp_path / "m.mat"))
    back = sio.loadmat(str(tmp_path / "m.mat"))["Y"]
    assert back.shape == (1, 12) and back.dtype == np.complex64 and np.array_equal(back.ravel(), v.ravel().astype(np.complex64))


# ------------------------------------------------------------------------------ free-instance hint
def _brute_free_flags(rows, lengths, total):
    """Every packet-local sample unsummed somewhere: coverage counted sample by sample."""
    cov = np.zeros(total, np.int32)
    for (s, p, c, _), n in zip(rows, lengths):
        for k in range(c):
            cov[s + k * p:s + k * p + n] += 1
    flags = []
    for (s, p, c, _), n in zip(rows, lengths):
        seen = np.zeros(n, bool)
        for k in range(c):
            seen |= cov[s + k * p:s + k * p + n] == 1
        flags.append(bool(c > 0 and seen.all()))
    return flags


def test_free_instance_flags_match_brute_force():
    rng = np.random.default_rng(42)
    for _ in range(200):
        npk = int(rng.integers(1, 5))
        total = int(rng.integers(200, 3000))
        rows, lengths = [], []
        for i in range(npk):
            n = int(rng.integers(5, 200))
            period = int(rng.integers(max(1, n // 2), 4 * n))       # sometimes shorter than the packet: self-overlap
            start = int(rng.integers(0, 300))
            rows.append((start, period, H.instance_count(start, period, n, total), i))
            lengths.append(n)
        assert H.free_instance_flags(rows, lengths, total) == _brute_free_flags(rows, lengths, total)
    assert H.free_instance_flags([(0, 10, 0, 0)], [5], 100) == [False]
    assert H.free_instance_flags([(0, 10, 3, 0)], [5], 100) == [True]
    assert H.free_instance_flags([(0, 4, 3, 0)], [5], 100) == [True]                      # neighbours overlap by one sample: head free in the first, tail in the last
    assert H.free_instance_flags([(0, 2, 3, 0)], [5], 100) == [False]                     # sample 2 is always under a neighbour
    assert H.free_instance_flags([(0, 10, 3, 0), (2, 10, 3, 1)], [5, 5], 100) == [False, False]
    assert H.free_instance_flags([(0, 10, 3, 0)], [5], 100, max_instances=2) == [False]   # conservative cap

"""GPU parity tests (run on the B200 box): the CUDA path, called through the C-ABI
(ctypes -> libb200vec.so), against the CPU oracle on the same seeded inputs and against the
golden fixtures the real reference produced.

Bars (BASELINE.json north_star / SURVEY.md §8c): placement integers exact; vector rel-L2
<= 1e-5 (we expect ~1e-7 and bit-exactness wherever only float32 adds are involved);
spectrogram within 0.01 dB on bins >= -80 dBc, band-wise triangulation against an fp64
truth below that; order statistics exact.
"""
import os

import numpy as np
import pytest
import scipy.signal

from oracle import ref_port as O
from tests import parity as P
from tests.golden import cases

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from vector_b200.engine import get_engine
    return get_engine(0)


def _build(name, eng, normalize=None):
    from vector_b200 import generate
    case = cases.build_cases()[name]
    pk = O.packets_for(case["configs"])
    n = O.total_samples_unified(case["length_ms"])
    norm = case["normalize"] if normalize is None else normalize
    want, mk_ref, counts_ref, max_ref = O.build_vector(pk, case["configs"], n, norm)
    vec, mk, counts, maxsq = generate.build_vector_device(case["configs"], n, norm, packets=pk, engine=eng)
    return dict(n=n, want=want, got=vec.cpu().numpy(), dev=vec, mk=mk, mk_ref=mk_ref, counts=counts,
                counts_ref=counts_ref, max_ref=max_ref, maxsq=maxsq, case=case, packets=pk)


# ------------------------------------------------------------------------------ K1 shift
@pytest.mark.parametrize("kind,shift", [("tone", 2e6), ("tone", -12.3456e6), ("chirp", -3e6),
                                        ("noise", 27.999e6), ("bpsk", 1.0)])
def test_frequency_shift_matches_oracle(eng, kind, shift):
    import vector_b200 as vb
    y = O.synth_packet(kind, seed=7).astype(np.complex64)
    want = O.apply_frequency_shift(y, shift, 56e6)
    got = vb.apply_frequency_shift(y, shift, 56e6)
    assert got.dtype == np.complex64 and got.shape == want.shape
    assert P.rel_l2(got, want) < 2e-7
    same = np.mean(got.view(np.float32) == want.view(np.float32))
    assert same > 0.999, same                       # fp64 phase/sincos, one rounding: ~all bit-equal
    assert P.ulp_diff_f32(got, want) <= 1.0
    assert vb.apply_frequency_shift(y, 0, 56e6) is y                    # utils.py:122-123


def test_frequency_shift_matches_reference_golden(eng, golden_arrays):
    import vector_b200 as vb
    y, _ = O.synth_bank(n=1)[0]
    got = vb.apply_frequency_shift(y, 2e6, 56e6)
    ref = golden_arrays["shift_tone_p2MHz"]
    assert P.rel_l2(got, ref) < 2e-7 and P.ulp_diff_f32(got, ref) <= 1.0


def test_reference_shift_test_peak_moves(eng):
    """tests/test_utils.py:43-60 of the reference, against the engine."""
    import vector_b200 as vb
    sr, base, shift = 8000, 1000, 500
    t = np.linspace(0, 0.1, int(sr * 0.1), endpoint=False)
    sig = np.exp(2j * np.pi * base * t)

    def peak(x):
        sp = np.fft.fftshift(np.fft.fft(x))
        fr = np.fft.fftshift(np.fft.fftfreq(len(x), d=1 / sr))
        return fr[np.argmax(np.abs(sp))]
    for s in (shift, -shift):
        out = vb.apply_frequency_shift(sig, s, sr)
        assert out.dtype == np.complex64
        assert abs(peak(out) - (base + s)) < 1


# ------------------------------------------------------------------------------ K0 resample
@pytest.mark.parametrize("n_in,sr_in,sr_out", [(2020, 20e6, 56e6), (4020, 40e6, 56e6),
                                               (56100, 56e6, 200e6), (5000, 56e6, 20e6),
                                               (4096, 56e6, 28e6), (101, 3.0, 7.0), (1, 1.0, 3.0)])
def test_resample_matches_oracle(eng, n_in, sr_in, sr_out, golden_arrays):
    import vector_b200 as vb
    if sr_in > 100:
        x = O.synth_packet("chirp", fs=sr_in, body_s=(n_in - 20) / sr_in, pre=20,
                           seed=300 + n_in).astype(np.complex64)[:n_in]
    else:
        rng = np.random.default_rng(n_in)
        x = (rng.standard_normal(n_in) + 1j * rng.standard_normal(n_in)).astype(np.complex64)
    assert len(x) == n_in
    want = O.resample_signal(x, sr_in, sr_out)
    got = vb.resample_signal(x, sr_in, sr_out)
    assert got.dtype == np.complex64 and got.shape == want.shape
    truth = O.resample_fourier_f64(x, len(want))
    e_ref = P.rel_l2(want, truth)
    e_gpu = P.rel_l2(got, truth)
    assert P.rel_l2(got, want) <= P.VEC_REL_L2
    assert e_gpu <= max(2 * e_ref, 2e-7), (e_gpu, e_ref)      # fp64 Bluestein: closer to exact than scipy-c64
    key = f"resample_{n_in}_{int(sr_in/1e6)}to{int(sr_out/1e6)}"
    if key + "__head" in golden_arrays.files:
        assert P.rel_l2(got[:1024], golden_arrays[key + "__head"]) <= P.VEC_REL_L2
    assert vb.resample_signal(x, 56e6, 56e6) is x                       # utils.py:109-110


# ------------------------------------------------------------------------------ K2/K3 build
@pytest.mark.parametrize("name", ["wf_2p5ms_plain", "wf_2p5ms_shift_norm", "overlap_5ms",
                                  "empty_0p5ms", "cfg1_100ms"])
def test_build_matches_oracle_and_reference(eng, name, golden_meta, golden_arrays):
    r = _build(name, eng)
    g = golden_meta["builds"][name]
    assert r["counts"] == r["counts_ref"]
    assert [m[0] for m in r["mk"]] == g["marker_times"] == [m[0] for m in r["mk_ref"]]
    assert [m[2] for m in r["mk"]] == g["marker_names"]
    got, want = r["got"], r["want"]
    assert got.dtype == np.complex64 and got.shape == want.shape == (g["n"],)
    assert int(np.count_nonzero(got)) == g["nonzero"]
    if not np.any(want):
        assert not np.any(got)
        return
    assert P.rel_l2(got, want) <= 3e-7
    assert P.max_err_rel_peak(got, want) <= 4e-7           # a few float32 ulps of the peak
    for s, w in zip(golden_arrays[f"{name}__starts"], golden_arrays[f"{name}__windows"]):
        assert P.rel_l2(got[s:s + w.shape[0]], w) <= P.VEC_REL_L2 or not np.any(w)
    if r["maxsq"] is not None:
        # numpy's complex64 abs is a SIMD float32 hypot that is not correctly rounded (its last bit
        # depends on the host CPU's dispatch); ours is fl32(sqrt(exact fp64 sum)): allow 1 ulp
        m = np.float32(np.sqrt(r["maxsq"].item()))
        assert abs(m - r["max_ref"]) <= np.spacing(r["max_ref"])


def test_unshifted_build_is_bit_exact(eng):
    """No shift, no normalisation: only float32 adds in the reference's own order."""
    r = _build("wf_2p5ms_plain", eng)
    assert np.array_equal(r["got"], r["want"])


def test_overlapping_instances_bit_exact_without_shift(eng):
    from vector_b200 import generate
    cfgs = O._rows_to_configs([("noise", 0, 0.01, 0.003), ("multitone", 0, 0.4, 0.0), ("bpsk", 0, 0.13, 0.05)])
    pk = O.packets_for(cfgs)
    n = 300_001                                           # odd length, ragged last tile
    want = O.build_vector(pk, cfgs, n, False)[0]
    vec, _, counts, _ = generate.build_vector_device(cfgs, n, False, packets=pk, engine=eng)
    assert counts[0] > 400                                # period 560 samples << packet length
    assert np.array_equal(vec.cpu().numpy(), want)


def test_normalisation_matches_numpy_division(eng):
    r = _build("overlap_5ms", eng, normalize=True)
    raw = _build("overlap_5ms", eng, normalize=False)
    m = np.float32(np.sqrt(r["maxsq"].item()))
    m_np = np.max(np.abs(raw["got"]))                     # numpy's SIMD hypot: may be 1 ulp off the
    assert abs(m - m_np) <= np.spacing(m_np)              # correctly rounded value we compute
    v = raw["got"].astype(np.complex128)
    exact = np.sqrt(np.max(v.real ** 2 + v.imag ** 2))
    assert m == np.float32(exact)                         # ours IS the correctly rounded max
    assert np.array_equal(r["got"], raw["got"] / m)       # complex64 / float32 scalar, bit for bit


@pytest.mark.parametrize("lo,hi", [(0, 4096), (16, 100_000), (4095, 4097), (131_071, 262_145),
                                   (5_599_999, 5_600_000), (1_000_000, 1_000_000)])
def test_slab_equals_slice_of_full_build(eng, lo, hi):
    from vector_b200 import generate
    case = cases.build_cases()["cfg1_100ms"]
    pk = O.packets_for(case["configs"])
    n = 5_600_000
    full, _, _, _ = generate.build_vector_device(case["configs"], n, False, packets=pk, engine=eng)
    part, _, _, _ = generate.build_vector_device(case["configs"], n, False, packets=pk, engine=eng, lo=lo, hi=hi)
    assert part.numel() == hi - lo
    assert torch.equal(part, full[lo:hi])


@pytest.mark.slow
def test_cfg2_full_size(eng, golden_meta, golden_arrays):
    name = "cfg2_1000ms"
    r = _build(name, eng)
    g = golden_meta["builds"][name]
    assert r["n"] == 56_000_000 and r["counts"] == [10, 20, 49, 97, 192, 475]
    assert [m[0] for m in r["mk"]] == g["marker_times"]
    got, want = r["got"], r["want"]
    assert int(np.count_nonzero(got)) == g["nonzero"] == 32_295_500
    assert P.rel_l2(got, want) <= 3e-7
    assert abs(float(np.abs(got).sum(dtype=np.float64)) - g["sum_abs"]) <= 2e-6 * g["sum_abs"]
    for s, w in zip(golden_arrays[f"{name}__starts"], golden_arrays[f"{name}__windows"]):
        if np.any(w):
            assert P.rel_l2(got[s:s + w.shape[0]], w) <= P.VEC_REL_L2


def test_full_size_properties(eng):
    """Size-independent properties at BASELINE size (N = 56 M): linearity over packets and
    periodicity of a single packet's placement; |v| <= 1 with equality after normalisation."""
    from vector_b200 import generate
    cfgs = O.cfg2_configs()
    pk = O.packets_for(cfgs)
    n = 56_000_000
    full, _, _, _ = generate.build_vector_device(cfgs, n, False, packets=pk, engine=eng)
    acc = torch.zeros_like(full)
    for c, p in zip(cfgs, pk):
        one, _, _, _ = generate.build_vector_device([c], n, False, packets=[p], engine=eng)
        acc += one
    assert torch.equal(acc, full)                          # same float32 add order
    c, p = cfgs[-1], pk[-1]
    one, _, counts, _ = generate.build_vector_device([c], n, False, packets=[p], engine=eng)
    per = int(c["period"] * 56e6)
    st = max(0, int(round(c["start_time"] * 56e6)) - 100)
    k = counts[0]
    a = one[st: st + (k - 1) * per]
    b = one[st + per: st + k * per]
    assert torch.equal(a, b)
    nv, _, _, maxsq = generate.build_vector_device(cfgs, n, True, packets=pk, engine=eng)
    mx = torch.abs(nv).max().item()
    assert 1 - 2e-7 <= mx <= 1 + 2e-7
    del full, acc, one, nv


# ------------------------------------------------------------------------------ K4 STFT
def _check_spec(S, S_ref, S_true):
    assert S.shape == S_ref.shape and S.dtype == np.float32
    assert P.spec_db_error(S, S_ref) <= P.SPEC_DB
    ours, refs = P.band_errors(S, S_true), P.band_errors(S_ref, S_true)
    for band, (emax, e99) in ours.items():
        if band[0] <= -80 and band in refs:
            assert e99 <= 2 * refs[band][1] + 1e-4, (band, e99, refs[band])
    zero_cols = ~np.any(S_ref, axis=0)
    assert not np.any(S[:, zero_cols])                     # exact-zero frames stay exact zeros


def test_full_rate_stft_matches_oracle(eng):
    import vector_b200 as vb
    r = _build("wf_2p5ms_shift_norm", eng)
    f, t, S = vb.full_rate_spectrogram_device(r["dev"], 56e6, engine=eng)
    fr, tr, Sr = O.full_rate_spectrogram(r["want"], 56e6)
    _, _, St = O.stft_psd_f64(r["want"], 56e6, "hann", 1024, 512, 1024, "density")
    assert np.array_equal(f, fr) and np.array_equal(t, tr)
    _check_spec(S.cpu().numpy(), Sr, St)


@pytest.mark.parametrize("nperseg,noverlap,stride,win", [(1024, 512, 1, "hann"), (1024, 768, 1, "hann"),
                                                         (1024, 0, 1, "blackmanharris"), (1024, 513, 1, "hann"),
                                                         (1024, 1023, 1, "boxcar"), (1024, 990, 1, "hann"),
                                                         (512, 412, 56, "hann"), (934, 0, 6, "hann"),
                                                         (1000, 499, 3, "blackmanharris")])
def test_stft1024_kernels_match_scipy(eng, nperseg, noverlap, stride, win):
    """The nfft-1024 kernels (direct-store tiles for nperseg 1024 / even hop <= 512 / aligned input, the gather
    kernel for everything else) over hops, strides and windows, odd origins."""
    rng = np.random.default_rng(nperseg + noverlap + stride)
    n = 40_000 * stride + 7
    t = np.arange(n) / 56e6
    x = (np.exp(2j * np.pi * 3.3e6 * t / stride) + 1e-3 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))).astype(np.complex64)
    d = torch.from_numpy(x).cuda()
    fs = 56e6 / stride
    S = eng.stft(d, win, nperseg, noverlap, 1024, fs, scaling="spectrum", stride=stride).cpu().numpy()
    xs = x[::stride]
    _, _, Sr = scipy.signal.spectrogram(xs, fs=fs, window=win, nperseg=nperseg, noverlap=noverlap, nfft=1024,
                                        return_onesided=False, detrend=False, scaling="spectrum")
    Sr = np.fft.fftshift(Sr, axes=0)
    assert S.shape == Sr.shape
    assert P.spec_db_error(S, Sr) <= P.SPEC_DB
    # a slab-relative call (origin != 0, frame_lo != 0) gives the same columns
    hop = nperseg - noverlap
    f_lo = 5
    origin = (f_lo * hop * stride) // 2 * 2
    part = eng.stft(d[origin:], win, nperseg, noverlap, 1024, fs, scaling="spectrum", stride=stride,
                    in_origin=origin, frame_lo=f_lo, n_frames=S.shape[1] - f_lo, total_len=n).cpu().numpy()
    assert np.array_equal(part, S[:, f_lo:])


def test_full_rate_stft_cfg1(eng):
    import vector_b200 as vb
    r = _build("cfg1_100ms", eng)
    f, t, S = vb.full_rate_spectrogram_device(r["dev"], 56e6, engine=eng)
    fr, tr, Sr = O.full_rate_spectrogram(r["want"], 56e6)
    _, _, St = O.stft_psd_f64(r["want"], 56e6, "hann", 1024, 512, 1024, "density")
    assert S.shape == (1024, 10936)
    _check_spec(S.cpu().numpy(), Sr, St)


@pytest.mark.parametrize("n", [1024, 1535, 1536, 5000, 70_000, 300_001])
def test_full_rate_stft_frame_major_equals_rows(eng, n):
    """b200vec_stft1024_fm (warp-per-frame kernel, frame-major memory) == the C-order tile kernel, bit
    for bit, on noise (no exact zeros), whole-vector and slab-relative calls; the oracle bar holds."""
    rng = np.random.default_rng(n)
    x = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    x[n // 3: n // 3 + 2000] = 0                                       # an all-zero stretch: exact-zero frames
    d = torch.from_numpy(x).to(eng.device)
    A = eng.stft(d, "hann", 1024, 512, 1024, 56e6)                     # frame-major default
    B = eng.stft(d, "hann", 1024, 512, 1024, 56e6, frame_major=False)
    assert A.shape == B.shape == (1024, (n - 1024) // 512 + 1)
    assert A.stride(0) == 1 and (B.stride(1) == 1 or B.shape[1] == 1)
    assert torch.equal(A, B)
    _, _, Sr = O.full_rate_spectrogram(x, 56e6)
    assert P.spec_db_error(A.cpu().numpy(), Sr) <= P.SPEC_DB
    if A.shape[1] > 3:                                                 # slab-relative: origin != 0, frame_lo != 0
        f_lo = 2
        part = eng.stft(d[f_lo * 512 - 10:], "hann", 1024, 512, 1024, 56e6, in_origin=f_lo * 512 - 10,
                        frame_lo=f_lo, n_frames=A.shape[1] - f_lo, total_len=n)
        assert torch.equal(part, A[:, f_lo:])
    if n > 1025:                                                       # input 8-byte aligned only: register-load path
        odd = d[1:]
        assert odd.data_ptr() % 16 == 8
        assert torch.equal(eng.stft(odd, "hann", 1024, 512, 1024, 56e6), eng.stft(odd.clone(), "hann", 1024, 512, 1024, 56e6))


# ------------------------------------------------------------------------------ fused build + STFT
def _fused(eng, cfgs, pk, n, normalize, world=1, rank=0, frame_major=True):
    """frame_major=True: warp-per-frame kernel (spectrogram memory [frames, 1024], handed out as its
    transpose view); False: the 16-frame tile kernel writing C-order [1024, frames]."""
    from vector_b200 import dist as vdist, generate
    info = generate.prepare_bank(eng, cfgs, pk, 56e6)
    rows = generate.placement_rows(cfgs, info, n, 56e6)
    plan = vdist.plan_slab(n, world, rank, 1024, 512, 1, align=512)
    out_spec = eng.alloc_spectrogram(1024, plan.frame_hi - plan.frame_lo, frame_major=frame_major)
    vec, spec, _ = vdist.build_stft_slab(eng, rows, plan, normalize, None, fs=56e6, out_spec=out_spec)
    assert spec.shape == (1024, plan.frame_hi - plan.frame_lo)
    return plan, rows, vec, spec


@pytest.mark.parametrize("frame_major", [True, False])
@pytest.mark.parametrize("name", ["wf_2p5ms_shift_norm", "overlap_5ms", "empty_0p5ms", "cfg1_100ms"])
def test_fused_build_stft_matches_oracle(eng, name, frame_major):
    """One kernel for placement + scale + STFT: the vector is bit-identical to the two-pass build,
    the spectrogram meets the oracle bar and is bit-identical to the stand-alone STFT kernel."""
    import vector_b200 as vb
    r = _build(name, eng)
    case = r["case"]
    plan, rows, vec, spec = _fused(eng, case["configs"], r["packets"], r["n"], case["normalize"],
                                   frame_major=frame_major)
    assert torch.equal(vec, r["dev"])
    if r["n"] < 1024:
        assert spec.shape[1] == 0
        return
    fr, tr, Sr = O.full_rate_spectrogram(r["want"], 56e6)
    S = spec.cpu().numpy()
    if not np.any(r["want"]):
        assert S.shape == Sr.shape and not np.any(S)
        return
    _, _, St = O.stft_psd_f64(r["want"], 56e6, "hann", 1024, 512, 1024, "density")
    _check_spec(S, Sr, St)
    _, _, S2 = vb.full_rate_spectrogram_device(r["dev"], 56e6, engine=eng)
    assert torch.equal(spec, S2)


@pytest.mark.parametrize("frame_major", [True, False])
@pytest.mark.parametrize("n", [1, 511, 512, 513, 1023, 1024, 1025, 4096, 4607, 4608, 4609, 8192, 8193, 8704, 8705,
                               16 * 512 * 3 + 511, 300_001])
def test_fused_ragged_lengths(eng, n, frame_major):
    """Tile edges: vectors shorter than a frame, exactly one tile, odd tails, partial last tiles."""
    import vector_b200 as vb
    from vector_b200 import generate
    cfgs = O._rows_to_configs([("noise", 0.5, 0.01, 0.003), ("multitone", -2, 0.4, 0.0), ("bpsk", 1, 0.13, 0.05)])
    for c in cfgs:
        c["period"] = max(c["period"], 1e-5)
    pk = [(y[:700], pre) for y, pre in O.packets_for(cfgs)]          # short packets so that they fit
    want = O.build_vector(pk, cfgs, n, True)[0]
    plan, rows, vec, spec = _fused(eng, cfgs, pk, n, True, frame_major=frame_major)
    ref, _, _, _ = generate.build_vector_device(cfgs, n, True, packets=pk, engine=eng)
    assert vec.numel() == n and torch.equal(vec, ref)
    assert P.rel_l2(vec.cpu().numpy(), want) <= 3e-7
    nf = 0 if n < 1024 else (n - 1024) // 512 + 1
    assert spec.shape == (1024, nf)
    if nf:
        _, _, S2 = vb.full_rate_spectrogram_device(ref, 56e6, engine=eng)
        assert torch.equal(spec, S2)
        _, _, Sr = O.full_rate_spectrogram(want, 56e6)
        assert P.spec_db_error(spec.cpu().numpy(), Sr) <= P.SPEC_DB


@pytest.mark.parametrize("frame_major", [True, False])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_fused_slabs_concatenate(eng, world, frame_major):
    """Time-slab sharding of the fused step, ranks emulated one after another on one GPU: slabs and
    spectrogram columns concatenate to the single-GPU result bit for bit."""
    from vector_b200 import dist as vdist, generate
    from vector_b200._lib import BUILD_MAXONLY, BUILD_SCALE
    case = cases.build_cases()["cfg1_100ms"]
    cfgs, n = case["configs"], 5_600_000 + 300                         # not a multiple of 512
    pk = O.packets_for(cfgs)
    plan1, rows, vec1, spec1 = _fused(eng, cfgs, pk, n, True)
    descs, nd = eng.make_descs(rows)
    plans = [vdist.plan_slab(n, world, r, 1024, 512, 1, align=512) for r in range(world)]
    assert plans[0].lo == 0 and plans[-1].hi == n and plans[-1].frame_hi == plan1.frame_hi
    maxsq = torch.zeros(1, dtype=torch.float64, device=eng.device)
    for pl in plans:                                                  # pass A per rank, MAX-combined
        m = torch.zeros(1, dtype=torch.float64, device=eng.device)
        eng.build_pass(descs, nd, pl.lo, pl.hi, None, BUILD_MAXONLY, maxsq=m)
        maxsq = torch.maximum(maxsq, m)
    vecs, specs = [], []
    for pl in plans:
        v = torch.empty(pl.hi - pl.lo, dtype=torch.complex64, device=eng.device)
        sp = eng.alloc_spectrogram(1024, pl.frame_hi - pl.frame_lo, frame_major=frame_major)
        eng.build_stft1024(descs, nd, pl.lo, pl.hi, v, BUILD_SCALE, pl.frame_lo, pl.frame_hi - pl.frame_lo, sp,
                           fs=56e6, maxsq=maxsq)
        vecs.append(v); specs.append(sp)
    assert torch.equal(torch.cat(vecs), vec1)
    assert torch.equal(torch.cat(specs, dim=1), spec1)


@pytest.mark.slow
def test_fused_full_size(eng):
    """BASELINE size (56 M samples, 109 374 frames): fused == two-pass build + stand-alone STFT."""
    import vector_b200 as vb
    from vector_b200 import generate
    cfgs = O.cfg2_configs()
    pk = O.packets_for(cfgs)
    n = 56_000_000
    plan, rows, vec, spec = _fused(eng, cfgs, pk, n, True)
    _, _, vec_t, spec_t = _fused(eng, cfgs, pk, n, True, frame_major=False)
    assert torch.equal(vec, vec_t) and torch.equal(spec, spec_t)       # warp-per-frame == tile kernel
    del vec_t, spec_t
    ref, _, counts, _ = generate.build_vector_device(cfgs, n, True, packets=pk, engine=eng)
    assert counts == [10, 20, 49, 97, 192, 475]
    assert torch.equal(vec, ref)
    _, _, S2 = vb.full_rate_spectrogram_device(ref, 56e6, engine=eng)
    assert spec.shape == (1024, 109_374) and torch.equal(spec, S2)
    # Parseval per frame: sum_k |X_k|^2 * scale == sum_n |w_n x_n|^2 * 1024 * scale
    w = torch.from_numpy(scipy.signal.get_window("hann", 1024)).to(eng.device)
    scale = 1.0 / (56e6 * float((w * w).sum()))
    for f in (0, 1, 54_321, 109_373):
        seg = ref[f * 512: f * 512 + 1024].to(torch.complex128) * w
        want = float((seg.abs() ** 2).sum()) * 1024 * scale
        got = float(spec[:, f].double().sum())
        assert abs(got - want) <= 1e-5 * max(want, 1e-30)
    del vec, spec, ref, S2


@pytest.mark.parametrize("name", ["wf_2p5ms_shift_norm", "overlap_5ms", "cfg1_100ms"])
def test_create_spectrogram_matches_oracle_and_reference(eng, name, golden_meta, golden_arrays):
    import vector_b200 as vb
    r = _build(name, eng)
    f, t, S = vb.create_spectrogram(r["want"], 56e6, time_resolution_us=1)     # numpy in / numpy out
    fr, tr, Sr = O.create_spectrogram(r["want"], 56e6, time_resolution_us=1)
    assert np.array_equal(f, fr) and np.array_equal(t, tr)
    assert S.flags["C_CONTIGUOUS"]
    p = O.spectrogram_plan(len(r["want"]), 56e6, time_resolution_us=1)
    x = r["want"][::p.factor]
    _, _, St = O.stft_psd_f64(x, p.fs, p.window, p.nperseg, p.noverlap, p.nfft, "spectrum")
    _check_spec(S, Sr, St)
    g = golden_meta["builds"][name].get("spectrogram")
    if g:
        assert list(S.shape) == g["shape"]
        sub = S[::16, ::max(1, S.shape[1] // 64)].astype(np.float64)
        ref = golden_arrays[f"{name}__spec_sub"].astype(np.float64)
        m = ref >= g["peak"] * P.SPEC_FLOOR
        assert np.max(np.abs(10 * np.log10(sub[m]) - 10 * np.log10(ref[m]))) <= P.SPEC_DB


@pytest.mark.parametrize("n,tr", [(56100, 1), (56100, 50), (2000, 1), (500, 1), (33, 1), (280001, 1),
                                  (2_000_001, 1), (5_000_001, 1)])
def test_create_spectrogram_parameter_branches(eng, n, tr):
    """Every duration branch of utils.py:206-235 plus decimation and nfft 512..16384."""
    import vector_b200 as vb
    rng = np.random.default_rng(n)
    tt = np.arange(n) / 56e6
    x = (np.exp(2j * np.pi * 3e6 * tt) + 0.05 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))).astype(np.complex64)
    f, t, S = vb.create_spectrogram(x, 56e6, time_resolution_us=tr)
    fr, t_r, Sr = O.create_spectrogram(x, 56e6, time_resolution_us=tr)
    assert S.shape == Sr.shape and np.array_equal(f, fr) and np.array_equal(t, t_r)
    assert P.spec_db_error(S, Sr) <= P.SPEC_DB


def test_create_spectrogram_fallbacks_and_errors(eng, golden_meta):
    import vector_b200 as vb
    f, t, S = vb.create_spectrogram(np.zeros(20000, np.complex64), 56e6)
    fr, tr, Sr = O.create_spectrogram(np.zeros(20000, np.complex64), 56e6)
    assert list(S.shape) == golden_meta["misc"]["spec_zeros_shape"] == list(Sr.shape)
    assert np.array_equal(t, tr) and np.array_equal(f, fr) and not np.any(S)
    with pytest.raises(ValueError, match="Signal is empty"):
        vb.create_spectrogram(np.zeros(0, np.complex64), 56e6)
    sig64 = np.exp(2j * np.pi * 0.1 * np.arange(4000))               # complex128 in -> float64 out
    assert vb.create_spectrogram(sig64, 8000.0)[2].dtype == np.float64


def test_reference_spectrogram_tests(eng):
    """tests/test_utils.py:69-83 of the reference: the peak follows the shift; decimation
    preserves the +-fs/2 axis."""
    import vector_b200 as vb
    sr = 8000
    t = np.linspace(0, 0.1, 800, endpoint=False)
    sig = np.exp(2j * np.pi * 1000 * t)
    shifted = vb.apply_frequency_shift(sig, 500, sr)
    f, _, S = vb.create_spectrogram(shifted, sr)
    assert abs(f[np.argmax(S.mean(axis=1))] - 1500) < sr / 256
    big = np.exp(2j * np.pi * 1000 * np.arange(3_000_000) / sr).astype(np.complex64)
    f2, _, _ = vb.create_spectrogram(big, sr, max_samples=1_000_000)
    assert abs(f2[0] + sr / 2) < sr / 200 and f2[-1] < sr / 2


# ------------------------------------------------------------------------------ K5 dB/percentiles
def test_select_is_exact(eng):
    rng = np.random.default_rng(11)
    for n in (1, 2, 31, 1000, 4097, 1_000_003):
        a = rng.standard_normal(n).astype(np.float32)
        a[rng.integers(0, n, n // 3)] = 0.0                    # heavy duplicates
        a[rng.integers(0, n, n // 7)] = np.float32(1e-20)
        srt = np.sort(a)
        d = torch.from_numpy(a).cuda()
        ranks = sorted({0, n - 1, n // 2, n // 3, min(n - 1, n // 2 + 1)})
        got = eng.select_ranks(d, ranks)
        assert np.array_equal(got, srt[ranks]), (n, got, srt[ranks])
        pos = np.sort(a[a > 0])
        assert eng.select_count(d, True) == pos.size
        if pos.size:
            rk = sorted({0, pos.size - 1, pos.size // 20})
            assert np.array_equal(eng.select_ranks(d, rk, True), pos[rk])
        for q in (5, 10.0, 95.0):
            assert eng.percentile(d, q) == np.percentile(a, q)


def test_select_f64_is_exact(eng):
    """b200vec_select_f64 (np.median of detect_packet_bounds' smoothed energies, utils.py:818): exact order
    statistics of float64 data — duplicates, negatives, values that differ only in their lowest bits."""
    rng = np.random.default_rng(12)
    for n in (1, 2, 33, 5000, 400_001):
        a = rng.standard_normal(n) * 10.0 ** rng.integers(-12, 3, n)
        a[rng.integers(0, n, n // 4)] = 0.0
        a[rng.integers(0, n, n // 5)] = 1.0 + np.finfo(np.float64).eps * rng.integers(0, 4, n // 5)
        srt = np.sort(a)
        d = torch.from_numpy(a).cuda()
        for ranks in ([0], [n - 1], [n // 2], sorted({max(0, n // 2 - 1), n // 2}), sorted({0, n // 3, n // 2, n - 1})):
            got = eng.select_f64(d, ranks)
            assert got == [float(srt[r]) for r in ranks], (n, ranks, got)


@pytest.mark.parametrize("name", ["wf_2p5ms_shift_norm", "cfg1_100ms"])
def test_normalize_spectrogram_matches_oracle(eng, name):
    import vector_b200 as vb
    r = _build(name, eng)
    _, _, S = O.create_spectrogram(r["want"], 56e6, time_resolution_us=1)
    db_ref, vmin_ref, vmax_ref = O.normalize_spectrogram(S, 5, 95, 25)
    db, vmin, vmax = vb.normalize_spectrogram(S, 5, 95, 25)
    assert db.shape == db_ref.shape and db.dtype == db_ref.dtype
    assert np.max(np.abs(db - db_ref)) <= 1e-4               # same floor, log10f round-off only
    assert abs(vmin - vmin_ref) <= 1e-4 and abs(vmax - vmax_ref) <= 1e-4
    out = vb.normalize_spectrogram(np.zeros((0, 0), np.float32))
    assert out[0].size == 0 and out[1:] == (0, 0)            # utils.py:359-360


# ------------------------------------------------------------------------------ K6 .wv
def test_wv_payload_and_file_match_oracle(eng, tmp_path, golden_meta):
    import vector_b200 as vb
    r = _build("wf_2p5ms_shift_norm", eng)
    q = eng.wv_quantize(torch.from_numpy(r["want"]).cuda()).cpu().numpy()
    assert np.array_equal(q, O.wv_quantize(r["want"]))
    p = tmp_path / "v.wv"
    vb.save_vector_wv(r["want"], str(p), 56e6)
    raw = p.read_bytes()
    assert raw.startswith(b"{TYPE: SMU-WV,0}")                              # tests/test_utils.py:163-170
    hdr_ref = golden_meta["misc"]["wv_header_wf_2p5ms_shift_norm"]
    date = hdr_ref.split("{DATE: ")[1].split("}")[0]
    hdr = raw[:raw.index(b":#") + 2].decode()
    assert hdr.split("{DATE: ")[0] == hdr_ref.split("{DATE: ")[0]
    assert hdr.split("}", 3)[3] == hdr_ref.split("}", 3)[3]                  # everything after DATE
    assert raw[-1:] == b"}" and len(raw) == len(hdr) + 4 * r["n"] + 1
    assert np.array_equal(np.frombuffer(raw[len(hdr):-1], np.int16), O.wv_quantize(r["want"]))


# ------------------------------------------------------------------------------ drop-in boundary
def test_install_rebinds_and_generate_vector_runs(eng, tmp_path, monkeypatch):
    """install() on stand-in modules shaped like the reference's (same names, inline
    generate_vector): names are rebound and the replaced method writes the same .mat the
    oracle's build gives."""
    import types
    import scipy.io as sio
    from unittest.mock import MagicMock
    from vector_b200 import generate
    fake_utils = types.ModuleType("utils")
    for nme in generate._PATCHED_NAMES:
        setattr(fake_utils, nme, lambda *a, **k: (_ for _ in ()).throw(AssertionError("not rebound")))
    fake_gui = types.ModuleType("unified_gui")
    fake_gui.TARGET_SAMPLE_RATE = 56e6
    fake_gui.messagebox = MagicMock()
    fake_gui.plot_spectrogram = MagicMock()
    fake_gui.create_spectrogram = fake_utils.create_spectrogram

    class UnifiedVectorApp:
        def validate_packet_timing(self, markers, pcs):
            from vector_b200 import validate_packet_timing
            return validate_packet_timing(markers, [pc.get_config() for pc in pcs])

        def generate_vector(self, output_format="mat"):
            raise AssertionError("not replaced")
    fake_gui.UnifiedVectorApp = UnifiedVectorApp
    done = generate.install(fake_utils, fake_gui, None)
    assert ("unified_gui", "UnifiedVectorApp.generate_vector") in done
    import vector_b200.api as api
    assert fake_utils.create_spectrogram is api.create_spectrogram and fake_gui.create_spectrogram is api.create_spectrogram

    case = cases.build_cases()["wf_2p5ms_shift_norm"]
    pk = O.packets_for(case["configs"])
    cfgs = []
    for c, (y, pre) in zip(case["configs"], pk):
        path = tmp_path / c["file"]
        sio.savemat(str(path), {"Y": y.astype(np.complex128), "pre_samples": pre})
        cfgs.append(dict(c, file=str(path)))
    app = UnifiedVectorApp()
    app.packet_configs = [types.SimpleNamespace(get_config=lambda c=c: dict(c)) for c in cfgs]
    app.vector_length_var = types.SimpleNamespace(get=lambda: "2.5")
    app.normalize = types.SimpleNamespace(get=lambda: True)
    monkeypatch.chdir(tmp_path)
    app.generate_vector("mat")
    assert not fake_gui.messagebox.showerror.called, fake_gui.messagebox.showerror.call_args
    files = [f for f in os.listdir(tmp_path) if f.startswith("vector_") and f.endswith(".mat")]
    assert len(files) == 1
    y, pre = api.load_packet_info(str(tmp_path / files[0]))
    want = O.build_vector(pk, case["configs"], 140000, True)[0]
    assert pre == 0 and P.rel_l2(y, want) <= 3e-7
    kw = fake_gui.plot_spectrogram.call_args.kwargs
    assert "Timing Validation" in kw["title"] and len(kw["packet_markers"]) == 4
    app.generate_vector("wv")
    assert any(f.endswith(".wv") for f in os.listdir(tmp_path))


def test_main_flavour_resamples(eng, golden_meta, golden_arrays):
    """main.py:214-321 flavour: packets at 20/40/56 MHz resampled to 56 MHz first."""
    from vector_b200 import generate
    g = golden_meta["builds"]["main_resample"]
    pk, cfgs = [], []
    for i, c in enumerate(g["configs"]):
        kind = c["file"].split("_")[1].split(".")[0]
        y = O.synth_packet(kind, fs=c["sample_rate"], body_s=1e-4, pre=20, seed=201 + i)
        pk.append((y.astype(np.complex64), 20))
        cfgs.append(c)
    n = O.total_samples_main("0.004")
    want = O.build_vector(pk, cfgs, n, True, resample=True)[0]
    vec, _, _, _ = generate.build_vector_device(cfgs, n, True, packets=pk, resample=True, engine=eng)
    got = vec.cpu().numpy()
    assert P.rel_l2(got, want) <= P.VEC_REL_L2
    assert P.rel_l2(got, golden_arrays["main_resample__full"]) <= P.VEC_REL_L2


def test_invalid_arguments_fail_loudly(eng):
    from vector_b200._lib import B200VecError
    with pytest.raises(B200VecError):
        eng.build_vector([(0, 0, 1, 0)], 1000, False)               # period must be positive
    with pytest.raises(B200VecError):
        eng.build_vector([(0, 10, 1, 40)], 1000, False)             # uncommitted slot
    x = torch.zeros(4096, dtype=torch.complex64, device="cuda")
    with pytest.raises(B200VecError):
        eng.stft(x, "hann", 1000, 0, 1000, 1.0)                     # nfft not a power of two


# ------------------------------------------------------------------------------ ranged max pass
@pytest.mark.parametrize("name,n", [("cfg1_100ms", 5_600_000), ("overlap_5ms", 280_000)])
def test_ranged_max_equals_full_max(eng, name, n):
    """Pass A over hostmath.max_eval_ranges (one common period + head/tail, split over ranks) gives
    the same max|v|^2 double as pass A over the whole vector, and the oracle's max."""
    from vector_b200 import generate
    from vector_b200._lib import BUILD_MAXONLY
    case = cases.build_cases()[name]
    pk = O.packets_for(case["configs"])
    info = generate.prepare_bank(eng, case["configs"], pk, 56e6)
    rows = generate.placement_rows(case["configs"], info, n, 56e6)
    descs, nd = eng.make_descs(rows)
    full = torch.zeros(1, dtype=torch.float64, device=eng.device)
    eng.build_pass(descs, nd, 0, n, None, BUILD_MAXONLY, maxsq=full)
    ranges = eng.max_ranges(rows, n)
    if name == "cfg1_100ms":
        assert sum(hi - lo for lo, hi in ranges) < n // 2          # periods 1.0/1.2/1.5 ms: lcm 6 ms
    for world in (1, 3):
        m = torch.zeros(1, dtype=torch.float64, device=eng.device)
        for r in range(world):
            eng.max_pass(descs, nd, eng.max_ranges(rows, n, world, r), maxsq=m)
        assert torch.equal(m, full)
    want = O.build_vector(pk, case["configs"], n, False)[0]
    assert abs(np.sqrt(full.item()) - float(np.max(np.abs(want)))) <= 2e-7 * np.sqrt(full.item())   # np.abs is a float32 hypot


@pytest.mark.parametrize("n", [2, 3, 4, 7, 4097, 560_003])
def test_save_vector_from_device_matches_host_writer(eng, tmp_path, n):
    """save_vector(CUDA tensor): device plane split (K7) + pinned D2H writes the same bytes as the host path."""
    from vector_b200 import api
    rng = np.random.default_rng(n)
    v = (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)
    api.save_vector(v, str(tmp_path / "h.mat"))
    api.save_vector(torch.from_numpy(v).to(eng.device), str(tmp_path / "d.mat"))
    h, d = (tmp_path / "h.mat").read_bytes(), (tmp_path / "d.mat").read_bytes()
    assert len(h) == len(d) and h[128:] == d[128:]
    y, pre = api.load_packet_info(str(tmp_path / "d.mat"))
    assert np.array_equal(y, v) and pre == 0


# ------------------------------------------------------------------------------ cfg4 / cfg5 shapes
def _cfg4_case(ms):
    """SURVEY.md §8d cfg4 shortened to `ms` milliseconds: 200 MSa/s, the six packets made at 56 MSa/s and
    resampled to 200 MSa/s (L' = int(56100*200/56) = 200357, K0), shifts +20/-35/+10/-60/+45/-80 MHz,
    dense periods 2.0/1.5/1.0/0.8/0.5/0.4 ms (shorter than the 1 ms packets: instances of one packet overlap)."""
    from vector_b200 import synth
    fs = 200e6
    cfgs, pk = synth.scenario(synth.CFG4_ROWS, 56e6)
    for c in cfgs:
        c["sample_rate"] = 56e6
    return fs, cfgs, pk, int(ms * 1e-3 * fs)


def test_cfg4_shape_matches_oracle_and_shards(eng):
    from vector_b200 import dist as vdist, generate
    from vector_b200._lib import BUILD_MAXONLY, BUILD_SCALE
    fs, cfgs, pk, n = _cfg4_case(30)                                   # 6 M samples; the trains share a 12 ms period
    want, _, counts_ref, _ = O.build_vector(pk, cfgs, n, True, fs=fs, resample=True)
    info = generate.prepare_bank(eng, cfgs, pk, fs, resample=True)
    assert [i[0] for i in info] == [200357] * 6
    rows = generate.placement_rows(cfgs, info, n, fs)
    assert [r[2] for r in rows] == counts_ref
    plan = vdist.plan_slab(n, 1, 0, 1024, 512, 1, align=512)
    vec, spec, _ = vdist.build_stft_slab(eng, rows, plan, True, None, fs=fs)          # ONE launch, in-kernel max
    assert P.rel_l2(vec.cpu().numpy(), want) <= P.VEC_REL_L2
    _, _, Sr = O.full_rate_spectrogram(want, fs)
    assert P.spec_db_error(spec.cpu().numpy(), Sr) <= P.SPEC_DB
    # time-sharded over 2 / 4 / 8 ranks (emulated one after another): pass A shares MAX-combine to the
    # same double, slabs and spectrogram columns concatenate bit for bit
    descs, nd = eng.make_descs(rows)
    for world in (2, 4, 8):
        plans = [vdist.plan_slab(n, world, r, 1024, 512, 1, align=512) for r in range(world)]
        maxsq = torch.zeros(1, dtype=torch.float64, device=eng.device)
        for r in range(world):
            eng.max_pass(descs, nd, eng.max_ranges(rows, n, world, r), maxsq=maxsq)
        assert torch.equal(maxsq, eng._maxsq)
        vecs, specs = [], []
        for pl in plans:
            v = torch.empty(pl.hi - pl.lo, dtype=torch.complex64, device=eng.device)
            sp = eng.alloc_spectrogram(1024, pl.frame_hi - pl.frame_lo, frame_major=True)
            eng.build_stft1024(descs, nd, pl.lo, pl.hi, v, BUILD_SCALE, pl.frame_lo, pl.frame_hi - pl.frame_lo, sp,
                               fs=fs, maxsq=maxsq)
            vecs.append(v); specs.append(sp)
        assert torch.equal(torch.cat(vecs), vec)
        assert torch.equal(torch.cat(specs, dim=1), spec)


@pytest.mark.parametrize("v", [0, 1, 2, 3, 100, 255])
def test_cfg5_scenarios_match_oracle(eng, v):
    """vector v of the 256-vector parameter sweep (3 packets, 100 ms @ 56 MSa/s, shift /
    period / start drawn from default_rng(5000 + v)); independent vectors, no collective."""
    from vector_b200 import generate, synth
    cfgs, pk = synth.scenario(synth.cfg5_rows(v), 56e6)
    n = 5_600_000
    want, _, counts_ref, _ = O.build_vector(pk, cfgs, n, True)
    vec, _, counts, _ = generate.build_vector_device(cfgs, n, True, packets=pk, engine=eng)
    assert counts == counts_ref
    assert P.rel_l2(vec.cpu().numpy(), want) <= P.VEC_REL_L2
    plan, rows, fvec, fspec = _fused(eng, cfgs, pk, n, True)
    assert torch.equal(fvec, vec)
    _, _, Sr = O.full_rate_spectrogram(want, 56e6)
    assert P.spec_db_error(fspec.cpu().numpy(), Sr) <= P.SPEC_DB


def test_peer_max_protocol_single_rank(eng):
    """b200vec_build_stft1024_fm_norm_peer with world = 1 (the block is local memory): the epoch /
    slot-parity / arrival protocol runs for several steps and gives the in-kernel-max result every time."""
    from vector_b200 import dist as vdist, generate
    case = cases.build_cases()["cfg1_100ms"]
    pk = O.packets_for(case["configs"])
    n = 1_000_000
    info = generate.prepare_bank(eng, case["configs"], pk, 56e6)
    rows = generate.placement_rows(case["configs"], info, n, 56e6)
    plan = vdist.plan_slab(n, 1, 0, 1024, 512, 1, align=512)
    vec0, spec0, m0 = vdist.build_stft_slab(eng, rows, plan, True, None, fs=56e6)
    m0 = m0.clone()
    block = torch.zeros(16, dtype=torch.int32, device=eng.device)
    peer = vdist.PeerMax(blocks=([block.data_ptr()], 1, 0))
    for step in range(1, 5):
        vec, spec, m = vdist.build_stft_slab(eng, rows, plan, True, None, fs=56e6, peer=peer)
        assert peer.epoch == step
        assert torch.equal(m, m0) and torch.equal(vec, vec0) and torch.equal(spec, spec0)
    assert int(block[4]) == 4                                           # one arrival per epoch


# ------------------------------------------------------------------------------ either side of the path (§8f rank 4)
@pytest.mark.parametrize("kind,ratio", [("tone", 0.2), ("chirp", 0.2), ("bpsk", 0.2), ("noise", 0.2), ("multitone", 0.5),
                                        ("tone", 0.05)])
def test_detect_packet_bounds_matches_reference_statement(eng, kind, ratio):
    """utils.detect_packet_bounds: packet embedded in a noise floor, indices exact."""
    import vector_b200 as vb
    rng = np.random.default_rng(7)
    body = O.synth_packet(kind, body_s=2e-4, pre=0, seed=11)
    sig = 0.01 * (rng.standard_normal(40_000) + 1j * rng.standard_normal(40_000))
    sig[17_003:17_003 + len(body)] += body
    sig = sig.astype(np.complex64)
    want = O.detect_packet_bounds(sig, 56e6, ratio)
    got = vb.detect_packet_bounds(sig, 56e6, ratio)
    assert tuple(int(v) for v in got) == want
    got_dev = vb.detect_packet_bounds(torch.from_numpy(sig).to(eng.device), 56e6, ratio)
    assert tuple(got_dev) == want
    assert 17_003 - 56 <= want[0] <= 17_003 + 56


def test_detect_packet_bounds_smoothing_and_degenerate_cases(eng):
    import vector_b200 as vb
    rng = np.random.default_rng(3)
    sig = (rng.standard_normal(5000) + 1j * rng.standard_normal(5000)).astype(np.complex64)
    sm, mx = eng.box_energy(torch.from_numpy(sig).to(eng.device), 56)
    ref = np.convolve(np.abs(sig) ** 2, np.ones(56) / 56, mode="same")
    assert np.max(np.abs(sm.cpu().numpy() - ref)) <= 1e-6 * ref.max()            # numpy's |complex64| is a 2-ulp float32 hypot
    assert abs(mx.item() - ref.max()) <= 1e-6 * ref.max()
    z = np.zeros(1000, np.complex64)
    assert tuple(vb.detect_packet_bounds(z, 56e6)) == O.detect_packet_bounds(z, 56e6)     # threshold 0: every sample
    short = np.zeros(10, np.complex64)                                                   # shorter than the 1 us window
    assert tuple(vb.detect_packet_bounds(short, 56e6)) == O.detect_packet_bounds(short, 56e6)
    with pytest.raises(ValueError):
        vb.detect_packet_bounds(np.zeros(0, np.complex64), 56e6)                          # np.max of an empty array


@pytest.mark.parametrize("shape", [(1024, 37), (2048, 5), (1, 9), (7, 1)])
def test_median_filter_rows2_matches_scipy(eng, shape):
    import vector_b200 as vb
    rng = np.random.default_rng(shape[0])
    a = (rng.standard_normal(shape) * 20 - 60).astype(np.float32)
    want = O.median_filter_rows2(a)
    assert np.array_equal(vb.median_filter_rows2(a), want)
    d = torch.from_numpy(a).to(eng.device)
    assert np.array_equal(vb.median_filter_rows2(d).cpu().numpy(), want)                 # C-order rows
    fm = torch.from_numpy(np.ascontiguousarray(a.T)).to(eng.device).t()                  # frame-major memory, same values
    out = vb.median_filter_rows2(fm)
    assert out.stride() == fm.stride() and np.array_equal(out.cpu().numpy(), want)

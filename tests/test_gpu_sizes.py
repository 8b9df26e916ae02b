"""Parity at the BASELINE sizes (VERDICT r1 "parity gaps"): the oracle — not the engine's own other
kernels — is the judge of the full-size results.

  cfg2 + cfg3(full)   56 M samples, all 109 374 frames of the fused step vs scipy (0.01 dB, band rule,
                      exact-zero frames)
  cfg2 preview        create_spectrogram + normalize_spectrogram(5, 95, 25) vs the goldens the REAL
                      reference produced (tests/golden: spec_sub, db_sub, vmin, vmax, shape, axes)
  cfg3 normalisation  normalize_spectrogram on the 112 M-element full-rate map vs the oracle
                      (np.percentile's float32 rank arithmetic above 2^24)
  cfg4                10 s @ 200 MSa/s (2e9 samples) by windows: the oracle builds [lo, hi) windows with the
                      reference's own adds; the GPU slab matches to 1e-5 and its spectrogram columns to 0.01 dB
  cfg5                all 256 scenario vectors of the sweep, built by the batch entry point, vs the oracle
"""
import numpy as np
import pytest
import scipy.signal

from oracle import ref_port as O
from tests import parity as P
from tests.golden import cases

pytestmark = [pytest.mark.gpu, pytest.mark.slow]

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from vector_b200.engine import get_engine
    return get_engine(0)


@pytest.fixture(scope="module")
def cfg2(eng):
    """cfg2 built once: oracle vector, fused GPU vector + full-rate spectrogram."""
    from vector_b200 import dist as vdist, generate
    cfgs = O.cfg2_configs()
    pk = O.packets_for(cfgs)
    n = 56_000_000
    want, _, counts, _ = O.build_vector(pk, cfgs, n, True)
    info = generate.prepare_bank(eng, cfgs, pk, 56e6)
    rows = generate.placement_rows(cfgs, info, n, 56e6)
    plan = vdist.plan_slab(n, 1, 0, 1024, 512, 1, align=512)
    vec, spec, _ = vdist.build_stft_slab(eng, rows, plan, True, None, fs=56e6)
    assert [r[2] for r in rows] == counts == [10, 20, 49, 97, 192, 475]
    return dict(n=n, want=want, vec=vec, spec=spec)


def _psd_f64_chunked(x, fs, frames_per_chunk=4096):
    """O.stft_psd_f64 for the hann 1024/512 'density' shape, a few thousand frames at a time."""
    nf = (len(x) - 1024) // 512 + 1
    for f0 in range(0, nf, frames_per_chunk):
        f1 = min(nf, f0 + frames_per_chunk)
        seg = x[f0 * 512: (f1 - 1) * 512 + 1024]
        yield f0, f1, O.stft_psd_f64(seg, fs, "hann", 1024, 512, 1024, "density")[2]


def test_cfg3_full_size_matches_oracle(eng, cfg2):
    """All 109 374 frames of the fused step against scipy.signal.spectrogram on the oracle's vector."""
    assert P.rel_l2(cfg2["vec"].cpu().numpy(), cfg2["want"]) <= 3e-7
    _, _, Sr = O.full_rate_spectrogram(cfg2["want"], 56e6)
    S = cfg2["spec"].cpu().numpy()
    assert S.shape == Sr.shape == (1024, 109_374) and S.dtype == np.float32
    assert P.spec_db_error(S, Sr) <= P.SPEC_DB
    zero_cols = ~np.any(Sr, axis=0)
    assert zero_cols.sum() > 30_000 and not np.any(S[:, zero_cols])            # exact-zero frames stay exact zeros
    # band rule below -80 dBc, vs an fp64 truth, accumulated chunk by chunk (max and a pooled p99 estimate)
    peak = float(Sr.max())                               # band edges: the reference's own peak (1e-6 from the truth's)
    worst = {}
    for f0, f1, St in _psd_f64_chunked(cfg2["want"], 56e6, 8192):
        with np.errstate(divide="ignore", invalid="ignore"):
            lvl = 10 * np.log10(St / peak)
            eo = np.abs(10 * np.log10(S[:, f0:f1].astype(np.float64)) - 10 * np.log10(St))
            er = np.abs(10 * np.log10(Sr[:, f0:f1].astype(np.float64)) - 10 * np.log10(St))
        for top in (-80, -100, -120):
            m = (lvl < top) & (lvl >= top - 20) & np.isfinite(eo) & np.isfinite(er)
            if m.sum() >= 1000:
                a, b = np.percentile(eo[m], 99), np.percentile(er[m], 99)
                w = worst.setdefault(top, [0.0, 0.0])
                w[0], w[1] = max(w[0], a), max(w[1], b)
    assert worst, "no bins below -80 dBc?"
    for top, (ours, refs) in worst.items():
        assert ours <= 2 * refs + 1e-4, (top, ours, refs)


def test_cfg2_preview_matches_reference_goldens(eng, cfg2, golden_meta, golden_arrays):
    """create_spectrogram(vector, 56e6, time_resolution_us=1) + normalize_spectrogram(., 5, 95, 25) — the calls
    generate_vector / plot_spectrogram make (unified_gui.py:1800, utils.py:431) — against what the REAL reference
    returned for cfg2 (tests/golden/make_golden.py)."""
    import vector_b200 as vb
    g = golden_meta["builds"]["cfg2_1000ms"]["spectrogram"]
    f, t, S_dev = vb.create_spectrogram_device(cfg2["vec"], 56e6, time_resolution_us=1, engine=eng)
    S = S_dev.cpu().numpy()
    assert list(S.shape) == g["shape"] == [1024, 9995] and S.dtype == np.float32
    assert (float(f[0]), float(f[-1]), float(t[0]), float(t[-1])) == (g["f0"], g["f1"], g["t0"], g["t1"])
    assert cases.sha1(f) == g["freqs_sha1"] and cases.sha1(t) == g["times_sha1"]
    sub = S[::16, ::max(1, S.shape[1] // 64)].astype(np.float64)
    ref = golden_arrays["cfg2_1000ms__spec_sub"].astype(np.float64)
    m = ref >= g["peak"] * P.SPEC_FLOOR
    assert m.sum() > 500
    assert np.max(np.abs(10 * np.log10(sub[m]) - 10 * np.log10(ref[m]))) <= P.SPEC_DB
    assert abs(float(S.sum(dtype=np.float64)) - g["total"]) <= 1e-5 * g["total"]
    # the oracle on the oracle's vector agrees too (same shape, 0.01 dB)
    _, _, Sr = O.create_spectrogram(cfg2["want"], 56e6, time_resolution_us=1)
    assert P.spec_db_error(S, Sr) <= P.SPEC_DB
    db, vmin, vmax = eng.normalize_spectrogram(S_dev, 5, 95, 25)
    assert abs(vmin - g["vmin"]) <= 0.01 and abs(vmax - g["vmax"]) <= 0.01
    dsub = db.cpu().numpy()[::16, ::max(1, S.shape[1] // 64)]
    assert np.max(np.abs(dsub - golden_arrays["cfg2_1000ms__db_sub"])) <= 0.01
    db_r, vmin_r, vmax_r = O.normalize_spectrogram(Sr, 5, 95, 25)
    mask = Sr >= Sr.max() * P.SPEC_FLOOR                       # SURVEY §8c rule (iv): the dB map on the >= -80 dBc bins
    assert np.max(np.abs(db.cpu().numpy()[mask] - db_r[mask])) <= 0.01
    assert np.max(np.abs(db.cpu().numpy() - db_r)) <= 0.1      # everywhere else the floor term dominates
    assert abs(vmin - vmin_r) <= 0.01 and abs(vmax - vmax_r) <= 0.01


def test_normalize_full_rate_map_matches_oracle(eng, cfg2):
    """normalize_spectrogram on the 112 M-element cfg3 map: np.percentile evaluates its virtual index in float32
    (hostmath.percentile_ranks), which above 2^24 elements decides the neighbours — must match numpy exactly."""
    S_dev = cfg2["spec"]
    Sr = np.ascontiguousarray(S_dev.cpu().numpy())            # same values, C order, for the oracle
    db_r, vmin_r, vmax_r = O.normalize_spectrogram(Sr, 5, 95, 25)
    db, vmin, vmax = eng.normalize_spectrogram(S_dev, 5, 95, 25)
    assert db.shape == S_dev.shape
    assert abs(vmin - vmin_r) <= 1e-3 and abs(vmax - vmax_r) <= 1e-3
    assert float(np.max(np.abs(db.cpu().numpy() - db_r))) <= 1e-3
    # the floor itself: exact order statistics of the positive elements
    pos = Sr[Sr > 0]
    floor_np = np.percentile(pos, 5)
    n, n_pos = Sr.size, pos.size
    flat = S_dev.contiguous().view(-1)
    assert eng.select_begin(flat, use_abs=True) == n_pos
    from vector_b200 import hostmath as H
    lo, hi, g = H.percentile_ranks(n_pos, 5, np.float32)
    v = eng.select_finish(flat, [n - n_pos + lo, n - n_pos + hi, 0, n - 1, n // 2])
    assert H.lerp(v[0], v[1], g) == floor_np
    assert v[2] == 0.0 and v[3] == Sr.max() and v[4] == np.partition(Sr.reshape(-1), n // 2)[n // 2]


def test_cfg4_full_length_by_windows(eng):
    """cfg4 at FULL size on one GPU (2e9 samples: 16 GB vector + 16 GB spectrogram), checked window by window
    against the oracle: head, tail, the boundaries of the periodic tiling and of 8-rank slabs (+- halo), interior."""
    from vector_b200 import dist as vdist, generate, hostmath as H, synth
    free, _ = torch.cuda.mem_get_info()
    if free < 40e9:
        pytest.skip("needs ~34 GB of free device memory")
    fs = 200e6
    n = int(10.0 * fs)
    cfgs, pk = synth.scenario(synth.CFG4_ROWS, 56e6)
    for c in cfgs:
        c["sample_rate"] = 56e6
    info = generate.prepare_bank(eng, cfgs, pk, fs, resample=True)
    rows = generate.placement_rows(cfgs, info, n, fs)
    prep = O.prepared_packets(pk, cfgs, fs, resample=True)
    assert [len(y) for y, _ in prep] == [i[0] for i in info] == [200357] * 6
    plan = vdist.plan_slab(n, 1, 0, 1024, 512, 1, align=512)
    vec, spec, maxsq = vdist.build_stft_slab(eng, rows, plan, True, None, fs=fs)
    tp = H.tile_plan(rows, [i[0] for i in info], n, 0, n)
    assert tp is not None and tp["T"] == 2_400_000           # periods 2/1.5/1/0.8/0.5/0.4 ms: lcm 12 ms
    # the global max, from the oracle: one common period + head/tail (hostmath.max_eval_ranges)
    m_ref = np.float32(0)
    for lo, hi in H.max_eval_ranges(rows, [i[0] for i in info], n):
        for a in range(lo, hi, 4_000_000):
            m_ref = max(m_ref, np.max(np.abs(O.build_vector_window(prep, cfgs, n, a, min(hi, a + 4_000_000), fs))))
    m = np.float32(np.sqrt(maxsq.item()))
    assert abs(m - m_ref) <= 2e-6 * m_ref                      # the resampled packets differ by ~2e-7 (fp64 Bluestein vs scipy's complex64 FFTs)
    w = 1 << 20
    starts = {0, n - w, tp["x1"] - w // 2, tp["x2"] - w // 2, 1_234_567_890 // 512 * 512, 700_000_000 // 512 * 512}
    for r in range(1, 8):
        starts.add((H.slab_bounds(n, 8, r, 512)[0] - w // 2) // 512 * 512)
    rng = np.random.default_rng(4)
    for s in rng.integers(0, n - w, 4):
        starts.add(int(s) // 512 * 512)
    for lo in sorted({max(0, min(n - w, s)) // 512 * 512 for s in starts}):
        want = O.build_vector_window(prep, cfgs, n, lo, lo + w, fs) / m_ref
        got = vec[lo:lo + w].cpu().numpy()
        assert P.rel_l2(got, want) <= P.VEC_REL_L2, lo
        _, _, Sr = O.full_rate_spectrogram(want, fs)
        S = spec[:, lo // 512: lo // 512 + Sr.shape[1]].cpu().numpy()
        assert P.spec_db_error(S, Sr) <= P.SPEC_DB, lo
    del vec, spec


def test_cfg5_all_256_vectors_match_oracle(eng):
    """The whole 256-vector sweep through the batch entry point (two device calls), every vector vs the oracle:
    instance counts exact, rel-L2 <= 1e-5."""
    from vector_b200 import generate, synth
    n = 5_600_000
    bank = synth.bank(56e6)
    scen, packets = [], {}
    for v in range(256):
        cfgs = synth.configs(synth.cfg5_rows(v))
        scen.append(cfgs)
        for c in cfgs:
            packets[c["file"]] = bank[c["kind"]]
    vecs, markers, counts, maxsq = generate.build_batch_device(scen, n, True, packets=packets, engine=eng)
    assert vecs.shape == (256, n) and maxsq.shape == (256,)
    worst = 0.0
    for v in range(256):
        pk = [packets[c["file"]] for c in scen[v]]
        want, mk_ref, counts_ref, max_ref = O.build_vector(pk, scen[v], n, True)
        assert counts[v] == counts_ref, v
        assert [m[0] for m in markers[v]] == [m[0] for m in mk_ref]
        w = torch.from_numpy(want).to(eng.device)
        err = (torch.linalg.vector_norm((vecs[v] - w).to(torch.complex128)) /
               torch.linalg.vector_norm(w.to(torch.complex128))).item()
        worst = max(worst, err)
        assert err <= P.VEC_REL_L2, (v, err)
        mm = np.float32(np.sqrt(maxsq[v].item()))
        assert abs(mm - max_ref) <= np.spacing(max_ref), v
    assert worst <= 1e-6

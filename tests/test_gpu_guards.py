"""Out-of-bounds guards: compute-sanitizer is closed on this pool, so every output-writing entry point is
called with its output INSIDE a larger buffer whose margins hold a sentinel pattern; a kernel that stores one
element before or past its output changes the pattern.  Sizes are ragged on purpose (partial tiles, partial
frames, odd lengths, misaligned tails).  All calls go through the C-ABI (Engine is a ctypes shim)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")

PAD = 4096                      # guard elements either side (a multiple of every alignment the ABI asks for)
SENT = 0x5A                     # every guard byte


@pytest.fixture(scope="module")
def eng():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from vector_b200.engine import get_engine
    return get_engine(0)


class Guarded:
    """`n` elements of `dtype` with PAD sentinel elements before and after; .view is the part a kernel may write."""

    def __init__(self, n, dtype, device, fill_view=True):
        self.esz = torch.empty(0, dtype=dtype).element_size()
        self.n = n
        self.raw = torch.full(((n + 2 * PAD) * self.esz,), SENT, dtype=torch.uint8, device=device)
        self.view = self.raw[PAD * self.esz:(PAD + n) * self.esz].view(dtype)
        if not fill_view:
            self.view.zero_()

    def check(self, what=""):
        torch.cuda.synchronize()
        head = self.raw[:PAD * self.esz]
        tail = self.raw[(PAD + self.n) * self.esz:]
        assert bool((head == SENT).all()), f"{what}: stored before the output"
        assert bool((tail == SENT).all()), f"{what}: stored past the output"

    def written(self):
        """True when no element of the output still holds the sentinel pattern (the kernel covered all of it)."""
        b = self.raw[PAD * self.esz:(PAD + self.n) * self.esz].view(-1, self.esz)
        return not bool((b == SENT).all(dim=1).any())


def _rows(eng, n, fs=56e6):
    from vector_b200 import generate, synth
    cfgs, pk = synth.scenario(synth.CFG2_ROWS[:4], fs)
    info = generate.prepare_bank(eng, cfgs, pk, fs)
    return generate.placement_rows(cfgs, info, n, fs)


@pytest.mark.parametrize("n", [1, 2, 511, 513, 4097, 300_001, 1_234_567])
def test_vector_builders_stay_inside_the_output(eng, n):
    from vector_b200._lib import BUILD_STORE
    rows = _rows(eng, n)
    descs, nd = eng.make_descs(rows)
    g = Guarded(n, torch.complex64, eng.device)
    eng.build_pass(descs, nd, 0, n, g.view, BUILD_STORE)
    g.check("b200vec_build STORE"); assert g.written()
    g = Guarded(n, torch.complex64, eng.device)
    eng.build_norm(descs, nd, 0, n, g.view, [(0, n)])
    g.check("b200vec_build_norm"); assert g.written()
    lo = min(n - 1, 300)                                   # a slab that starts inside the vector
    g = Guarded(n - lo, torch.complex64, eng.device)
    eng.build_pass(descs, nd, lo, n, g.view, BUILD_STORE)
    g.check("b200vec_build STORE, slab"); assert g.written()


@pytest.mark.parametrize("n", [1024, 1025, 1535, 1536, 2047, 300_001, 1_000_448])
def test_fused_step_stays_inside_vector_and_spectrogram(eng, n):
    from vector_b200._lib import BUILD_STORE
    rows = _rows(eng, n)
    descs, nd = eng.make_descs(rows)
    nf = (n - 1024) // 512 + 1
    for norm in (False, True):
        gv = Guarded(n, torch.complex64, eng.device)
        gs = Guarded(nf * 1024, torch.float32, eng.device)
        spec = gs.view.view(nf, 1024).t()                  # frame-major [1024, nf]
        eng.build_stft1024(descs, nd, 0, n, gv.view, BUILD_STORE, 0, nf, spec,
                           norm_ranges=[(0, n)] if norm else None)
        gv.check(f"fused step (norm={norm}), vector"); gs.check(f"fused step (norm={norm}), spectrogram")
        assert gv.written() and gs.written()
    # the C-order tile kernel: rows of `pitch` floats, only the first nf of each row belong to the output
    pitch = (nf + 31) // 32 * 32 + 32
    gv = Guarded(n, torch.complex64, eng.device)
    gs = Guarded(1024 * pitch, torch.float32, eng.device)
    spec = gs.view.view(1024, pitch)[:, :nf]
    eng.build_stft1024(descs, nd, 0, n, gv.view, BUILD_STORE, 0, nf, spec)
    gv.check("tile kernel, vector"); gs.check("tile kernel, spectrogram")
    torch.cuda.synchronize()
    margin = gs.raw.view(torch.float32)[PAD:PAD + 1024 * pitch].view(1024, pitch)[:, nf:]
    assert bool((margin.contiguous().view(torch.uint8) == SENT).all()), "tile kernel wrote past a row's frames"


@pytest.mark.parametrize("n,nperseg,hop,nfft,stride", [(3000, 256, 56, 512, 1), (40_001, 512, 8, 1024, 1), (700_003, 934, 934, 1024, 6),
                                                        (100_000, 64, 48, 128, 1), (9_000, 1024, 512, 1024, 1), (70_000, 6720, 3360, 8192, 1)])
def test_stft_kernels_stay_inside_the_spectrogram(eng, n, nperseg, hop, nfft, stride):
    rng = np.random.default_rng(n)
    x = torch.from_numpy((rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)).to(eng.device)
    n_dec = (n + stride - 1) // stride
    nf = (n_dec - nperseg) // hop + 1
    pitch = (nf + 31) // 32 * 32 + 32
    gs = Guarded(nfft * pitch, torch.float32, eng.device)
    out = gs.view.view(nfft, pitch)[:, :nf]
    eng.stft(x, "hann", nperseg, nperseg - hop, nfft, 56e6, stride=stride, out=out, frame_major=False)
    gs.check("b200vec_stft")
    torch.cuda.synchronize()
    margin = gs.raw.view(torch.float32)[PAD:PAD + nfft * pitch].view(nfft, pitch)[:, nf:]
    assert bool((margin.contiguous().view(torch.uint8) == SENT).all()), "b200vec_stft wrote past a row's frames"
    if (nperseg, hop, nfft, stride) == (1024, 512, 1024, 1):                # the frame-major warp-per-frame kernel
        gs = Guarded(nf * 1024, torch.float32, eng.device)
        eng.stft(x, "hann", 1024, 512, 1024, 56e6, out=gs.view.view(nf, 1024).t())
        gs.check("b200vec_stft1024_fm"); assert gs.written()
    if nfft <= 8192:                                                        # the float64 kernel
        xd = x.to(torch.complex128)
        S = eng.stft_f64(xd, "hann", nperseg, nperseg - hop, nfft, 56e6, stride=stride)
        assert S.shape == (nfft, nf) and bool(torch.isfinite(S).all())


@pytest.mark.parametrize("n", [1, 3, 1023, 4099, 1_000_003])
@pytest.mark.parametrize("off", [0, 1, 3])
def test_elementwise_kernels_stay_inside_their_outputs(eng, n, off):
    rng = np.random.default_rng(n + off)
    base = torch.from_numpy(rng.standard_normal(n + 4).astype(np.float32) ** 2).to(eng.device)
    s = base[off:off + n]                                                   # input misaligned by `off` floats
    g = Guarded(n + off, torch.float32, eng.device)
    eng.db(s, 1e-9, out=g.view[off:])
    g.check("b200vec_db")
    assert bool((g.raw[PAD * 4:(PAD + off) * 4] == SENT).all())             # and nothing before a misaligned output
    xc = torch.from_numpy((rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)).to(eng.device)
    g = Guarded(2 * n, torch.int16, eng.device)
    eng.wv_quantize(xc * 0.2, out=g.view)
    g.check("b200vec_wv_quantize")                                          # (0x5A5A is a legitimate int16: no coverage check)
    pitch = (n + 3) // 4 * 4
    g = Guarded(2 * max(pitch, 4), torch.float32, eng.device)
    planes = eng.split_planes(xc, out=g.view.view(2, max(pitch, 4)))
    g.check("b200vec_split_planes")
    assert torch.equal(planes[0], xc.real) and torch.equal(planes[1], xc.imag)
    if pitch > n:                                                           # the pad of row 0 is not output either
        torch.cuda.synchronize()
        assert bool((g.view.view(2, max(pitch, 4))[0, n:].contiguous().view(torch.uint8) == SENT).all())
    for stride in (2, 7, 56):
        n_out = (n + stride - 1) // stride
        g = Guarded(n_out, torch.complex64, eng.device)
        check_rc = eng.lib.b200vec_decimate(eng._ctx, xc.data_ptr(), n, stride, g.view.data_ptr(), eng._stream())
        assert check_rc == 0
        g.check("b200vec_decimate"); assert g.written()
        assert torch.equal(g.view, xc[::stride])
    win = 56
    g = Guarded(max(n, 1), torch.float64, eng.device)
    mx = torch.zeros(1, dtype=torch.float64, device=eng.device)
    if n >= win:
        assert eng.lib.b200vec_box_energy(eng._ctx, xc.data_ptr(), n, win, g.view.data_ptr(), mx.data_ptr(), eng._stream()) == 0
        g.check("b200vec_box_energy"); assert g.written()


@pytest.mark.parametrize("rows,cols", [(1, 1), (2, 5), (1024, 33), (1024, 999), (37, 4001)])
def test_median_filter_and_replicate_stay_inside(eng, rows, cols):
    rng = np.random.default_rng(rows * cols)
    s = torch.from_numpy(rng.standard_normal((rows, cols)).astype(np.float32)).to(eng.device)
    g = Guarded(rows * cols, torch.float32, eng.device)
    out = g.view.view(rows, cols)
    assert eng.lib.b200vec_median2_rows(eng._ctx, s.data_ptr(), rows, cols, s.stride(0), s.stride(1), out.data_ptr(),
                                        out.stride(0), out.stride(1), eng._stream()) == 0
    g.check("b200vec_median2_rows"); assert g.written()
    want = s.clone(); want[1:] = torch.maximum(s[1:], s[:-1])
    assert torch.equal(out, want)
    if rows == 1024:                                                        # frame-major spectrogram [1024, cols]: tile a period
        g = Guarded(cols * 1024, torch.float32, eng.device)
        spec = g.view.view(cols, 1024).t()
        per = max(1, cols // 5)
        spec[:, :per] = s[:, :per]
        spec[:, per:] = 0
        eng.replicate_frames(spec, 0, per, per, cols - per)
        g.check("b200vec_replicate_frames")
        idx = torch.arange(cols, device=eng.device) % per
        assert torch.equal(spec, s[:, :per][:, idx])


def test_normalize_and_batch_stay_inside(eng):
    rng = np.random.default_rng(77)
    n = 1_000_003
    s = torch.from_numpy(rng.standard_normal(n).astype(np.float32) ** 2).to(eng.device)
    db, vmin, vmax = eng.normalize_spectrogram(s, 5, 95, 25)
    assert db.numel() == n and bool(torch.isfinite(db).all())
    from vector_b200 import generate, synth
    fs, total = 56e6, 123_457
    cfgs, pk = synth.scenario(synth.CFG2_ROWS[:3], fs)
    info = generate.prepare_bank(eng, cfgs, pk, fs)
    rows = generate.placement_rows(cfgs, info, total, fs)
    scen = [[(st, per, cnt, slot, 1e6 * (v + 1)) for (st, per, cnt, slot) in rows] for v in range(5)]
    eng.batch_prepare(scen, fs)
    ld = (total + 1) // 2 * 2 + 6                                           # rows longer than the vectors
    g = Guarded(5 * ld, torch.complex64, eng.device)
    out, _ = eng.batch_build(total, normalize=True, out=g.view.view(5, ld))
    g.check("b200vec_batch_build")
    torch.cuda.synchronize()
    assert bool((g.view.view(5, ld)[:, total:].contiguous().view(torch.uint8) == SENT).all()), "batch build wrote past a vector's end"
    assert bool(torch.isfinite(torch.view_as_real(out)).all())

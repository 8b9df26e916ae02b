"""Reference-named helpers: same names, signatures, defaults, return types and exceptions as
ramiyako/vector's ``utils.py`` functions on the vector-build path, with the arithmetic done
by the CUDA engine.  numpy in, numpy out — device tensors never leak through these names
(``*_device`` variants return torch CUDA tensors for callers that stay on the GPU).

Citations are into the reference.
"""
from __future__ import annotations

import os
import struct
import time
from datetime import datetime

import numpy as np
import scipy.io as sio
import torch

from . import hostmath as H
from .engine import get_engine

# scratch bank slots used by the single-array helpers (the vector build uses slots 0..31)
_SLOT_SCRATCH = 63


# ------------------------------------------------------------------------------ file formats
_MI_DTYPES = {1: "<i1", 2: "<u1", 3: "<i2", 4: "<u2", 5: "<i4", 6: "<u4", 7: "<f4", 9: "<f8", 12: "<i8", 13: "<u8"}
_MX_NUMERIC = {6: np.float64, 7: np.float32, 8: np.int8, 9: np.uint8, 10: np.int16, 11: np.uint16,
               12: np.int32, 13: np.uint32, 14: np.int64, 15: np.uint64}


def _mat5_element(buf, pos):
    """(type, payload offset, payload bytes, offset of the next element) of the data element at pos."""
    word = struct.unpack_from("<I", buf, pos)[0]
    if word >> 16:                                         # small data element: 2-byte type, 2-byte size, 4 data bytes
        return word & 0xFFFF, pos + 4, word >> 16, pos + 8
    nbytes = struct.unpack_from("<I", buf, pos + 4)[0]
    return word, pos + 8, nbytes, pos + 8 + (nbytes + 7) // 8 * 8


def _read_mat5_numeric(path):
    """Fast path of the MAT v5 reader for what the vector-build path loads (utils.py:89-105): a file of
    uncompressed little-endian numeric arrays (real or complex; `Y`, `pre_samples`, ...).  Returns
    {name: (real ndarray view, imag ndarray view or None, dims)} without copying sample data, or None when
    the file holds anything else (compressed elements, cells, structs, chars, sparse, big-endian) — the
    caller then goes through scipy.io.loadmat like the reference."""
    try:
        buf = np.memmap(path, dtype=np.uint8, mode="r")
    except (OSError, ValueError):
        return None
    if buf.size < 136 or bytes(buf[126:128]) != b"IM" or bytes(buf[:4]) == b"\0\0\0\0":
        return None
    out, pos, end = {}, 128, buf.size
    try:
        while pos + 8 <= end:
            mtype, off, nbytes, nxt = _mat5_element(buf, pos)
            if mtype != 14 or nxt > end:                   # miMATRIX only (15 = zlib-compressed)
                return None
            p, stop = off, off + nbytes
            t, o, nb, p = _mat5_element(buf, p)            # array flags
            if t != 6 or nb != 8:
                return None
            flags = struct.unpack_from("<I", buf, o)[0]
            klass, is_complex = flags & 0xFF, bool(flags & 0x0800)
            if klass not in _MX_NUMERIC or flags & 0x0600:  # numeric classes only; no logical / global
                return None
            t, o, nb, p = _mat5_element(buf, p)            # dimensions
            if t != 5:
                return None
            dims = struct.unpack_from("<%di" % (nb // 4), buf, o)
            t, o, nb, p = _mat5_element(buf, p)            # name
            if t != 1:
                return None
            name = bytes(buf[o:o + nb]).decode("latin1")
            count = int(np.prod(dims)) if dims else 0
            parts = []
            for _ in range(2 if is_complex else 1):
                t, o, nb, p = _mat5_element(buf, p)
                if t not in _MI_DTYPES:
                    return None
                dt = np.dtype(_MI_DTYPES[t])
                if nb != count * dt.itemsize:
                    return None
                parts.append(np.frombuffer(buf, dtype=dt, count=count, offset=o) if count else np.zeros(0, dt))
            if p > stop + 8:
                return None
            out[name] = (parts[0], parts[1] if is_complex else None, dims, _MX_NUMERIC[klass])
            pos = nxt
    except (struct.error, ValueError, IndexError):
        return None
    return out if pos == end else None


def _load_packet_fast(file_path):
    """load_packet_info through the fast reader; None when the file needs scipy's general reader."""
    arrays = _read_mat5_numeric(file_path)
    if arrays is None:
        return None
    if "Y" in arrays:
        key = "Y"
    elif len(arrays) == 1:
        key = next(iter(arrays))
    else:
        return None                                        # ambiguous: let the reference path raise its message
    re, im, dims, _ = arrays[key]
    if sum(1 for d in dims if d != 1) > 1:
        return None                                        # a true matrix: its flatten order needs the general path
    y = np.empty(re.shape[0], dtype=np.complex64)
    y.real = re                                            # per-component cast to float32 == astype(complex64)
    y.imag = im if im is not None else 0
    pre = 0
    if "pre_samples" in arrays:
        pr, pi, pdims, _ = arrays["pre_samples"]
        if pr.shape[0] != 1 or pi is not None:
            return None
        pre = int(pr[0])
    return y, pre


def load_packet_info(file_path):
    """(complex64 samples, pre_samples) — utils.py:89-105.  Key 'Y', else the single
    non-dunder key, else ValueError; flattened; pre_samples defaults to 0.  Plain numeric MAT v5 files
    (what save_vector and the packet extractor write) are parsed directly from a memory map — one pass
    from the file's planes to complex64; anything else goes through scipy.io.loadmat."""
    fast = _load_packet_fast(file_path)
    if fast is not None:
        return fast
    mat = sio.loadmat(file_path, squeeze_me=True, struct_as_record=False)
    if "Y" in mat:
        data = mat["Y"]
    else:
        keys = [k for k in mat.keys() if not k.startswith("__")]
        if len(keys) == 1:
            data = mat[keys[0]]
        else:
            raise ValueError(f"Ambiguous packet data in {file_path}. Available keys: {list(mat.keys())}")
    if data.ndim > 1:
        data = data.flatten()
    return data.astype(np.complex64), int(mat.get("pre_samples", 0))


def load_packet(file_path):
    """complex64 samples only — utils.py:48-87 (same key rules; errors are printed and
    re-raised like the reference does)."""
    try:
        size_mb = os.path.getsize(file_path) / (1024 * 1024)
        if size_mb > 50:
            print(f"📁 Loading large file: {size_mb:.1f}MB")
        pkt, _ = load_packet_info(file_path)
        if len(pkt) > 20_000_000:
            print(f"⚠️ Heavy packet loaded: {len(pkt):,} samples ({len(pkt)/56e6:.2f}s, "
                  f"{pkt.nbytes/(1024*1024):.1f}MB)")
        return pkt
    except Exception as e:
        print(f"Error loading packet from {file_path}: {e}")
        raise


def _mat5_tag(mtype, nbytes):
    return struct.pack("<II", mtype, nbytes)


def _mat5_y_header(n):
    """Everything of the miMATRIX element 'Y' (complex single, 1 x n) up to the real-part data, as
    scipy.io.savemat lays it out (oned_as='row', no compression) for n >= 2."""
    plane = 8 + (4 * n + 7) // 8 * 8
    body = (_mat5_tag(6, 8) + struct.pack("<II", 7 | 0x0800, 0)          # array flags: mxSINGLE_CLASS, complex
            + _mat5_tag(5, 8) + struct.pack("<ii", 1, n)                  # dimensions
            + struct.pack("<HH4s", 1, 1, b"Y\0\0\0"))                    # name, small data element
    return _mat5_tag(14, len(body) + 2 * plane) + body


_MAT5_PRE_SAMPLES = (_mat5_tag(14, 0x48) + _mat5_tag(6, 8) + struct.pack("<II", 14, 0)   # mxINT64_CLASS
                     + _mat5_tag(5, 8) + struct.pack("<ii", 1, 1)
                     + _mat5_tag(1, 11) + b"pre_samples".ljust(16, b"\0")
                     + _mat5_tag(12, 8) + struct.pack("<q", 0))


def _mat5_file_header():
    text = ("MATLAB 5.0 MAT-file Platform: %s, Created on: %s" % (os.name, time.asctime())).encode("latin1")
    return text[:116].ljust(116, b"\0") + b"\0" * 8 + struct.pack("<H", 0x0100) + b"IM"


def _write_mat5_vector(output_path, re_plane, im_plane, n):
    """MAT v5 file {'Y': complex single 1 x n, 'pre_samples': int64 0} written straight from the two
    float32 planes — byte-identical to scipy.io.savemat's output after the 128-byte text header.
    A plane is a host array or a CUDA tensor (then it is streamed, _stream_to_file)."""
    pad = b"\0" * ((-4 * n) % 8)
    with open(output_path, "wb", buffering=0) as fh:
        fh.write(_mat5_file_header() + _mat5_y_header(n))
        for plane in (re_plane, im_plane):
            fh.write(_mat5_tag(7, 4 * n))
            if isinstance(plane, torch.Tensor):
                _stream_to_file(fh, plane)
            else:
                fh.write(memoryview(plane).cast("B"))
            fh.write(pad)
        fh.write(_MAT5_PRE_SAMPLES)


_RING = {}
_CHUNK_BYTES = 32 << 20


def _ring(device):
    """Two pinned staging blocks + two events per device, allocated once."""
    key = device.index
    if key not in _RING:
        _RING[key] = ([torch.empty(_CHUNK_BYTES, dtype=torch.uint8).pin_memory() for _ in range(2)],
                      [torch.cuda.Event() for _ in range(2)])
    return _RING[key]


def _stream_to_file(fh, dev):
    """Write a contiguous CUDA tensor's bytes to the open file: chunked device-to-host copies into a
    two-block pinned ring, the copy of chunk k+1 in flight (DMA engine) while chunk k is being written —
    the save path's D2H costs no wall time beyond the file write itself."""
    flat = dev.reshape(-1).view(torch.uint8)
    total = flat.numel()
    if total == 0:
        return
    bufs, evs = _ring(dev.device)
    stream = torch.cuda.current_stream(dev.device)
    n_chunks = (total + _CHUNK_BYTES - 1) // _CHUNK_BYTES

    def issue(k):
        lo = k * _CHUNK_BYTES
        hi = min(total, lo + _CHUNK_BYTES)
        bufs[k & 1][: hi - lo].copy_(flat[lo:hi], non_blocking=True)
        evs[k & 1].record(stream)
        return hi - lo

    sizes = {0: issue(0)}
    for k in range(n_chunks):
        if k + 1 < n_chunks:
            sizes[k + 1] = issue(k + 1)                   # block (k+1)&1 was written out one iteration ago
        evs[k & 1].synchronize()
        fh.write(memoryview(bufs[k & 1].numpy())[: sizes.pop(k)])


def save_vector(vector, output_path):
    """MAT v5 {'Y': complex64 1-D, 'pre_samples': 0} — utils.py:659-670.  `vector` is a numpy array
    (the reference's argument) or a complex64 CUDA tensor.  The file is written directly in
    savemat's layout (two float32 planes) instead of through scipy's generic writer; a CUDA tensor is
    split into its planes on the device (K7) and streamed through a pinned ring while the file is being
    written.  Vectors shorter than 2 samples or too long for one MAT v5 element (>= 2^29 samples) go
    through scipy.io.savemat, which then produces the very output / exception the reference produces."""
    if isinstance(vector, torch.Tensor) and vector.is_cuda:
        dev = vector.reshape(-1)
        if dev.dtype != torch.complex64:
            dev = dev.to(torch.complex64)
        n = dev.numel()
        if 2 <= n < (1 << 29):
            eng = get_engine(dev.device)
            planes = eng.split_planes(dev)
            _write_mat5_vector(output_path, planes[0], planes[1], n)
            return
        vector = dev.cpu().numpy()
    vector = np.asarray(vector)
    if vector.ndim > 1:
        vector = vector.flatten()
    vector = vector.astype(np.complex64)
    n = vector.shape[0]
    if 2 <= n < (1 << 29):
        _write_mat5_vector(output_path, np.ascontiguousarray(vector.real), np.ascontiguousarray(vector.imag), n)
    else:
        sio.savemat(output_path, {"Y": vector, "pre_samples": 0})


def wv_header(n_samples, sample_rate, rms_dbfs=0.0, peak_dbfs=0.0, date=None):
    """Tag block of an R&S .wv file as mat2wv writes it (mat_to_wv_converter.py:56-63)."""
    date = date or datetime.now().strftime("%Y-%m-%d;%H:%M:%S")
    return (f"{{TYPE: SMU-WV,0}}{{COMMENT: Generated by mat2wv.py}}{{DATE: {date}}}"
            f"{{LEVEL OFFS: {rms_dbfs}, {peak_dbfs}}}{{CLOCK: {sample_rate}}}"
            f"{{SAMPLES: {n_samples}}}{{WAVEFORM-{4 * n_samples + 1}:#").encode()


def wv_payload_device(vector_dev, engine=None):
    """int16 I/Q interleave of a complex64 CUDA tensor (K6; mat_to_wv_converter.py:42-50)."""
    eng = engine or get_engine(vector_dev.device)
    return eng.wv_quantize(vector_dev)


def wv_levels_device(dev, engine=None):
    """mat2wv's bNormalize branch (mat_to_wv_converter.py:30-36) for a complex64 CUDA tensor:
    (pre_scale, rms_dbfs, peak_dbfs) with pre_scale = fl32(1 / max |x|) — numpy divides a complex64 array
    by a float32 scalar by multiplying with its float32 reciprocal — and the two header values as the
    float32 scalars the reference formats.  max |x| is the correctly rounded float32 of the exact maximum
    (numpy's SIMD hypot may differ in the last bit); the mean power is accumulated in fp64 (the reference's
    float32 pairwise mean differs from it by ~1e-7 relative, < 1e-6 dB in the header)."""
    eng = engine or get_engine(dev.device)
    n = dev.numel()
    mx2, _ = eng.power_stats(dev, 1.0)
    m = np.float32(np.sqrt(mx2))
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.float32(1.0) / m
        pk2, sm = eng.power_stats(dev, float(r))
        a = np.float32(np.sqrt(pk2))
        peak = np.float32(-10) * np.log10(a * a)
        rms = np.float32(-10) * np.log10(np.float32(sm / n))
    return float(r), rms, peak


def save_vector_wv(vector, output_path, sample_rate, normalize=False, _date=None):
    """R&S .wv writer — utils.py:672-677 -> mat2wv(vector, path, fs, bNormalize=normalize)
    (vector_analyzer/mat_to_wv_converter.py:7-65).  `vector` may be a numpy array or a
    complex64 CUDA tensor: quantisation (and, with normalize=True, max |x| and the level metrics) run on
    the device, and only the 4 B/sample int16 payload crosses PCIe, streamed through a pinned ring while
    the file is being written."""
    eng = get_engine(vector.device) if isinstance(vector, torch.Tensor) and vector.is_cuda else get_engine()
    if isinstance(vector, torch.Tensor):
        dev = vector.reshape(-1).to(device=eng.device, dtype=torch.complex64)
    else:
        host = np.asarray(vector).flatten().astype(np.complex64)
        dev = torch.from_numpy(host).to(eng.device)
    n = dev.numel()
    pre_scale, rms, peak = 1.0, 0.0, 0.0
    if normalize:
        if n == 0:
            raise ValueError("zero-size array to reduction operation maximum which has no identity")   # np.max([])
        print("Normalize signal")
        pre_scale, rms, peak = wv_levels_device(dev, eng)
    payload = eng.wv_quantize(dev, pre_scale=pre_scale) if n else None
    with open(output_path, "wb", buffering=0) as fh:
        fh.write(wv_header(n, sample_rate, rms, peak, date=_date))
        if n:
            _stream_to_file(fh, payload)
        fh.write(b"}")


# ------------------------------------------------------------------------------ DSP helpers
def resample_signal(signal, orig_sr, target_sr):
    """utils.py:107-118: unchanged object when the rates match, else scipy.signal.resample's
    Fourier method to int(len*ratio) samples, complex64 out (K0)."""
    if orig_sr == target_sr:
        return signal
    ratio = target_sr / orig_sr
    new_length = int(len(signal) * ratio)
    if new_length <= 0:
        raise ValueError("resample: new length must be positive")
    eng = get_engine()
    eng.bank_upload(_SLOT_SCRATCH, np.asarray(signal).astype(np.complex64))
    eng.bank_resample(_SLOT_SCRATCH, new_length)
    eng.bank_commit(_SLOT_SCRATCH, 0.0, target_sr)
    return eng.bank_download(_SLOT_SCRATCH)


def apply_frequency_shift(signal, freq_shift, sample_rate):
    """utils.py:120-127: same object back when freq_shift == 0, else
    complex64(signal * exp(2j*pi*f*arange(L)/fs)) with the phase formed in fp64 (K1)."""
    if freq_shift == 0:
        return signal
    eng = get_engine()
    eng.bank_upload(_SLOT_SCRATCH, np.asarray(signal).astype(np.complex64))
    eng.bank_commit(_SLOT_SCRATCH, float(freq_shift), float(sample_rate))
    return eng.bank_download(_SLOT_SCRATCH)


compute_freq_ranges = H.compute_freq_ranges


# ------------------------------------------------------------------------------ spectrogram
def _run_plan(eng, x_dev, plan, flag, stride):
    plan = H.resolve_plan(plan)                      # may raise ValueError like scipy would
    flag.zero_()
    if x_dev.dtype == torch.complex128 and plan.nfft <= 8192:        # scipy's float64 path, in float64
        S = eng.stft_f64(x_dev, plan.window, plan.nperseg, plan.noverlap, plan.nfft, plan.fs,
                         scaling=plan.scaling, stride=stride, flag=flag)
        return plan, S
    dbl = x_dev.dtype == torch.complex128                            # nfft 16384 does not fit in float64: float32, widened
    if dbl:
        x_dev, stride = x_dev[::stride].to(torch.complex64), 1
    S = eng.stft(x_dev, plan.window, plan.nperseg, plan.noverlap, plan.nfft, plan.fs,
                 scaling=plan.scaling, stride=stride, flag=flag)
    return plan, (S.to(torch.float64) if dbl else S)


def create_spectrogram_device(sig_dev, sr, center_freq=0, max_samples=2_000_000,
                              time_resolution_us=1, adaptive_resolution=True, engine=None):
    """create_spectrogram on a complex64 CUDA tensor; returns (freqs f64, times f64, Sxx CUDA
    float32 [nfft, frames]).  A complex128 tensor takes scipy's double-precision path (float64 window, FFT
    and PSD: b200vec_stft_f64) and returns a float64 Sxx.  Follows utils.py:161-353 including the fallback chain.  A decimated
    preview (heavy signals, utils.py:192-195) gathers sig[::factor] ONCE into a compact buffer
    (b200vec_decimate) that the overlapping frames then read — the same samples and arithmetic as
    transforming the strided view."""
    n = sig_dev.numel()
    if n == 0:
        raise ValueError("Signal is empty")
    eng = engine or get_engine(sig_dev.device)
    plan0 = H.plan_spectrogram(n, sr, max_samples, time_resolution_us, adaptive_resolution)
    if plan0.heavy:
        print(f"🔍 Heavy packet detected: {n:,} samples ({n/sr:.2f}s)")
        print(f"📉 Downsampled by factor {plan0.factor}: {plan0.n_dec:,} samples")
    x, stride = sig_dev, plan0.factor
    if stride > 1 and plan0.nperseg >= 2 * plan0.hop and sig_dev.dtype == torch.complex64:
        x, stride = eng.decimate(sig_dev, stride), 1             # frames overlap: read the strided samples once
    flag = eng._flag
    fb = H.fallback_plans(plan0)
    try:
        plan, S = _run_plan(eng, x, plan0, flag, stride)
    except Exception:                                                    # utils.py:298-313 catches everything
        plan, S = _run_plan(eng, x, fb["on_error"], flag, stride)
    if int(flag.item()) == 0:                                            # utils.py:316
        print("Warning: Spectrogram is all zeros, trying with reduced window size...")
        try:
            plan, S = _run_plan(eng, x, fb["on_zero"], flag, stride)     # :318-334
        except Exception:
            plan, S = _run_plan(eng, x, fb["last_resort"], flag, stride)  # :335-347
    freqs = np.fft.fftshift(plan.freqs()) * plan.factor + center_freq    # :350
    return freqs, plan.times(), S


def create_spectrogram(sig, sr, center_freq=0, max_samples=2_000_000, time_resolution_us=1,
                       adaptive_resolution=True):
    """Drop-in for utils.create_spectrogram (utils.py:161-353): numpy in, (freqs, times, Sxx)
    numpy out, Sxx C-contiguous [nfft, frames] with fftshifted rows.  The arithmetic is that of
    scipy for the input's precision: float32 for complex64 / float32 data (the reference's own case),
    float64 for complex128 / float64 data (b200vec_stft_f64), with a float64 Sxx as scipy returns."""
    if len(sig) == 0:
        raise ValueError("Signal is empty")
    a = np.asarray(sig)
    eng = get_engine()
    wide = np.result_type(a.dtype, np.complex64) == np.complex128           # scipy's outdtype rule
    dev = torch.from_numpy(np.ascontiguousarray(a.astype(np.complex128 if wide else np.complex64, copy=False))).to(eng.device)
    f, t, S = create_spectrogram_device(dev, sr, center_freq, max_samples, time_resolution_us,
                                        adaptive_resolution, engine=eng)
    return f, t, S.cpu().numpy()


def full_rate_spectrogram_device(sig_dev, fs, center_freq=0.0, engine=None, **kw):
    """The fixed analysis STFT of vector_analyzer/spectrogram_analysis.py:32-53 (BASELINE
    cfg3): hann, nperseg 1024, noverlap 512, nfft 1024, two-sided, no detrend, 'density'."""
    eng = engine or get_engine(sig_dev.device)
    n = sig_dev.numel()
    plan = H.SpectrogramPlan(n, 1, n, fs, False, "hann", 1024, 512, 1024, "density")
    plan = H.resolve_plan(plan)
    S = eng.stft(sig_dev, "hann", plan.nperseg, 512, 1024, fs, scaling="density", **kw)
    return np.fft.fftshift(plan.freqs()) + center_freq, plan.times(), S


def normalize_spectrogram(Sxx, low_percentile=10.0, high_percentile=95.0, max_dynamic_range=25):
    """Drop-in for utils.normalize_spectrogram (utils.py:356-404): (Sxx_db, vmin, vmax)."""
    Sxx = np.asarray(Sxx)
    if Sxx.size == 0:
        return np.array([]), 0, 0
    eng = get_engine()
    dev = torch.from_numpy(np.ascontiguousarray(np.abs(Sxx).astype(np.float32, copy=False))).to(eng.device)
    db, vmin, vmax = eng.normalize_spectrogram(dev, low_percentile, high_percentile, max_dynamic_range)
    out = db.cpu().numpy()
    if Sxx.dtype == np.float64:
        out, vmin, vmax = out.astype(np.float64), np.float64(vmin), np.float64(vmax)
    return out, vmin, vmax


# ------------------------------------------------------------------------------ either side of the path
def detect_packet_bounds(signal, sample_rate, threshold_ratio=0.2):
    """Drop-in for utils.detect_packet_bounds (utils.py:811-825): (start, end) sample indices of the
    packet — |x|^2 box-smoothed over 1 us (np.convolve 'same' in float64), threshold at noise +
    ratio * (peak - noise) with noise = median of the first tenth, first / last sample above it;
    (0, len(signal)) when nothing is above.  `signal`: numpy array or complex64 CUDA tensor.
    Smoothing, peak and the index search run on the device (K8/K9); the median is an exact order
    statistic of the device array (radix select on 64-bit keys, one launch), the two scalars are combined
    on the host in double precision exactly like the reference's numpy scalars."""
    eng = get_engine(signal.device) if isinstance(signal, torch.Tensor) and signal.is_cuda else get_engine()
    if isinstance(signal, torch.Tensor):
        dev = signal.reshape(-1).to(device=eng.device, dtype=torch.complex64)
    else:
        dev = torch.from_numpy(np.ascontiguousarray(np.asarray(signal).reshape(-1).astype(np.complex64))).to(eng.device)
    n = dev.numel()
    window = max(1, int(sample_rate // 1_000_000))
    if n == 0:
        raise ValueError("v cannot be empty")                  # np.convolve on an empty signal, utils.py:816
    if window > 4096:
        raise ValueError(f"detect_packet_bounds: smoothing window of {window} samples (sample rate above 4 GHz)")
    if n < window:
        # np.convolve(energy, kernel, 'same') returns max(n, window) samples centred on the LONGER operand:
        # the same numbers as smoothing the signal zero-padded to `window` samples from this offset
        pad = torch.zeros(window, dtype=torch.complex64, device=eng.device)
        off = (window - 1) // 2 - (n - 1) // 2
        pad[off:off + n] = dev
        dev = pad
    smoothed, max_en = eng.box_energy(dev, window)
    head = smoothed[: max(1, smoothed.numel() // 10)]         # len(smoothed) = max(n, window)
    m = head.numel()
    if m % 2:                                                 # np.median: exact order statistics (b200vec_select_f64)
        noise = eng.select_f64(head, [m // 2])[0]
    else:
        lo_v, hi_v = eng.select_f64(head, [m // 2 - 1, m // 2])
        noise = float((np.float64(lo_v) + np.float64(hi_v)) / 2)
    max_en = float(max_en)
    threshold = noise + threshold_ratio * (max_en - noise)
    first, last = eng.threshold_bounds(smoothed, threshold)
    if last < 0:
        return 0, n                                     # (0, len(signal)), utils.py:822-823
    return first, last


def median_filter_rows2(Sxx_db, engine=None):
    """scipy.ndimage.median_filter(Sxx_db, size=(2, 1)) as plot_spectrogram applies it before display
    (utils.py:435-440): each bin takes the larger of itself and the bin one row up (row 0 unchanged).
    numpy in -> numpy out; CUDA tensor in -> CUDA tensor out (same layout)."""
    if isinstance(Sxx_db, torch.Tensor) and Sxx_db.is_cuda:
        eng = engine or get_engine(Sxx_db.device)
        return eng.median2_rows(Sxx_db)
    a = np.asarray(Sxx_db)
    if a.ndim != 2:
        raise ValueError("median_filter_rows2 expects a 2-D array")
    if a.size == 0:
        return a.copy()
    eng = engine or get_engine()
    dev = torch.from_numpy(np.ascontiguousarray(a.astype(np.float32, copy=False))).to(eng.device)
    return eng.median2_rows(dev).cpu().numpy().astype(a.dtype, copy=False)

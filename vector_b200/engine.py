"""Device engine: torch owns device memory and streams, libb200vec.so (hand-written sm_100a
kernels behind the C-ABI of include/b200vec.h) does the arithmetic.

One Engine per GPU (one process per GPU under torch.distributed).  Every method takes and
returns torch CUDA tensors; the numpy-facing, reference-named API lives in api.py.  There is
no CPU path: constructing an Engine without a CUDA device or without the built library
raises.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib, hostmath as H
from ._lib import BUILD_MAXONLY, BUILD_SCALE, BUILD_STORE, PktDesc, check


class Engine:
    def __init__(self, device=None):
        if not torch.cuda.is_available():
            raise _lib.B200VecError(
                "vector_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.lib = _lib.load()
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None
                                   else (device.index if isinstance(device, torch.device) else int(device)))
        ctx = C.c_void_p()
        check(self.lib.b200vec_create(self.device.index, C.byref(ctx)))
        self._ctx = ctx
        self._windows = {}
        self._maxsq = torch.zeros(1, dtype=torch.float64, device=self.device)
        self._flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        self.launches = 0          # kernels launched through this engine (bench.py's gpu_launches)

    def close(self):
        if getattr(self, "_ctx", None):
            self.lib.b200vec_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ helpers
    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    @staticmethod
    def _ptr(t):
        return C.c_void_p(t.data_ptr())

    # ------------------------------------------------------------------ packet bank
    def bank_upload(self, slot, samples):
        """samples: numpy complex64 [L] (host) or torch complex64 CUDA tensor."""
        if isinstance(samples, torch.Tensor):
            # device-to-device copy in stream order; torch's allocator frees `t` in stream order too
            t = samples.to(self.device, torch.complex64).contiguous()
            check(self.lib.b200vec_bank_upload(self._ctx, slot, self._ptr(t), t.numel(), 1, self._stream()))
        else:
            # No synchronize: a copy from pageable memory returns once the source has been read, and an
            # array that needed no conversion is the caller's own (pinned) buffer, which the caller keeps
            # alive until the stream has passed — like any buffer handed to an asynchronous copy.
            a = np.ascontiguousarray(samples, dtype=np.complex64)
            check(self.lib.b200vec_bank_upload(self._ctx, slot, a.ctypes.data_as(C.c_void_p), a.size, 0,
                                               self._stream()))

    def bank_resample(self, slot, n_out):
        check(self.lib.b200vec_bank_resample(self._ctx, slot, int(n_out), self._stream()))

    def bank_commit(self, slot, freq_hz=0.0, fs_hz=H.TARGET_SAMPLE_RATE):
        check(self.lib.b200vec_bank_shift(self._ctx, slot, float(freq_hz), float(fs_hz), self._stream()))
        self.launches += 1

    def bank_attach(self, slot, samples):
        """Make a complex64 CUDA tensor a committed bank slot WITHOUT copying it (b200vec_bank_attach):
        the caller keeps `samples` alive and unchanged while launches that name the slot run."""
        assert samples.is_cuda and samples.dtype == torch.complex64 and samples.is_contiguous()
        check(self.lib.b200vec_bank_attach(self._ctx, slot, self._ptr(samples), samples.numel(), self._stream()))

    def bank_len(self, slot):
        n = C.c_int64()
        check(self.lib.b200vec_bank_len(self._ctx, slot, C.byref(n)))
        return n.value

    def bank_download(self, slot):
        n = self.bank_len(slot)
        out = np.empty(n, dtype=np.complex64)
        check(self.lib.b200vec_bank_download(self._ctx, slot, out.ctypes.data_as(C.c_void_p), n, self._stream()))
        return out

    # ------------------------------------------------------------------ vector build
    @staticmethod
    def make_descs(rows, free_flags=None):
        """rows: iterable of (start, period, count, slot) -> ctypes array of b200vec_pkt_desc.
        free_flags (hostmath.free_instance_flags): sets bit 0 of `reserved`, the hint that lets the
        in-kernel max pass skip spans inside a single instance."""
        rows = list(rows)
        arr = (PktDesc * max(1, len(rows)))()
        for i, (start, period, count, slot) in enumerate(rows):
            arr[i] = PktDesc(int(start), int(period), int(count), int(slot),
                             1 if (free_flags is not None and free_flags[i]) else 0)
        return arr, len(rows)

    def free_flags(self, rows, total):
        """hostmath.free_instance_flags for these rows, cached (the interval sweep is host work that
        would otherwise be repeated for every step of an unchanged scenario)."""
        rows = [tuple(int(v) for v in r) for r in rows]
        key = (tuple(rows), tuple(self.bank_len(r[3]) for r in rows), int(total))
        cache = self.__dict__.setdefault("_free_flag_cache", {})
        if key not in cache:
            if len(cache) > 64:
                cache.clear()
            cache[key] = H.free_instance_flags(rows, list(key[1]), total)
        return cache[key]

    def build_pass(self, descs, n_descs, lo, hi, out, mode, maxsq=None):
        maxsq = self._maxsq if maxsq is None else maxsq
        if hi <= lo:
            return
        check(self.lib.b200vec_build(self._ctx, descs, n_descs, int(lo), int(hi),
                                     self._ptr(out) if out is not None else None, mode,
                                     self._ptr(maxsq), self._stream()))
        self.launches += 1

    def max_ranges(self, rows, total, world=1, rank=0):
        """This rank's share of the sample ranges whose max|v| is the max over the whole vector
        (hostmath.max_eval_ranges: periodic trains repeat bit-exactly, so one common period plus
        the head and tail of the vector is enough)."""
        rows = list(rows)
        lengths = [self.bank_len(r[3]) for r in rows]
        return H.split_ranges(H.max_eval_ranges(rows, lengths, total), world, rank)

    def max_pass(self, descs, n_descs, ranges, maxsq=None):
        """Pass A: max |v|^2 over the given sample ranges into maxsq (atomic max; zero it first)."""
        for lo, hi in ranges:
            self.build_pass(descs, n_descs, lo, hi, None, BUILD_MAXONLY, maxsq=maxsq)

    def build_norm(self, descs, n_descs, lo, hi, out, ranges, maxsq=None):
        """The normalised build in ONE launch (b200vec_build_norm): max |v|^2 over `ranges`, grid barrier,
        scale pass.  maxsq receives max |v|^2."""
        maxsq = self._maxsq if maxsq is None else maxsq
        flat = [int(v) for r in ranges for v in r]
        arr = (C.c_int64 * max(len(flat), 1))(*flat)
        check(self.lib.b200vec_build_norm(self._ctx, descs, n_descs, int(lo), int(hi),
                                          self._ptr(out) if out is not None else None, arr, len(flat) // 2,
                                          self._ptr(maxsq), self._stream()))
        self.launches += 1

    def build_vector(self, rows, total, normalize=True, lo=0, hi=None, out=None, reduce_max=None,
                     world=1, rank=0):
        """Samples [lo, hi) of the composite vector as a complex64 CUDA tensor.

        normalize=True: max |v|^2 and the scaled stores are ONE launch (pass A over the sample ranges
        that hold the global max, grid barrier, pass B) unless `reduce_max(tensor)` is given — then pass
        A is its own launch, reduce_max all-reduces that one double across ranks, and pass B recomputes
        and stores v * (1/max).  Returns (vector, maxsq tensor)."""
        hi = total if hi is None else hi
        n = hi - lo
        if out is None:
            out = torch.empty(max(n, 0), dtype=torch.complex64, device=self.device)
        if not normalize:
            descs, nd = self.make_descs(rows)
            self.build_pass(descs, nd, lo, hi, out, BUILD_STORE)
            return out, None
        if reduce_max is None:
            if lo != 0 or hi != total:
                descs, nd = self.make_descs(rows)
                ranges = [(lo, hi)]                                          # a lone sub-range: its own max
            else:
                descs, nd = self.make_descs(rows, self.free_flags(rows, total))
                ranges = self.max_ranges(rows, total, 1, 0)
                if len(ranges) > 4:
                    ranges = [(0, total)]
            if n > 0:
                self.build_norm(descs, nd, lo, hi, out, ranges)
            else:
                self._maxsq.zero_()
            return out, self._maxsq
        descs, nd = self.make_descs(rows)
        self._maxsq.zero_()
        self.max_pass(descs, nd, self.max_ranges(rows, total, world, rank))
        reduce_max(self._maxsq)
        self.build_pass(descs, nd, lo, hi, out, BUILD_SCALE)
        return out, self._maxsq

    def build_stft1024(self, descs, n_descs, lo, hi, vec_out, mode, frame_lo, n_frames, spec_out,
                       window="hann", fs=H.TARGET_SAMPLE_RATE, scaling="density", maxsq=None, flag=None,
                       norm_ranges=None, peer=None):
        """Fused pass B + full-rate STFT (nperseg = nfft = 1024, hop 512): samples [lo, hi) of the
        vector into vec_out, frames [frame_lo, frame_lo+n_frames) into spec_out, one kernel, the
        vector never re-read.  lo must equal frame_lo*512.  spec_out is a float32 [1024, n_frames]
        tensor; its strides pick the kernel: frame-major memory (stride(0) == 1, what
        alloc_spectrogram(..., frame_major=True) returns) runs the warp-per-frame kernel
        (b200vec_build_stft1024_fm), row-major memory the 16-frame tile kernel.
        norm_ranges (frame-major only): [(lo, hi), ...] sample ranges whose max |v| is the max of the
        whole vector (max_ranges()) — the kernel then finds the max itself, scales by it and writes
        max |v|^2 to maxsq: the whole normalised step in ONE launch (`mode` is ignored).
        peer (with norm_ranges = this rank's share): a dist.PeerMax — the per-rank maxima are then
        MAX-combined inside the kernel over NVLink peer memory (no NCCL call in the step)."""
        maxsq = self._maxsq if maxsq is None else maxsq
        taps, scale = self.window(window, 1024, fs, scaling)
        frame_major = spec_out.dim() == 2 and spec_out.stride(0) == 1 and spec_out.stride(1) >= 1024
        ld = spec_out.stride(1) if frame_major else spec_out.stride(0)
        fl = self._ptr(flag) if flag is not None else None
        if norm_ranges is not None:
            if not frame_major:
                raise ValueError("in-kernel normalisation needs frame-major spectrogram memory")
            flat = [int(v) for r in norm_ranges for v in r]
            arr = (C.c_int64 * max(len(flat), 1))(*flat)
            if peer is not None:
                blocks = (C.c_void_p * peer.world)(*peer.ptrs)
                check(self.lib.b200vec_build_stft1024_fm_norm_peer(
                    self._ctx, descs, n_descs, int(lo), int(hi), self._ptr(vec_out), arr, len(flat) // 2,
                    blocks, peer.world, peer.rank, peer.next_epoch(),
                    self._ptr(maxsq), self._ptr(taps), scale, int(frame_lo), int(n_frames), self._ptr(spec_out),
                    ld, fl, self._stream()))
            else:
                check(self.lib.b200vec_build_stft1024_fm_norm(
                    self._ctx, descs, n_descs, int(lo), int(hi), self._ptr(vec_out), arr, len(flat) // 2,
                    self._ptr(maxsq), self._ptr(taps), scale, int(frame_lo), int(n_frames), self._ptr(spec_out),
                    ld, fl, self._stream()))
        else:
            fn = self.lib.b200vec_build_stft1024_fm if frame_major else self.lib.b200vec_build_stft1024
            check(fn(self._ctx, descs, n_descs, int(lo), int(hi), self._ptr(vec_out), mode, self._ptr(maxsq),
                     self._ptr(taps), scale, int(frame_lo), int(n_frames), self._ptr(spec_out), ld, fl,
                     self._stream()))
        self.launches += 1

    # ------------------------------------------------------------------ spectrogram
    def window(self, name, nperseg, fs, scaling):
        key = (name, nperseg, float(fs), scaling)
        if key not in self._windows:
            taps, scale = H.window_and_scale(name, nperseg, fs, scaling)
            self._windows[key] = (torch.from_numpy(taps.astype(np.float32)).to(self.device), float(scale))
        return self._windows[key]

    def alloc_spectrogram(self, nfft, n_frames, frame_major=False):
        """float32 [nfft, n_frames] tensor for a spectrogram.
        frame_major=False: C-order rows whose starts sit on 128-byte boundaries (row pitch rounded
        up to 32 floats), so the kernels' per-row runs never straddle sectors.
        frame_major=True: the transpose view of a contiguous [n_frames, nfft] buffer — same shape,
        dtype and values for every consumer, one frame's bins contiguous in memory."""
        if frame_major:
            return torch.empty((max(n_frames, 0), nfft), dtype=torch.float32, device=self.device).t()
        pitch = (n_frames + 31) // 32 * 32
        return torch.empty((nfft, max(pitch, 32)), dtype=torch.float32, device=self.device)[:, :n_frames]

    def stft(self, x, window, nperseg, noverlap, nfft, fs, scaling="density", stride=1,
             in_origin=0, frame_lo=0, n_frames=None, total_len=None, out=None, flag=None, frame_major=None):
        """|STFT|^2*scale of x[::stride] -> float32 [nfft, n_frames], rows fftshifted.
        x is a complex64 CUDA tensor whose element 0 is vector sample `in_origin`.
        The full-rate analysis shape (nperseg = nfft = 1024, hop 512, stride 1) defaults to
        frame-major memory (the result is the transpose view of a [n_frames, 1024] buffer) and the
        warp-per-frame kernel; pass frame_major=False (or a C-order `out`) for C-order rows."""
        hop = nperseg - noverlap
        total_len = (in_origin + x.numel()) if total_len is None else total_len
        n_dec = (total_len + stride - 1) // stride
        if n_frames is None:
            n_frames = 0 if n_dec < nperseg else (n_dec - nperseg) // hop + 1 - frame_lo
        fm_shape = nperseg == 1024 and nfft == 1024 and hop == 512 and stride == 1
        if out is None:
            fm = fm_shape if frame_major is None else (frame_major and fm_shape)
            out = self.alloc_spectrogram(nfft, max(n_frames, 0), frame_major=fm)
        taps, scale = self.window(window, nperseg, fs, scaling)
        fl = self._ptr(flag) if flag is not None else None
        if fm_shape and out.dim() == 2 and out.stride(0) == 1 and out.stride(1) >= 1024 and n_frames > 0:
            check(self.lib.b200vec_stft1024_fm(self._ctx, self._ptr(x), int(in_origin), x.numel(), self._ptr(taps),
                                               scale, int(frame_lo), int(n_frames), self._ptr(out), out.stride(1),
                                               fl, self._stream()))
        else:
            if n_frames > 1 and out.stride(1) != 1:
                raise ValueError("this STFT shape needs C-order spectrogram memory")
            check(self.lib.b200vec_stft(self._ctx, self._ptr(x), int(in_origin), x.numel(), int(stride),
                                        int(nperseg), int(hop), int(nfft), self._ptr(taps), scale,
                                        int(frame_lo), int(n_frames), self._ptr(out), out.stride(0),
                                        fl, self._stream()))
        self.launches += 1
        return out

    def stft_f64(self, x, window, nperseg, noverlap, nfft, fs, scaling="density", stride=1, flag=None):
        """|STFT|^2*scale of x[::stride] for a complex128 CUDA tensor -> float64 [nfft, n_frames], rows fftshifted:
        scipy's double-precision path (b200vec_stft_f64), window taps and scale in float64."""
        x = x.reshape(-1)
        hop = nperseg - noverlap
        n_dec = (x.numel() + stride - 1) // stride
        n_frames = 0 if n_dec < nperseg else (n_dec - nperseg) // hop + 1
        key = (window, nperseg, float(fs), scaling, "f64")
        if key not in self._windows:
            taps, scale = H.window_and_scale(window, nperseg, fs, scaling, dtype=np.complex128)
            self._windows[key] = (torch.from_numpy(np.ascontiguousarray(taps, dtype=np.float64)).to(self.device),
                                  float(scale))
        taps, scale = self._windows[key]
        pitch = max((n_frames + 15) // 16 * 16, 16)
        out = torch.empty((nfft, pitch), dtype=torch.float64, device=self.device)[:, :n_frames]
        if n_frames > 0:
            check(self.lib.b200vec_stft_f64(self._ctx, self._ptr(x), 0, x.numel(), int(stride), int(nperseg), int(hop),
                                            int(nfft), self._ptr(taps), scale, 0, int(n_frames), self._ptr(out),
                                            out.stride(0), self._ptr(flag) if flag is not None else None,
                                            self._stream()))
            self.launches += 1
        return out

    def replicate_frames(self, spec, frame_src0, period_frames, frame_dst0, n_frames):
        """spec[:, frame_dst0 + i] = spec[:, frame_src0 + (frame_dst0 + i - frame_src0) % period_frames] for a
        frame-major spectrogram tensor (b200vec_replicate_frames): the periodic tiling of a dense vector's map."""
        if not (spec.dim() == 2 and spec.stride(0) == 1 and spec.stride(1) >= 1024):
            raise ValueError("replicate_frames needs frame-major spectrogram memory")
        check(self.lib.b200vec_replicate_frames(self._ctx, self._ptr(spec), spec.stride(1), int(frame_src0),
                                                int(period_frames), int(frame_dst0), int(n_frames), self._stream()))
        self.launches += 1

    # ------------------------------------------------------------------ dB / percentiles
    def db(self, s, floor_val, out=None):
        s = s.contiguous()
        out = torch.empty_like(s) if out is None else out
        check(self.lib.b200vec_db(self._ctx, self._ptr(s), s.numel(), float(floor_val), self._ptr(out),
                                  self._stream()))
        self.launches += 1
        return out

    def select_count(self, x, positive_only=False):
        n = C.c_int64()
        check(self.lib.b200vec_select_count(self._ctx, self._ptr(x), x.numel(), int(positive_only),
                                            C.byref(n), self._stream()))
        self.launches += 3
        return n.value

    def select_ranks(self, x, ranks, positive_only=False):
        ranks = [int(r) for r in ranks]
        arr = (C.c_int64 * len(ranks))(*ranks)
        vals = (C.c_float * len(ranks))()
        check(self.lib.b200vec_select_ranks(self._ctx, self._ptr(x), x.numel(), int(positive_only), arr,
                                            len(ranks), vals, self._stream()))
        self.launches += 7
        return np.asarray(list(vals), dtype=np.float32)

    def percentile(self, x, q, positive_only=False, count=None):
        """np.percentile(x[x>0] if positive_only else x, q) for a float32 CUDA tensor,
        bit-for-bit (exact order statistics + numpy's own float32 index/lerp arithmetic)."""
        x = x.contiguous().view(-1)
        if count is None:
            count = self.select_count(x, positive_only) if positive_only else x.numel()
        if count == 0:
            return None
        lo, hi, gamma = H.percentile_ranks(count, q, np.float32)
        v = self.select_ranks(x, [lo, hi], positive_only)
        return H.lerp(v[0], v[1], gamma)

    def select_begin(self, x, use_abs=False):
        """13-bit key histogram of a float32 CUDA tensor in one read (b200vec_select_begin).
        Returns the number of elements > 0 (after the optional abs)."""
        n_pos = C.c_int64()
        check(self.lib.b200vec_select_begin(self._ctx, self._ptr(x), x.numel(), int(use_abs), C.byref(n_pos),
                                            self._stream()))
        self.launches += 1
        return n_pos.value

    def select_finish(self, x, ranks):
        """Exact values at `ranks` (positions in the ascending sort of ALL elements of the tensor
        select_begin was called on): bin pick on the host, one compaction read, exact select on the
        candidates (b200vec_select_finish)."""
        ranks = [int(r) for r in ranks]
        arr = (C.c_int64 * len(ranks))(*ranks)
        vals = (C.c_float * len(ranks))()
        check(self.lib.b200vec_select_finish(self._ctx, self._ptr(x), x.numel(), arr, len(ranks), vals,
                                             self._stream()))
        self.launches += 8
        return np.asarray(list(vals), dtype=np.float32)

    def normalize_spectrogram(self, s, low_percentile=10.0, high_percentile=95.0, max_dynamic_range=25):
        """Device form of utils.normalize_spectrogram (utils.py:356-404) for a float32 CUDA
        tensor: returns (Sxx_db tensor, vmin, vmax).

        10*log10(|S| + floor) is monotone in |S|, so every order statistic the function needs — the 5th
        percentile of the positive elements (the floor) and the low / high percentiles, minimum and maximum
        of the dB image — is an order statistic of |S|: all eight ranks come out of ONE selection (two
        full reads of S), the dB image is written once, and the six dB values are the dB of six |S| values."""
        # order statistics and an elementwise map do not care about the element order: frame-major memory
        # (the transpose view the fused step hands out) is used as it lies, no 448 MB re-layout
        transposed = s.dim() == 2 and not s.is_contiguous() and s.t().is_contiguous()
        base = s.t() if transposed else s.contiguous()
        flat = base.view(-1)
        n = flat.numel()
        if n == 0:
            return torch.empty_like(s), 0, 0
        db = torch.empty_like(flat)
        vals = (C.c_float * 9)()
        n_pos = C.c_int64()
        check(self.lib.b200vec_normalize_db(self._ctx, self._ptr(flat), n, float(low_percentile), float(high_percentile),
                                            self._ptr(db), vals, C.byref(n_pos), self._stream()))
        self.launches += 5
        out = np.asarray(list(vals), dtype=np.float32)
        six, g1, g2 = out[1:7], out[7], out[8]
        db = db.view(base.shape)
        if transposed:
            db = db.t()
        vmin, vmax = H.lerp(six[0], six[1], g1), H.lerp(six[2], six[3], g2)
        vmin, vmax = H.colour_limits(vmin, vmax, six[4], six[5], max_dynamic_range)
        return db, vmin, vmax

    def decimate(self, x, stride):
        """x[::stride] as a compact complex64 tensor (b200vec_decimate; utils.py:192-195)."""
        x = x.reshape(-1)
        n_out = (x.numel() + stride - 1) // stride
        out = torch.empty(n_out, dtype=torch.complex64, device=self.device)
        check(self.lib.b200vec_decimate(self._ctx, self._ptr(x), x.numel(), int(stride), self._ptr(out), self._stream()))
        self.launches += 1
        return out

    # ------------------------------------------------------------------ scenario batch
    @staticmethod
    def pack_batch(scenarios):
        """The ctypes tables of b200vec_batch_prepare for `scenarios` (per vector a list of (start, period, count,
        src_slot, freq_hz[, free_flag])): built once, reusable for every step that re-prepares the same sweep."""
        from ._lib import BatchPkt
        n_pk = sum(len(sc) for sc in scenarios)
        pk = (BatchPkt * max(n_pk, 1))()
        first = (C.c_int32 * (len(scenarios) + 1))()
        i = 0
        for v, sc in enumerate(scenarios):
            first[v] = i
            for row in sc:
                start, period, count, slot, freq = row[:5]
                pk[i] = BatchPkt(int(start), int(period), int(count), int(slot),
                                 1 if (len(row) > 5 and row[5]) else 0, float(freq))
                i += 1
        first[len(scenarios)] = i
        jobs = len({(int(r[3]), float(r[4])) for sc in scenarios for r in sc})
        return {"pk": pk, "first": first, "n": len(scenarios), "jobs": jobs}

    def batch_prepare(self, scenarios, fs_hz):
        """scenarios: per vector a list of (start, period, count, src_slot, freq_hz[, free_flag]), or the
        tables pack_batch() made of it — every distinct (src_slot, freq_hz) is shifted once, in one launch
        (b200vec_batch_prepare)."""
        packed = scenarios if isinstance(scenarios, dict) else self.pack_batch(scenarios)
        check(self.lib.b200vec_batch_prepare(self._ctx, packed["pk"], packed["first"], packed["n"], float(fs_hz),
                                             self._stream()))
        self.launches += 1
        self._batch_n = packed["n"]
        self._last_batch = packed
        self._last_batch_jobs = packed["jobs"]

    def batch_build(self, total, normalize=True, out=None, maxsq=None):
        """All vectors of the prepared batch in ONE launch (b200vec_batch_build): returns
        (complex64 [n_vectors, total] CUDA tensor, maxsq float64 [n_vectors] or None)."""
        nv = self._batch_n
        ld = (total + 1) // 2 * 2
        if out is None:
            out = torch.empty((nv, max(ld, 2)), dtype=torch.complex64, device=self.device)
        if normalize and maxsq is None:
            maxsq = torch.empty(nv, dtype=torch.float64, device=self.device)
        check(self.lib.b200vec_batch_build(self._ctx, int(total), self._ptr(out), out.stride(0), int(bool(normalize)),
                                           self._ptr(maxsq) if normalize else None, self._stream()))
        self.launches += 1
        return out[:, :total], (maxsq if normalize else None)

    def peer_status(self):
        """1 when an in-kernel peer MAX exchange on this engine timed out since the last call."""
        st = C.c_int()
        check(self.lib.b200vec_peer_status(self._ctx, C.byref(st), self._stream()))
        return st.value

    # ------------------------------------------------------------------ .wv payload
    def wv_quantize(self, x, out=None, pre_scale=1.0):
        x = x.contiguous()
        out = torch.empty(2 * x.numel(), dtype=torch.int16, device=self.device) if out is None else out
        check(self.lib.b200vec_wv_quantize(self._ctx, self._ptr(x), x.numel(), float(pre_scale), self._ptr(out),
                                           self._stream()))
        self.launches += 1
        return out

    def power_stats(self, x, pre_scale=1.0):
        """(max |y|^2, sum fl32(|y|)^2) of y = x * pre_scale as Python floats (b200vec_power_stats)."""
        x = x.contiguous()
        st = (C.c_double * 2)()
        check(self.lib.b200vec_power_stats(self._ctx, self._ptr(x), x.numel(), float(pre_scale), st, self._stream()))
        self.launches += 1
        return float(st[0]), float(st[1])


_ENGINES = {}


def _split_planes(self, x, out=None):
    """complex64 CUDA tensor [n] -> float32 [2, n] (row 0 real, row 1 imaginary): the two planes a
    MAT v5 complex array is stored as (K7, b200vec_split_planes)."""
    x = x.reshape(-1)
    n = x.numel()
    pitch = (n + 3) // 4 * 4                                  # both rows 16-byte aligned
    if out is None:
        out = torch.empty((2, max(pitch, 4)), dtype=torch.float32, device=self.device)
    check(self.lib.b200vec_split_planes(self._ctx, self._ptr(x), n, self._ptr(out[0]), self._ptr(out[1]),
                                        self._stream()))
    self.launches += 1
    return out[:, :n]


Engine.split_planes = _split_planes


def _box_energy(self, x, window):
    """(smoothed float64 [n], max float64 [1]) — b200vec_box_energy (utils.py:813-816)."""
    x = x.reshape(-1)
    out = torch.empty(x.numel(), dtype=torch.float64, device=self.device)
    mx = torch.empty(1, dtype=torch.float64, device=self.device)
    check(self.lib.b200vec_box_energy(self._ctx, self._ptr(x), x.numel(), int(window), self._ptr(out),
                                      self._ptr(mx), self._stream()))
    self.launches += 1
    return out, mx


def _select_f64(self, x, ranks):
    """Exact values at `ranks` of the ascending sort of a float64 CUDA tensor (b200vec_select_f64)."""
    x = x.contiguous().view(-1)
    ranks = [int(r) for r in ranks]
    arr = (C.c_int64 * len(ranks))(*ranks)
    vals = (C.c_double * len(ranks))()
    check(self.lib.b200vec_select_f64(self._ctx, self._ptr(x), x.numel(), arr, len(ranks), vals, self._stream()))
    self.launches += 1
    return [float(v) for v in vals]


def _threshold_bounds(self, smoothed, threshold):
    """(first, last) index with smoothed >= threshold; (n, -1) when none — b200vec_threshold_bounds."""
    res = (C.c_int64 * 2)()
    check(self.lib.b200vec_threshold_bounds(self._ctx, self._ptr(smoothed), smoothed.numel(), float(threshold),
                                            res, self._stream()))
    self.launches += 1
    return int(res[0]), int(res[1])


def _median2_rows(self, s):
    """max(s[r-1], s[r]) per column, row 0 kept — b200vec_median2_rows; output has s's layout."""
    out = torch.empty_strided(s.shape, s.stride(), dtype=torch.float32, device=self.device)
    check(self.lib.b200vec_median2_rows(self._ctx, self._ptr(s), s.shape[0], s.shape[1], s.stride(0), s.stride(1),
                                        self._ptr(out), out.stride(0), out.stride(1), self._stream()))
    self.launches += 1
    return out


Engine.box_energy = _box_energy
Engine.threshold_bounds = _threshold_bounds
Engine.select_f64 = _select_f64
Engine.median2_rows = _median2_rows


def get_engine(device=None):
    """Process-wide Engine for `device` (default: torch's current CUDA device)."""
    if not torch.cuda.is_available():
        raise _lib.B200VecError(
            "vector_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    idx = torch.cuda.current_device() if device is None else (
        device.index if isinstance(device, torch.device) else int(device))
    if idx not in _ENGINES:
        _ENGINES[idx] = Engine(idx)
    return _ENGINES[idx]

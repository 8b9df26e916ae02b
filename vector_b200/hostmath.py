"""Host-side scalar logic of the vector-build path: everything the reference decides with
Python floats/ints before (or after) touching sample data.  It stays on the host verbatim
because the results are *defined* by Python double arithmetic and truncation
(SURVEY.md §7.3 item 3): placement integers, STFT parameter selection, window taps and
scale factors, np.percentile's index arithmetic, the timing-validation score.

Nothing here touches sample arrays; the numeric work is in the CUDA engine (engine.py).
Citations are into the reference (ramiyako/vector).
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass

import numpy as np

TARGET_SAMPLE_RATE = 56e6      # unified_gui.py:40 / main.py:17
MAX_PACKETS = 6                # unified_gui.py:41 / main.py:18


# --------------------------------------------------------------------------- placement
def total_samples(length_seconds, fs=TARGET_SAMPLE_RATE):
    """int(vector_length * TARGET_SAMPLE_RATE) — unified_gui.py:1711 (after ms/1000 at
    :1706) and main.py:217."""
    return int(length_seconds * fs)


def period_samples(period_s, fs=TARGET_SAMPLE_RATE):
    """int(cfg['period'] * fs), truncation — unified_gui.py:1748 / main.py:247."""
    return int(period_s * fs)


def start_offset(start_time_s, pre_buf, fs=TARGET_SAMPLE_RATE):
    """max(0, int(round(start*fs)) - pre_buf) — unified_gui.py:1754 / main.py:250
    (Python round = banker's rounding on the double product)."""
    return max(0, int(round(start_time_s * fs)) - pre_buf)


def instance_count(start, period, pkt_len, total):
    """How many instances `while pos + L <= total` places (unified_gui.py:1760-1769);
    main.py:253-257's range()/break form yields the same set."""
    if period <= 0:
        raise ValueError("period must be positive")
    last = total - pkt_len - start
    return 0 if last < 0 else last // period + 1


# --------------------------------------------------------------------------- windows
def _cosine_sum(n, coeffs):
    """scipy.signal.windows.general_cosine(n, coeffs, sym=False): evaluate on n+1 points
    of linspace(-pi, pi) and drop the last (the DFT-even / 'periodic' form get_window
    returns)."""
    if n <= 1:
        return np.ones(n, dtype=np.float64)
    grid = np.linspace(-np.pi, np.pi, n + 1, dtype=np.float64)
    w = np.zeros(n + 1, dtype=np.float64)
    for k, a in enumerate(np.asarray(coeffs, dtype=np.float64)):
        w += a * np.cos(k * grid)
    return w[:-1]


_WINDOW_COEFFS = {
    "hann": (0.5, 0.5),                                        # general_hamming(alpha=0.5)
    "blackmanharris": (0.35875, 0.48829, 0.14128, 0.01168),
}


def get_window(name, n):
    """float64 taps equal to scipy.signal.get_window(name, n) (fftbins=True) for the three
    windows create_spectrogram can reach: 'hann', 'blackmanharris', 'boxcar'
    (utils.py:276-279, 307, 330, 340)."""
    if name == "boxcar":
        return np.ones(n, dtype=np.float64)
    if name == "hann":
        a = 0.5
        return _cosine_sum(n, np.asarray([a, 1.0 - a]))
    if name in _WINDOW_COEFFS:
        return _cosine_sum(n, _WINDOW_COEFFS[name])
    raise ValueError(f"Unknown window type: {name}")


def window_and_scale(name, nperseg, fs, scaling, dtype=np.complex64):
    """(float32 taps, float32 scale) exactly as scipy's legacy _spectral_helper forms them
    for complex64 input: the window is cast to complex64 first and the scale is evaluated
    in that dtype — 1/(fs*sum(win*win)) for 'density', 1/sum(win)**2 for 'spectrum'."""
    win = get_window(name, nperseg).astype(dtype)
    if scaling == "density":
        scale = 1.0 / (fs * (win * win).sum())
    elif scaling == "spectrum":
        scale = 1.0 / win.sum() ** 2
    else:
        raise ValueError(f"Unknown scaling: {scaling!r}")
    real = np.float32 if np.dtype(dtype) == np.complex64 else np.float64     # complex128 input: everything in double
    return np.ascontiguousarray(win.real), np.asarray(scale).real.astype(real)[()]


# --------------------------------------------------------------------------- STFT plan
@dataclass(frozen=True)
class SpectrogramPlan:
    """What create_spectrogram settles on before calling scipy (utils.py:184-275)."""
    n_in: int
    factor: int          # decimation stride, sig[::factor]
    n_dec: int           # len(sig[::factor])
    fs: float            # sr / factor
    heavy: bool
    window: str
    nperseg: int
    noverlap: int
    nfft: int
    scaling: str = "spectrum"

    @property
    def hop(self):
        return self.nperseg - self.noverlap

    @property
    def n_frames(self):
        if self.n_dec < self.nperseg:
            return 0
        return (self.n_dec - self.nperseg) // self.hop + 1

    def times(self):
        """scipy: arange(nperseg/2, n - nperseg/2 + 1, nperseg - noverlap) / fs."""
        return np.arange(self.nperseg / 2, self.n_dec - self.nperseg / 2 + 1,
                         self.nperseg - self.noverlap) / float(self.fs)

    def freqs(self):
        """scipy.fft.fftfreq(nfft, 1/fs) — two-sided, un-shifted."""
        return np.fft.fftfreq(self.nfft, 1 / self.fs)


def plan_spectrogram(n, sr, max_samples=2_000_000, time_resolution_us=1,
                     adaptive_resolution=True):
    """Follow create_spectrogram's scalar decisions (utils.py:179-275) for a signal of n
    samples at rate sr."""
    if n == 0:
        raise ValueError("Signal is empty")                        # utils.py:180-181
    heavy = n > 5_000_000                                           # :184
    if heavy:
        max_samples = min(max_samples, 1_000_000)                   # :188
        time_resolution_us = max(time_resolution_us, 20)            # :189
    factor, n_dec, fs = 1, n, sr
    if n > max_samples:                                             # :192-195
        factor = int(np.ceil(n / max_samples))
        n_dec = (n + factor - 1) // factor
        fs = sr / factor
    duration_us = n_dec / fs * 1e6                                  # :203
    tr = time_resolution_us
    if adaptive_resolution:                                         # :206-231
        if duration_us <= 50:
            base, tr, stretch = max(32, min(n_dec // 12, 128)), min(tr, duration_us / 10), 1.2
        elif duration_us <= 500:
            base, tr, stretch = max(64, min(n_dec // 10, 256)), min(tr, duration_us / 20), 1.2
        elif duration_us <= 5000:
            base, tr, stretch = max(128, min(n_dec // 8, 512)), min(tr, 10), 1.5
        else:
            base, tr, stretch = max(256, min(n_dec // 6, 1024)), min(tr, 20), 1.5
            if heavy:
                base, tr, stretch = min(base, 512), max(tr, 50), 1.2
    else:                                                           # :232-235
        base, stretch = max(128, min(n_dec // 8, 512)), 1.2
    if tr is not None:                                              # :238-255
        step = max(1, int(round(fs * tr / 1e6)))
        step = max(1, min(step, n_dec // 10))
        nperseg = min(max(base, step * 2), n_dec)
        noverlap = max(0, nperseg - (2 * step if heavy else step))
    else:                                                           # :256-262
        nperseg = min(base, n_dec)
        noverlap = int(nperseg * (0.75 if heavy else 0.90))
    nfft = max(256, int(2 ** np.ceil(np.log2(nperseg * stretch))))  # :265
    nfft = min(nfft, 1024) if heavy else max(nfft, 512)             # :268-271
    return SpectrogramPlan(n, factor, n_dec, fs, heavy,
                           "hann" if heavy else "blackmanharris", nperseg, noverlap, nfft)


def fallback_plans(plan):
    """The degraded parameter sets of the fallback chain, in the order the reference tries
    them: exception fallback (utils.py:298-313), all-zero retry (:316-334), last resort
    (:335-347)."""
    n, fs = plan.n_dec, plan.fs
    w1 = min(256, n)
    w2 = min(64, n // 4)
    mk = lambda win, ns, no, nf: SpectrogramPlan(plan.n_in, plan.factor, n, fs, plan.heavy,
                                                 win, ns, no, nf)
    return dict(on_error=mk("hann", w1, w1 // 2, 512),
                on_zero=mk("hann", w2, w2 // 4, max(128, w2)),
                last_resort=mk("boxcar", 32, 16, 64))


def resolve_plan(plan):
    """Apply the argument handling scipy.signal.spectrogram performs before it computes
    (_triage_segments / _spectral_helper): an nperseg longer than the input is clamped to
    the input length (scipy warns), then the ValueErrors that drive the reference's
    except-branches are raised.  Returns the plan scipy would actually run."""
    nperseg = plan.nperseg
    if nperseg > plan.n_dec:
        nperseg = plan.n_dec
    if nperseg < 1:
        raise ValueError("nperseg must be a positive integer")
    if plan.nfft < nperseg:
        raise ValueError("nfft must be greater than or equal to nperseg.")
    if plan.noverlap >= nperseg:
        raise ValueError("noverlap must be less than nperseg.")
    if nperseg != plan.nperseg:
        plan = SpectrogramPlan(plan.n_in, plan.factor, plan.n_dec, plan.fs, plan.heavy,
                               plan.window, nperseg, plan.noverlap, plan.nfft, plan.scaling)
    return plan


# --------------------------------------------------------------------------- percentiles
def percentile_ranks(count, q, dtype=np.float32):
    """Index arithmetic of np.percentile(a, q) (method='linear') for a 1-D array of `count`
    elements of `dtype`: numpy divides q by dtype(100) and evaluates the virtual index
    (n-1)*q IN THE ARRAY'S DTYPE, so for float32 data the neighbours are decided by float32
    rounding (numpy/lib/_function_base_impl.py: percentile, _quantile, _get_indexes).
    Returns (lo_rank, hi_rank, gamma) with gamma of `dtype`."""
    qq = np.true_divide(q, dtype(100))
    vi = np.asanyarray((count - 1) * qq)                 # virtual index, array dtype
    prev = np.asanyarray(np.floor(vi))
    nxt = np.asanyarray(prev + 1)                        # may round back to prev above 2^24 (float32)
    if vi >= count - 1:                                  # _get_indexes: clamp to the last element
        prev, nxt = np.asanyarray(-1, dtype=vi.dtype), np.asanyarray(-1, dtype=vi.dtype)
    if vi < 0:
        prev, nxt = np.asanyarray(0, dtype=vi.dtype), np.asanyarray(0, dtype=vi.dtype)
    prev_i, nxt_i = prev.astype(np.intp), nxt.astype(np.intp)
    gamma = np.asanyarray(np.asanyarray(vi - prev_i), dtype=vi.dtype)   # _get_gamma
    lo, hi = int(prev_i), int(nxt_i)
    return (lo + count if lo < 0 else lo), (hi + count if hi < 0 else hi), gamma[()]


def lerp(a, b, t):
    """numpy's _lerp: a + (b-a)*t, switched to b - (b-a)*(1-t) for t >= 0.5, evaluated in
    the operands' dtype."""
    d = np.subtract(b, a)
    out = np.add(a, d * t)
    if t >= 0.5:
        out = np.subtract(b, d * (1 - t))
    return out


def colour_limits(vmin, vmax, db_min, db_max, max_dynamic_range=25):
    """Post-processing of the two dB percentiles (utils.py:375-397)."""
    if np.isnan(vmin) or np.isnan(vmax) or vmax <= vmin:
        vmin, vmax = db_min, db_max
        if vmax <= vmin:
            vmax = vmin + max_dynamic_range
    span = vmax - vmin
    if span > max_dynamic_range:
        vmin = vmax - max_dynamic_range
    elif span < 20:
        mid = (vmax + vmin) / 2
        vmin, vmax = mid - 10, mid + 10
    return max(vmin, -120), vmax


# --------------------------------------------------------------------------- misc helpers
def compute_freq_ranges(shifts, margin=1e6):
    """utils.py:129-159 — one (min-margin, max+margin) range, or None when every shift is 0
    / nothing numeric is left."""
    if not shifts or all(s == 0 for s in shifts):
        return None
    numeric = [s for s in shifts if isinstance(s, (int, float)) and not isinstance(s, bool)]
    if not numeric:
        return None
    return [(min(numeric) - margin, max(numeric) + margin)]


_STATUS = ((99.5, "✅ PERFECT"), (99.0, "✅ EXCELLENT"), (95.0, "⚠️ GOOD"), (90.0, "⚠️ FAIR"))


def validate_packet_timing(markers, configs, verbose=False):
    """Timing-validation score shown in the plot title (unified_gui.py:1496-1690).

    ``markers``: tuples whose first three fields are (time_s, freq_shift, packet_name);
    ``configs``: get_config() dicts (None entries are skipped).  Per packet: start-time,
    period, frequency and consistency accuracies, weighted 30/40/20/10, plus up to 5 points
    for > 2 instances; the mean picks the status word.  Returns (title, details)."""
    groups = {}
    for m in markers:
        groups.setdefault(m[2], []).append((m[0], m[1]))
    for g in groups.values():
        g.sort(key=lambda e: e[0])

    rows, details = [], []
    for cfg in configs:
        if not cfg:
            continue
        name = os.path.splitext(os.path.basename(cfg["file"]))[0]
        if name not in groups:
            continue
        t = [e[0] for e in groups[name]]
        fshift = [e[1] for e in groups[name]]
        n = len(t)
        exp_period = cfg["period"] * 1000
        exp_start = cfg["start_time"] * 1000
        exp_shift = cfg["freq_shift"]

        s_acc, s_msg, s_err = 100.0, "", 0.0
        if n > 0:
            s_err = abs(t[0] * 1000 - exp_start)
            limit = 10.0 * (1.0 if n <= 2 else 0.5)
            if s_err <= limit:
                s_msg = f"Start time accurate ({s_err:.2f}ms error)"
            else:
                s_acc = max(0, 100 - (s_err / limit * 50))
                s_msg = f"Start time inaccurate ({s_err:.2f}ms error)"

        p_acc, p_msg, p_err = 100.0, "", 0.0
        if n > 1:
            iv = [(t[i] - t[i - 1]) * 1000 for i in range(1, n)]
            mean_iv, std_iv = np.mean(iv), np.std(iv)
            p_err = abs(mean_iv - exp_period) / exp_period * 100
            if p_err <= 1.0:
                p_msg = f"Period accurate ({mean_iv:.2f}ms avg, {std_iv:.2f}ms std)"
            elif p_err <= 5.0:
                p_acc = 100 - (p_err - 1.0) * 5
                p_msg = f"Period good ({mean_iv:.2f}ms avg, {p_err:.1f}% error)"
            else:
                p_acc = max(0, 80 - (p_err - 5.0) * 2)
                p_msg = f"Period inaccurate ({mean_iv:.2f}ms avg, {p_err:.1f}% error)"
        elif n == 1:
            p_msg = "Single instance - no period to validate"

        f_msg = ""
        if fshift:
            distinct = set(fshift)
            if len(distinct) > 1:
                f_msg = f"Mixed frequency shifts ({len(distinct)} different shifts)"
            elif fshift[0] == exp_shift:
                f_msg = f"Frequency shift correct ({exp_shift/1e6:.1f}MHz)"
            elif abs(fshift[0] - exp_shift) / 1e6 <= 0.1:
                f_msg = f"Frequency shift close to expected ({fshift[0]/1e6:.1f}MHz)"
            else:
                f_msg = (f"Frequency shift differs from expected ({fshift[0]/1e6:.1f}MHz vs "
                         f"{exp_shift/1e6:.1f}MHz)")

        c_acc, c_msg = 100.0, ""
        if n >= 2:
            c_msg = f"Consistent instances ({n} instances)"
        elif n == 1:
            c_acc, c_msg = 80.0, "Single instance - cannot validate consistency"

        rows.append(dict(packet_name=name, start_accuracy=s_acc, period_accuracy=p_acc,
                         freq_accuracy=100.0, consistency_accuracy=c_acc, instances=n,
                         start_explanation=s_msg, period_explanation=p_msg,
                         freq_explanation=f_msg, consistency_explanation=c_msg))
        details.append(dict(packet_name=name, instances=n,
                            start_time_error_ms=s_err if n > 0 else 0,
                            period_error_percent=p_err if n > 1 else 0,
                            freq_shift_mhz=exp_shift / 1e6,
                            explanations=dict(start=s_msg, period=p_msg, freq=f_msg,
                                              consistency=c_msg)))
    if not rows:
        return "No packets to validate", []

    acc_sum = 0
    for r in rows:
        a = (r["period_accuracy"] * 0.4 + r["start_accuracy"] * 0.3 +
             r["freq_accuracy"] * 0.2 + r["consistency_accuracy"] * 0.1)
        if r["instances"] > 2:
            a = min(100.0, a + min(5.0, (r["instances"] - 2) * 1.0))
        acc_sum += a
    overall = acc_sum / len(rows)
    status = "❌ POOR"
    for floor_, word in _STATUS:
        if overall > floor_:
            status = word
            break
    if verbose:
        print("\n" + "=" * 80 + "\n🔍 ENHANCED PACKET TIMING VALIDATION RESULTS\n" + "=" * 80)
        for r in rows:
            print(f"\n📦 {r['packet_name']}: {r['instances']} instances")
            print(f"   🎯 Start Time: {r['start_accuracy']:.1f}% - {r['start_explanation']}")
            print(f"   ⏱️  Period: {r['period_accuracy']:.1f}% - {r['period_explanation']}")
            print(f"   📡 Frequency: {r['freq_accuracy']:.1f}% - {r['freq_explanation']}")
            print(f"   📊 Consistency: {r['consistency_accuracy']:.1f}% - {r['consistency_explanation']}")
        print(f"\n🏆 Overall Accuracy: {overall:.1f}% - {status}")
        print(f"📈 Total Packets Validated: {len(rows)}")
        print(f"📊 Total Packet Instances: {sum(r['instances'] for r in rows)}")
        print("=" * 80)
    return f"Timing Validation: {overall:.1f}% {status}", details


# --------------------------------------------------------------------------- sharding
def slab_bounds(total, world, rank, align=16):
    """Contiguous time slab of `rank` out of `world`: [g*N/G, (g+1)*N/G) rounded down to
    `align` samples (128 B of complex64); the last rank takes the remainder."""
    if world <= 1:
        return 0, total
    def cut(g):
        if g >= world:
            return total
        return min(total, (g * total // world) // align * align)
    return cut(rank), cut(rank + 1)


def frame_range_for_slab(lo, hi, total, stride, nperseg, hop, n_frames):
    """Frames whose FIRST sample (vector index f*hop*stride) lies in [lo, hi): the rank that
    owns a frame recomputes the <= (nperseg-1)*stride trailing halo locally instead of
    exchanging it.  Returns (f_lo, f_hi, halo_hi) with halo_hi the exclusive
    end of the vector samples those frames read."""
    step = hop * stride
    f_lo = min(n_frames, -(-lo // step))
    f_hi = min(n_frames, -(-hi // step))
    if f_hi <= f_lo:
        return f_lo, f_lo, hi
    last = ((f_hi - 1) * hop + nperseg - 1) * stride + 1
    return f_lo, f_hi, max(hi, min(total, last))


# --------------------------------------------------------------------------- max-abs pass
def common_period(rows, lengths, total):
    """(A, B, T) such that the composite repeats bit for bit with period T on [A, B + T):
    v(i + T) == v(i) for A <= i < B — or None when the trains share no usable period.

    The composite is a sum of periodic trains, v(i) = sum_p sum_k y_p[i - s_p - k P_p]
    (unified_gui.py:1757-1769).  Packet p's train repeats exactly — same y values added in the
    same (packet, instance) order, hence bit-identical float32 sums — under a shift by its period
    wherever no instance is entering or leaving:
        c_p(i + P_p) == c_p(i)   for   max(0, s_p + L_p - P_p) <= i < s_p + (K_p - 1) P_p.
    With T = lcm(P_p):  A = max_p max(0, s_p + L_p - P_p),   B = min_p (s_p + K_p P_p - T).
    rows: (start, period, count, slot); lengths: packet lengths."""
    live = [(int(s), int(p), int(k), int(n)) for (s, p, k, _), n in zip(rows, lengths) if k > 0]
    if not live or total <= 0 or any(k < 2 for _, _, k, _ in live):
        return None
    t = 1
    for _, p, _, _ in live:
        t = t * p // math.gcd(t, p)
        if t >= total:
            return None
    a = max(max(0, s + n - p) for s, p, _, n in live)
    b = min(s + k * p - t for s, p, k, _ in live)
    if b <= a or a + t >= total:
        return None
    return a, b, t


def max_eval_ranges(rows, lengths, total):
    """Sample ranges over which max|v| has to be evaluated to get the max over [0, total).

    With (A, B, T) = common_period(...): every value taken on [A, B + T) is already taken on
    [A, A + T), so   max over [0, total) = max over [0, A + T) U [B + T, total).
    Falls back to the whole vector when the trains share no usable period (a packet placed
    once, T too long, B <= A).  Returns a list of disjoint, ascending (lo, hi)."""
    whole = [(0, int(total))] if total > 0 else []
    cp = common_period(rows, lengths, total)
    if cp is None:
        return whole
    a, b, t = cp
    out = [(0, a + t)]
    if b + t < total:
        out.append((b + t, int(total)))
    return out


def tile_plan(rows, lengths, total, lo, hi, max_tile_samples=6_000_000, min_density=1.5, min_gain=1.0):
    """Periodic tiling of the slab [lo, hi) (multiples of 512 except at the vector's end): when the packet
    trains share a common period T that fits the L2 (max_tile_samples * 8 bytes) and overlap densely
    (sum_p L_p / P_p >= min_density instance reads per output sample), the steady part of the slab is the
    bit-exact repetition of its first T samples —
        v(i) == v(i - T)   for   A + T <= i < B + T      (common_period)
    — so it can be PLACED like one pseudo-packet of length and period T read from the vector itself
    (one read per sample instead of one per overlapping instance).  Returns None, or
    dict(x1, x2, T, count): [lo, x1) and [x2, hi) are built from the packet bank, [x1, x2) from
    out[x1 - T, x1) placed at x1, x1 + T, ... (`count` instances, the last also covering the 512-sample
    halo past x2); x1 and x2 are multiples of 512.  spec_period_frames = lcm(T, 512) / 512: frame f of the
    hop-512 spectrogram holds the same samples as frame f - spec_period_frames wherever both lie in [x1, x2)."""
    cp = common_period(rows, lengths, total)
    if cp is None:
        return None
    a, b, t = cp
    if t > max_tile_samples:
        return None
    density = sum(n / p for (_, p, k, _), n in zip(rows, lengths) if k > 0)
    if density < min_density:
        return None
    x1 = -(-(max(lo, a) + t) // 512) * 512
    lim = b + t                                       # v(i) == v(i - T) holds for i < lim
    x2 = min(hi, (lim - 512) // 512 * 512)
    if hi % 512 and x2 == hi:
        x2 = hi // 512 * 512
    if x2 - x1 < max(512, int(min_gain * t)):
        return None
    # the full-rate spectrogram (hop 512) of the periodic part repeats every lcm(T, 512) / 512 frames
    return dict(x1=x1, x2=x2, T=t, count=-(-(x2 + 512 - x1) // t), spec_period_frames=t // math.gcd(t, 512))


def split_ranges(ranges, world, rank):
    """Rank `rank`'s share of a list of sample ranges: the concatenated index space is cut into
    `world` equal contiguous pieces (any rank can evaluate any sample — placement is a pure
    function of the descriptors)."""
    size = sum(hi - lo for lo, hi in ranges)
    if world <= 1:
        return list(ranges)
    c0, c1 = rank * size // world, (rank + 1) * size // world
    out, pos = [], 0
    for lo, hi in ranges:
        n = hi - lo
        x0, x1 = max(c0, pos), min(c1, pos + n)
        if x1 > x0:
            out.append((lo + x0 - pos, lo + x1 - pos))
        pos += n
    return out


def free_instance_flags(rows, lengths, total, max_instances=4_000_000):
    """Per packet: does EVERY sample of the packet appear unsummed somewhere in [0, total) — for every packet-local
    index n, is there an instance k with no other instance (of any packet, itself included) over sample
    start + k*period + n?  Then max|y_p|^2 over the whole packet is attained by the vector, and the max pass may
    skip every sample lying under exactly one instance of it (b200vec_pkt_desc.reserved bit 0; the packet maximum
    enters the reduction from the bank).  One free-standing instance is the simple case; a train whose neighbours
    overlap by their pre-buffers (period slightly below the packet length) qualifies too: the head is unsummed in the
    first instance, the tail in the last, the middle in any.  rows: (start, period, count, slot).
    Conservative: all False when there are more than `max_instances` instances."""
    rows = [(int(s), int(p), int(k)) for s, p, k, _ in rows]
    n_inst = sum(k for _, _, k in rows)
    flags = [False] * len(rows)
    if n_inst == 0 or n_inst > max_instances:
        return flags
    starts = np.concatenate([s + p * np.arange(k, dtype=np.int64) for s, p, k in rows])
    lens = np.concatenate([np.full(k, int(n), np.int64) for (_, _, k), n in zip(rows, lengths)])
    owner = np.concatenate([np.full(k, i, np.int64) for i, (_, _, k) in enumerate(rows)])
    ends = starts + lens
    # sweep over the instance boundaries: coverage and the sum of the active instances' ids between them
    ids = np.arange(n_inst, dtype=np.int64)
    pos = np.concatenate([starts, ends])
    delta = np.concatenate([np.ones(n_inst, np.int64), -np.ones(n_inst, np.int64)])
    idd = np.concatenate([ids, -ids])
    order = np.lexsort((delta, pos))                      # ends before starts at equal positions
    pos, delta, idd = pos[order], delta[order], idd[order]
    cov = np.cumsum(delta)[:-1]
    act = np.cumsum(idd)[:-1]
    seg_a, seg_b = pos[:-1], pos[1:]
    single = (cov == 1) & (seg_b > seg_a) & (seg_b <= int(total))
    inst = act[single]                                    # the one active instance of each single-covered segment
    la = seg_a[single] - starts[inst]                     # packet-local ranges that appear unsummed
    lb = seg_b[single] - starts[inst]
    ow = owner[inst]
    for i in range(len(rows)):
        if rows[i][2] == 0:
            continue
        m = ow == i
        if not m.any():
            continue
        a_i, b_i = la[m], lb[m]
        o = np.argsort(a_i, kind="stable")
        a_i, b_i = a_i[o], b_i[o]
        reach = np.maximum.accumulate(b_i)
        # covered [0, L): starts at 0, no gap between consecutive (sorted) ranges, reaches L
        flags[i] = bool(a_i[0] == 0 and reach[-1] >= int(lengths[i]) and np.all(a_i[1:] <= reach[:-1]))
    return flags

"""Parity rules of SURVEY.md §8c, shared by the GPU tests, smoke() and bench.py's checks."""
import numpy as np

VEC_REL_L2 = 1e-5          # BASELINE.json north_star: vector samples within 1e-5 relative L2
SPEC_DB = 0.01             # spectrogram within 0.01 dB ...
SPEC_FLOOR = 1e-8          # ... on bins >= -80 dBc (power ratio 1e-8 below the global peak)


def rel_l2(got, want):
    want = np.asarray(want)
    d = np.linalg.norm(np.asarray(got, dtype=np.complex128) - want.astype(np.complex128))
    n = np.linalg.norm(want.astype(np.complex128))
    return d / n if n > 0 else d


def ulp_diff_f32(a, b):
    """max |a-b| in float32 ulps over interleaved float32 views (treats -0 == +0)."""
    a = np.ascontiguousarray(a).view(np.float32).astype(np.float64)
    b = np.ascontiguousarray(b).view(np.float32).astype(np.float64)
    scale = np.maximum(np.spacing(np.abs(b).astype(np.float32)).astype(np.float64), 1e-45)
    return float(np.max(np.abs(a - b) / scale)) if a.size else 0.0


def max_err_rel_peak(got, want):
    """max |got-want| / max |want|: the per-sample error relative to the vector's peak.  (ulps of
    the individual samples are meaningless once packets are summed: cancellation leaves tiny
    values whose last bits depend on 1-ulp differences in the operands.)"""
    want = np.asarray(want)
    peak = np.max(np.abs(want)) if want.size else 0.0
    d = np.max(np.abs(np.asarray(got).astype(np.complex128) - want.astype(np.complex128))) if want.size else 0.0
    return float(d / peak) if peak > 0 else float(d)


def spec_db_error(S, S_ref, floor=SPEC_FLOOR):
    """max |10log10 S - 10log10 S_ref| over bins with S_ref >= peak*floor."""
    S, S_ref = np.asarray(S, np.float64), np.asarray(S_ref, np.float64)
    peak = S_ref.max()
    if peak <= 0:
        return 0.0 if not np.any(S) else np.inf
    m = S_ref >= peak * floor
    with np.errstate(divide="ignore"):
        return float(np.max(np.abs(10 * np.log10(S[m]) - 10 * np.log10(S_ref[m]))))


def band_errors(S, S_true, lo_db=-140, step=20):
    """(max, p99) of |ddB| vs the fp64 truth per 20 dB band below the global peak."""
    S, S_true = np.asarray(S, np.float64), np.asarray(S_true, np.float64)
    peak = S_true.max()
    out = {}
    with np.errstate(divide="ignore", invalid="ignore"):
        lvl = 10 * np.log10(S_true / peak)
        err = np.abs(10 * np.log10(S) - 10 * np.log10(S_true))
    for top in range(0, lo_db, -step):
        m = (lvl < top) & (lvl >= top - step) & np.isfinite(err)
        if m.sum() >= 100:
            e = err[m]
            out[(top, top - step)] = (float(e.max()), float(np.percentile(e, 99)))
    return out

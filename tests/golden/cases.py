"""Case definitions shared by the golden generator (make_golden.py) and the tests.

All inputs are synthetic and seeded (oracle.ref_port.synth_packet), so the same packets can
be rebuilt anywhere; only the reference's OUTPUTS are stored as fixtures.
"""
import hashlib

import numpy as np

FS = 56e6


def rows(*r):
    """(kind, shift MHz, period ms, start ms) -> get_config()-style dicts
    (unified_gui.py:854-869 unit conversions)."""
    from oracle.ref_port import _rows_to_configs
    return _rows_to_configs(list(r))


def build_cases():
    """name -> dict(configs, length_ms, normalize)."""
    from oracle.ref_port import cfg1_configs, cfg2_configs
    return {
        # test_complete_workflow.py:124-129 timings, two packets, 2.5 ms, no shift
        "wf_2p5ms_plain": dict(configs=rows(("tone", 0, 1.0, 0.2), ("chirp", 0, 1.2, 0.6)),
                               length_ms="2.5", normalize=False),
        # same with shifts (test_unified_gui.py:247-251) and normalisation
        "wf_2p5ms_shift_norm": dict(configs=rows(("tone", 2, 1.0, 0.2), ("chirp", -3, 1.2, 0.6),
                                                 ("bpsk", 1, 1.5, 0.0)),
                                    length_ms="2.5", normalize=True),
        # self-overlap: period shorter than the packet (0.4 ms < 1.0018 ms)
        "overlap_5ms": dict(configs=rows(("noise", -8, 0.4, 0.05), ("multitone", 6, 0.7, 0.0)),
                            length_ms="5", normalize=True),
        # vector shorter than any packet -> all zeros, max 0, no normalisation
        "empty_0p5ms": dict(configs=rows(("tone", 2, 1.0, 0.2)), length_ms="0.5",
                            normalize=True),
        "cfg1_100ms": dict(configs=cfg1_configs(), length_ms="100", normalize=True),
        "cfg2_1000ms": dict(configs=cfg2_configs(), length_ms="1000", normalize=True),
    }


def plan_lengths():
    """Signal lengths whose derived STFT parameters are pinned (utils.py:184-275)."""
    return [1, 5, 33, 500, 2000, 2800, 5600, 28000, 56100, 140000, 280001, 1_000_000,
            2_000_001, 5_000_000, 5_000_001, 5_600_000, 56_000_000]


def sha1(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def window_starts(n, width=256, count=48):
    if n <= width:
        return np.array([0], dtype=np.int64)
    return np.unique(np.linspace(0, n - width, count).astype(np.int64))


def summarize(vec, width=256, count=48):
    """Compact, hash-anchored summary of a complex64 vector."""
    vec = np.ascontiguousarray(vec)
    n = vec.shape[0]
    starts = window_starts(n, width, count)
    win = np.stack([vec[s:s + min(width, n)] for s in starts]) if n else np.zeros((0, 0), vec.dtype)
    a = np.abs(vec)
    return dict(n=int(n), sha1=sha1(vec), sum_abs=float(a.sum(dtype=np.float64)),
                max_abs=float(a.max()) if n else 0.0,
                nonzero=int(np.count_nonzero(vec)), starts=starts, windows=win)

"""Seeded synthetic packets and the BASELINE scenario tables.

The reference ships its test packets as MAT fixtures made by a generator that is absent
from it (create_test_packets.py, referenced at test_complete_workflow.py:352); SURVEY.md §4.3
recovered their closed forms.  This module regenerates equivalent packets anywhere (the role
utils.generate_sample_packet / create_sample_packets, utils.py:679-686,766-782, play for the
reference) so benchmarks and scenario sweeps need no files.
"""
from __future__ import annotations

import numpy as np

FS = 56e6
KINDS = ("tone", "chirp", "bpsk", "noise", "multitone")


def packet(kind, fs=FS, body_s=1e-3, pre=100, seed=0):
    """`pre` samples of 0.01*(N+jN) followed by a `body_s` body: tone 5 MHz | chirp 1->19 MHz |
    BPSK 1 Msym/s on 7 MHz | noise sigma 0.125 | (2+5+8 MHz)/3.  complex128."""
    rng = np.random.default_rng(seed)
    n = int(round(body_s * fs))
    t = np.arange(n) / fs
    head = 0.01 * (rng.standard_normal(pre) + 1j * rng.standard_normal(pre))
    if kind == "tone":
        body = np.exp(2j * np.pi * 5e6 * t)
    elif kind == "chirp":
        body = np.exp(2j * np.pi * (1e6 * t + 0.5 * (19e6 - 1e6) / body_s * t * t))
    elif kind == "bpsk":
        sps = max(1, int(round(fs / 1e6)))
        sym = rng.integers(0, 2, size=-(-n // sps)) * 2 - 1
        body = np.repeat(sym, sps)[:n] * np.exp(2j * np.pi * 7e6 * t)
    elif kind == "noise":
        body = 0.125 * (rng.standard_normal(n) + 1j * rng.standard_normal(n))
    elif kind == "multitone":
        body = sum(np.exp(2j * np.pi * f * t) for f in (2e6, 5e6, 8e6)) / 3
    else:
        raise ValueError(kind)
    return np.concatenate([head, body]).astype(np.complex128)


def bank(fs=FS):
    """kind -> (complex64 samples, pre_samples=100), seeds 101..105."""
    return {k: (packet(k, fs, seed=101 + i).astype(np.complex64), 100) for i, k in enumerate(KINDS)}


def configs(rows):
    """(kind, shift MHz, period ms, start ms) -> get_config()-style dicts (unified_gui.py:854-869)."""
    return [dict(file=f"packet_{KINDS.index(k) + 1}_{k}.mat", kind=k, freq_shift=float(s) * 1e6,
                 period=float(p) / 1000.0, start_time=float(st) / 1000.0) for k, s, p, st in rows]


CFG1_ROWS = [("tone", 2, 1.0, 0.2), ("chirp", -3, 1.2, 0.6), ("bpsk", 1, 1.5, 0.0)]
CFG2_ROWS = [("tone", 2, 100, 0), ("chirp", -3, 50, 10), ("bpsk", 1, 20, 20),
             ("noise", -8, 10, 30), ("multitone", 6, 5, 40), ("tone", -12, 2, 50)]
CFG4_ROWS = [("tone", 20, 2.0, 0.0), ("chirp", -35, 1.5, 0.1), ("bpsk", 10, 1.0, 0.2),
             ("noise", -60, 0.8, 0.3), ("multitone", 45, 0.5, 0.4), ("tone", -80, 0.4, 0.5)]


def scenario(rows, fs=FS):
    cfgs = configs(rows)
    b = bank(fs)
    return cfgs, [b[c["kind"]] for c in cfgs]


def cfg5_rows(v):
    """Scenario v of the 256-vector sweep."""
    rng = np.random.default_rng(5000 + v)
    return [(k, round(float(rng.uniform(-20, 20)), 1), round(float(rng.uniform(1, 10)), 1),
             round(float(rng.uniform(0, 5)), 1)) for k in ("tone", "chirp", "bpsk")]

"""Per-call timing of the pieces of normalize_spectrogram (K5) on the full-rate cfg2 spectrogram."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vector_b200 import generate, synth, hostmath as H
from vector_b200.engine import get_engine

FS, N = 56e6, 56_000_000
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)
v, _ = eng.build_vector(rows, N, True)
S = eng.stft(v, "hann", 1024, 512, 1024, FS, frame_major=False)
flat = S.contiguous().view(-1)
n = flat.numel()


def t(name, fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r = fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b))
    print(f"{name:40s} {sorted(ts)[len(ts)//2]:8.3f} ms", flush=True)
    return r


cnt = t("select_count(positive)", lambda: eng.select_count(flat, True))
lo, hi, g = H.percentile_ranks(cnt, 5, np.float32)
t("select_ranks 2 (positive)", lambda: eng.select_ranks(flat, [lo, hi], True))
db = t("db", lambda: eng.db(flat, 1e-9))
t("select_ranks 6 (dB image)", lambda: eng.select_ranks(db, [n // 20, n // 20 + 1, n * 19 // 20, n * 19 // 20 + 1, 0, n - 1]))
t("select_ranks 1 (dB image, median)", lambda: eng.select_ranks(db, [n // 2]))
t("torch.sort for scale (library)", lambda: torch.sort(db), 2)

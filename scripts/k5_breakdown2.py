"""Per-call timing of normalize_spectrogram (K5) as the engine runs it now (one selection on |S| + one dB pass) on
the full-rate cfg2 spectrogram, frame-major as the fused step leaves it."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vector_b200 import dist as vdist, generate, synth, hostmath as H
from vector_b200.engine import get_engine

FS, N = 56e6, 56_000_000
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)
plan = vdist.plan_slab(N, 1, 0, 1024, 512, 1, align=512)
vec, S, _ = vdist.build_stft_slab(eng, rows, plan, True, None, fs=FS)
flat = S.t().view(-1)
n = flat.numel()


def t(name, fn, reps=7):
    fn(); torch.cuda.synchronize()
    ts, ws = [], []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w0 = time.perf_counter()
        a.record(); r = fn(); b.record(); b.synchronize(); ts.append(a.elapsed_time(b)); ws.append((time.perf_counter() - w0) * 1e3)
    print(f"{name:44s} device {sorted(ts)[len(ts)//2]:8.3f} ms   wall {sorted(ws)[len(ws)//2]:8.3f} ms", flush=True)
    return r


n_pos = t("select_begin(|S|)", lambda: eng.select_begin(flat, use_abs=True))
lo1, hi1, _ = H.percentile_ranks(n, 5, np.float32)
lo2, hi2, _ = H.percentile_ranks(n, 95, np.float32)
lo0, hi0, _ = H.percentile_ranks(n_pos, 5, np.float32)
ranks = [lo1, hi1, lo2, hi2, 0, n - 1, n - n_pos + lo0, n - n_pos + hi0]
t("select_finish(8 ranks)", lambda: (eng.select_begin(flat, use_abs=True), eng.select_finish(flat, ranks))[1])
t("db", lambda: eng.db(flat, 1e-9))
t("normalize_spectrogram(S, 5, 95, 25)", lambda: eng.normalize_spectrogram(S, 5, 95, 25))

"""Per-phase cycle accounting of pipe1024_kernel (timing build: nvcc -DB2_PIPE_TIMING -o libb200vec_timing.so)."""
import os, sys, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from vector_b200 import _lib
_lib.LIB_PATH = os.path.join(ROOT, "vector_b200", "csrc", "libb200vec_timing.so")
import numpy as np, torch
from vector_b200 import dist as vdist, generate, synth
from vector_b200.engine import get_engine
from vector_b200._lib import BUILD_MAXONLY, BUILD_SCALE
FS, N = 56e6, 56_000_000
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)
plan = vdist.plan_slab(N, 1, 0, 1024, 512, 1, align=512)
nf = plan.frame_hi - plan.frame_lo
vec = torch.empty(N, dtype=torch.complex64, device=eng.device)
spec = eng.alloc_spectrogram(1024, nf)
descs, nd = eng.make_descs(rows)
eng._maxsq.zero_()
eng.build_pass(descs, nd, plan.lo, plan.hi, None, BUILD_MAXONLY)
for _ in range(3):
    eng.build_stft1024(descs, nd, plan.lo, plan.hi, vec, BUILD_SCALE, plan.frame_lo, nf, spec, window="hann", fs=FS, scaling="density")
torch.cuda.synchronize()
buf = np.zeros((320, 16, 8), np.int64)
lib = _lib.load()
lib.b200vec_debug_pipe_prof.argtypes = [C.c_void_p]
assert lib.b200vec_debug_pipe_prof(buf.ctypes.data) == 0
b = buf[:(296 if os.environ.get("B200VEC_FUSED_W")=="8" else 148)]
tot = b.sum(axis=2)
print("cycles per warp (mean over CTAs):", tot.mean(axis=0).astype(int))
col, row = b[:, :8].mean(axis=(0, 1)), b[:, 8:].mean(axis=(0, 1))
if os.environ.get("B200VEC_PIPE") != "1":
    cta = np.zeros((320, 4), np.int64)
    lib.b200vec_debug_pipe_cta.argtypes = [C.c_void_p]
    lib.b200vec_debug_pipe_cta(cta.ctypes.data)
    ncta = len(b)
    for sm in sorted(set(cta[:ncta, 0]))[:3]:
        rows_ = [(int(r[1]), int(r[2]), i) for i, r in enumerate(cta[:ncta]) if r[0] == sm]
        t0 = min(r[0] for r in rows_)
        print(f"SM {sm}: " + ", ".join(f"cta {i}: [{a - t0}, {e - t0}]" for a, e, i in rows_))
    names = ["P2 (placement / span wait)", "barrier A", "P3 LDS + FFT", "P3 STS + setup", "barrier B", "P5 LDS + FFT", "P6 PSD + stores", "-"]
    allw = b.mean(axis=(0, 1))
    print("lock-step kernel, all warps:"); [print(f"  {n:32s} {v:10.0f}  {100*v/allw.sum():5.1f}%   per tile {v*len(b)/(6836*(2 if os.environ.get("B200VEC_FUSED_W")=="8" else 1)):7.0f}") for n, v in zip(names, allw) if v]
    sys.exit(0)
names_c = ["wait stage_full", "bulk store / zero / sweep", "LDS + FFT", "wait exch_free", "STS + arrive", "stage_free: arrive (+warp0: wait, setup, issue)", "-", "loop"]
names_r = ["wait exch_full", "LDS (+arrive)", "FFT + stores", "-", "-", "-", "-", "loop"]
print("column warps:"); [print(f"  {n:48s} {v:10.0f}  {100*v/col.sum():5.1f}%") for n, v in zip(names_c, col) if v]
print("row warps:"); [print(f"  {n:48s} {v:10.0f}  {100*v/row.sum():5.1f}%") for n, v in zip(names_r, row) if v]
print("warp 0 (producer):"); [print(f"  {n:48s} {v:10.0f}") for n, v in zip(names_c, b[:, 0].mean(axis=0)) if v]

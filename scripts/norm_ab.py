"""A/B timing of the ONE-launch step (in-kernel max + fused build + STFT) on the BASELINE shape.
Usage: python scripts/norm_ab.py [ENV=VAL,ENV=VAL ...]   (each argument is one variant; first = reference)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vector_b200 import dist as vdist, generate, synth
from vector_b200.engine import get_engine

FS, N = 56e6, int(os.environ.get("AB_N", 56_000_000))
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)
plan = vdist.plan_slab(N, 1, 0, 1024, 512, 1, align=512)
nf = plan.frame_hi - plan.frame_lo
vec = torch.empty(N, dtype=torch.complex64, device=eng.device)
spec = eng.alloc_spectrogram(1024, nf, frame_major=True)
KEYS = ("B200VEC_WF_ZW", "B200VEC_WF_TF", "B200VEC_WF_CHUNK", "B200VEC_WF_BULK", "B200VEC_WF_TAIL")


MODE = os.environ.get("AB_MODE", "norm")          # norm: the one-launch step; scale: the max is given (no max pass, no grid barrier)
if MODE == "scale":
    from vector_b200._lib import BUILD_SCALE
    _, _, _m = vdist.build_stft_slab(eng, rows, plan, True, None, fs=FS, out_vec=vec, out_spec=spec)
    descs, nd = eng.make_descs(rows)


def run():
    if MODE == "scale":
        eng.build_stft1024(descs, nd, plan.lo, plan.hi, vec, BUILD_SCALE, plan.frame_lo, nf, spec, window="hann", fs=FS,
                           scaling="density")
        return vec, spec, eng._maxsq
    return vdist.build_stft_slab(eng, rows, plan, True, None, fs=FS, out_vec=vec, out_spec=spec)


def timeit(label, env):
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update(env)
    import time
    torch.cuda.synchronize()
    time.sleep(float(os.environ.get("AB_SLEEP", "1.5")))     # let the power average fall: every variant is timed as a burst
    vec.fill_(7.0); spec.fill_(7.0)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    reps = int(os.environ.get("AB_REPS", 400))
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _, _, m = run()
    for _ in range(reps):
        g.replay()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        g.replay()
    b.record(); b.synchronize()
    us = a.elapsed_time(b) / reps * 1e3
    ck = (int(vec.view(torch.int32).sum(dtype=torch.int64).item()), int(spec.view(torch.int32).sum(dtype=torch.int64).item()))
    print(f"{label:50s} {us:7.1f} us/step -> {(8*N + 4*1024*nf)/us/1e3:6.0f} GB/s  maxsq {m.item():.9g}  checksums {ck[0]} {ck[1]}  lib {os.environ.get('B200VEC_LIB', 'default')}", flush=True)
    return vec.clone(), spec.clone()


variants = [dict(kv.split("=") for kv in a.split(",") if kv) for a in sys.argv[1:]] or [{}, {"B200VEC_WF_BULK": "0"}]
ref = None
for env in variants:
    out = timeit(str(env), env)
    if ref is None:
        ref = out
    else:
        print("   identical to first variant:", torch.equal(torch.view_as_real(out[0]), torch.view_as_real(ref[0])),
              torch.equal(out[1], ref[1]), flush=True)

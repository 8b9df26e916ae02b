"""A/B timing of the fused build+STFT kernel variants on the BASELINE shape (cfg2, 56 M samples).
Usage: python scripts/fused_ab.py [ENV=VAL,ENV=VAL ...]   (each argument is one variant)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
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
spec_rows = eng.alloc_spectrogram(1024, nf)
spec_fm = eng.alloc_spectrogram(1024, nf, frame_major=True)
spec = spec_rows
descs, nd = eng.make_descs(rows)
KEYS = ("LAYOUT", "B200VEC_WF_TF", "B200VEC_WF_V", "B200VEC_WF_CHUNK", "B200VEC_WF_BULK", "B200VEC_WF_TAIL", "B200VEC_FUSED_W", "B200VEC_FUSED_V", "B200VEC_PIPE", "B200VEC_FUSED_TS", "B200VEC_FUSED_ZTMA")


def run():
    eng.build_stft1024(descs, nd, plan.lo, plan.hi, vec, BUILD_SCALE, plan.frame_lo, nf, spec,
                       window="hann", fs=FS, scaling="density")


def timeit(label, env):
    for k in KEYS:
        os.environ.pop(k, None)
    os.environ.update(env)
    global spec
    spec = spec_fm if env.get("LAYOUT") == "fm" else spec_rows
    eng._maxsq.zero_()
    eng.build_pass(descs, nd, plan.lo, plan.hi, None, BUILD_MAXONLY)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(15):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    med = ts[len(ts) // 2]
    print(f"{label:40s} min {ts[0]*1e3:7.1f} us  med {med*1e3:7.1f} us  -> {(8*N + 4*1024*nf)/med/1e6:6.0f} GB/s", flush=True)
    return vec.clone(), spec.clone()


variants = [dict(kv.split("=") for kv in a.split(",") if kv) for a in sys.argv[1:]] or [{}, {"LAYOUT": "fm"}]
ref = None
for env in variants:
    out = timeit(str(env), env)
    if ref is None:
        ref = out
    else:
        print("   identical to first variant:", torch.equal(torch.view_as_real(out[0]), torch.view_as_real(ref[0])),
              torch.equal(out[1], ref[1]), flush=True)
timeit("first again " + str(variants[0]), variants[0])

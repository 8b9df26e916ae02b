"""Where does the in-kernel pass A's time go?  Same box, same buffers:
(a) fused launch with the max supplied (SCALE), (b) one-launch form with NO ranges (cooperative launch + grid
barrier only), (c) one-launch form with the real max ranges, (d) with the whole vector as the range."""
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
spec = eng.alloc_spectrogram(1024, nf, frame_major=True)
descs, nd = eng.make_descs(rows, eng.free_flags(rows, N) if os.environ.get('NOFLAGS') != '1' else None)
print('free-instance flags:', eng.free_flags(rows, N))
ranges = eng.max_ranges(rows, N)
eng._maxsq.zero_(); eng.max_pass(descs, nd, ranges)
keep = eng._maxsq.clone()
variants = {"(a) SCALE, max supplied": dict(),
            "(b) one launch, no ranges": dict(norm_ranges=[]),
            "(c) one launch, period ranges": dict(norm_ranges=ranges),
            "(d) one launch, whole vector": dict(norm_ranges=[(0, N)])}
for label, kw in variants.items():
    def run():
        eng.build_stft1024(descs, nd, 0, N, vec, BUILD_SCALE, 0, nf, spec, fs=FS, maxsq=eng._maxsq if kw else keep, **kw)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    ts = []
    for _ in range(15):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    print(f"{label:34s} min {ts[0]*1e3:7.1f} us  med {ts[7]*1e3:7.1f} us   samples in ranges: {sum(h-l for l,h in kw.get('norm_ranges', []))}", flush=True)

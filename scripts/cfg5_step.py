"""The cfg5 batch step on one GPU (256 scenario vectors, prepare + one build launch): timing, for ncu captures."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vector_b200 import generate, synth
from vector_b200.engine import get_engine

FS, n = 56e6, 5_600_000
nv = int(os.environ.get("CFG5_VECTORS", 256))
eng = get_engine(0)
bank = synth.bank(FS)
scen = [synth.configs(synth.cfg5_rows(v)) for v in range(nv)]
packets = {c["file"]: bank[c["kind"]] for cfgs in scen for c in cfgs}
vecs, _, counts, maxsq = generate.build_batch_device(scen, n, True, packets=packets, engine=eng)
batch = eng._last_batch
out = torch.empty((nv, (n + 1) // 2 * 2), dtype=torch.complex64, device=eng.device)
del vecs
ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
for it in range(int(os.environ.get("CFG5_REPS", 6))):
    ev[0].record(); eng.batch_prepare(batch, FS); ev[1].record(); eng.batch_build(n, True, out=out, maxsq=maxsq); ev[2].record()
    torch.cuda.synchronize()
    print(f"prepare {ev[0].elapsed_time(ev[1]):.3f} ms  build {ev[1].elapsed_time(ev[2]):.3f} ms", flush=True)
for mode in (False,):
    ev[1].record(); eng.batch_build(n, mode, out=out, maxsq=maxsq); ev[2].record(); torch.cuda.synchronize()
    print(f"build without normalisation {ev[1].elapsed_time(ev[2]):.3f} ms", flush=True)

"""Where do the 12 us between the store-only build and the one-launch normalised build go?  cfg2, 56 M samples:
STORE (no max), SCALE (max given: scaled stores, ordinary launch), build_norm (max pass over the planned spans +
grid barrier + scaled stores, cooperative launch), MAXONLY alone.  Graph replay of 20 launches, CUDA events."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vector_b200 import generate, synth
from vector_b200.engine import get_engine
from vector_b200._lib import BUILD_STORE, BUILD_SCALE, BUILD_MAXONLY

FS, N = 56e6, 56_000_000
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)
vec = torch.empty(N, dtype=torch.complex64, device=eng.device)
descs0, nd = eng.make_descs(rows)
descs, _ = eng.make_descs(rows, eng.free_flags(rows, N))
ranges = eng.max_ranges(rows, N, 1, 0)
eng.build_norm(descs, nd, 0, N, vec, ranges)
torch.cuda.synchronize()
print("ranges", ranges, "maxsq", float(eng._maxsq.item()))
cases = {
    "STORE": lambda: eng.build_pass(descs0, nd, 0, N, vec, BUILD_STORE),
    "SCALE (max given)": lambda: eng.build_pass(descs0, nd, 0, N, vec, BUILD_SCALE),
    "build_norm, planned spans": lambda: eng.build_norm(descs, nd, 0, N, vec, ranges),
    "build_norm, whole vector as range, no free flags": lambda: eng.build_norm(descs0, nd, 0, N, vec, [(0, N)]),
    "build_norm, NO ranges (cooperative launch + barrier only)": lambda: eng.build_norm(descs, nd, 0, N, vec, []),
    "build_norm, first range only": lambda: eng.build_norm(descs, nd, 0, N, vec, ranges[:1]),
}
for rep in range(2):
    for name, run in cases.items():
        for _ in range(3):
            run()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(20):
                run()
        time.sleep(1.0)
        g.replay(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); b.synchronize()
        print(f"{name:52s} {a.elapsed_time(b) / 20 * 1e3:7.1f} us", flush=True)

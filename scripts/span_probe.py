"""What bounds the placement (span) kernel?  Store-only builds of 56 M samples with different packet coverage:
no instance at all (pure zero-block bulk stores), the cfg2 scenario (58 % of the samples under 1.46 instances),
every sample under exactly one instance, and a dense scenario.  CUDA events, graph replay of 20 launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vector_b200 import generate, synth
from vector_b200.engine import get_engine
from vector_b200._lib import BUILD_STORE

FS, N = 56e6, 56_000_000
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)
L = info[0][0]
vec = torch.empty(N, dtype=torch.complex64, device=eng.device)
cases = {
    "empty (no instance)": [(0, 10 ** 12, 0, 0)],
    "cfg2": rows,
    "one instance everywhere (period = L)": [(0, L, N // L, 0)],
    "two instances everywhere (period = L/2)": [(0, L // 2, (N - L) // (L // 2), 1)],
    "four instances everywhere": [(0, L // 4, (N - L) // (L // 4), 2)],
}
for name, rw in cases.items():
    descs, nd = eng.make_descs(rw)
    run = lambda: eng.build_pass(descs, nd, 0, N, vec, BUILD_STORE)
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20):
            run()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); b.record(); b.synchronize()
    us = a.elapsed_time(b) / 20 * 1e3
    print(f"{name:44s} {us:7.1f} us  {8 * N / us / 1e3:6.0f} GB/s written", flush=True)

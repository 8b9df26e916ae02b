"""A/B timing of the stand-alone full-rate STFT (cfg3 full): frame-major warp-per-frame kernel vs C-order tile kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vector_b200.engine import get_engine
N = 56_000_000
eng = get_engine(0)
g = torch.Generator(device="cuda").manual_seed(1)
x = torch.view_as_complex(torch.randn(N, 2, device="cuda", generator=g))
nf = (N - 1024) // 512 + 1
outs = {"frame-major (wf1024 PLAIN)": eng.alloc_spectrogram(1024, nf, frame_major=True),
        "C-order rows (tile kernel)": eng.alloc_spectrogram(1024, nf)}
for label, out in outs.items():
    for _ in range(3):
        eng.stft(x, "hann", 1024, 512, 1024, 56e6, out=out)
    torch.cuda.synchronize()
    ts = []
    for _ in range(15):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eng.stft(x, "hann", 1024, 512, 1024, 56e6, out=out); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    med = ts[len(ts) // 2]
    print(f"{label:32s} min {ts[0]*1e3:7.1f} us  med {med*1e3:7.1f} us  -> {(8*N + 4*1024*nf)/med/1e6:6.0f} GB/s (8 B/Sa read + 8 B/Sa written)", flush=True)
vals = list(outs.values())
print("identical:", torch.equal(vals[0], vals[1]))

"""A/B timing of the STFT kernel variants on the BASELINE shape (56 M samples, hann 1024/512)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vector_b200 import generate, synth
from vector_b200.engine import get_engine

eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS)
vec, _, _, _ = generate.build_vector_device(cfgs, 56_000_000, True, packets=pk, engine=eng)
nf = (56_000_000 - 1024) // 512 + 1
outs = {"padded": eng.alloc_spectrogram(1024, nf), "dense": torch.empty((1024, nf), dtype=torch.float32, device="cuda")}


def timeit(label, env, out):
    for k in ("B200VEC_STFT_W", "B200VEC_STFT_TMA", "B200VEC_STFT1024"):
        os.environ.pop(k, None)
    os.environ.update(env)
    for _ in range(3):
        eng.stft(vec, "hann", 1024, 512, 1024, 56e6, out=out)
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); eng.stft(vec, "hann", 1024, 512, 1024, 56e6, out=out); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    print(f"{label:40s} min {ts[0]*1e3:7.1f} us  med {ts[len(ts)//2]*1e3:7.1f} us  -> {896e6/ts[len(ts)//2]/1e6:6.0f} GB/s", flush=True)


for oname, out in outs.items():
    timeit(f"tma W=16 x1 CTA/SM [{oname}]", {}, out)
    timeit(f"tma W=8 x2 CTA/SM  [{oname}]", {"B200VEC_STFT_W": "8"}, out)
    timeit(f"gather 8 warps x2  [{oname}]", {"B200VEC_STFT_TMA": "0"}, out)
ref = outs["dense"].clone()
timeit("tma W=16 again", {}, outs["padded"])
print("padded == dense:", torch.equal(outs["padded"], ref))

"""Timing of every kernel family of the path OTHER than the bench step, at the BASELINE size (cfg2,
56 M samples / 109 374 x 1024 spectrogram): algorithmic bytes / device time, fraction of the measured
HBM peak.  CUDA events on the current stream, 3 warm-ups, median of 9.  One JSON object per line."""
import json, os, sys, contextlib
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from vector_b200 import api, generate, synth
from vector_b200.engine import get_engine
from vector_b200._lib import BUILD_STORE

FS, N = 56e6, 56_000_000
PEAK = 6536.4
try:
    PEAK = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
eng = get_engine(0)
cfgs, pk = synth.scenario(synth.CFG2_ROWS, FS)
info = generate.prepare_bank(eng, cfgs, pk, FS)
rows = generate.placement_rows(cfgs, info, N, FS)


def timed(name, fn, alg_bytes, note=""):
    """Device time per call.  Calls that only launch kernels are captured in a CUDA graph (10 calls per replay) so that
    launch gaps do not count; calls that synchronise (host read-backs) are timed eagerly, host round trips included."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    how, per = "graph replay of 10 calls", 10
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(per):
                fn()
        run = g.replay
    except Exception:
        torch.cuda.synchronize()
        how, per, run = "eager, host round trips included", 1, fn
    run(); torch.cuda.synchronize()
    ts = []
    for _ in range(9):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); run(); b.record(); b.synchronize()
        ts.append(a.elapsed_time(b) / per)
    ts.sort()
    ms = ts[len(ts) // 2]
    gbs = alg_bytes / ms / 1e6
    print(json.dumps({"kernel": name, "ms": round(ms, 4), "alg_bytes": alg_bytes, "GBps": round(gbs, 1),
                      "frac_of_measured_peak": round(gbs / PEAK, 3), "timing": how, "note": note}), flush=True)


vec = torch.empty(N, dtype=torch.complex64, device=eng.device)
descs, nd = eng.make_descs(rows)
timed("K2 place_kernel, store only (a6)", lambda: eng.build_pass(descs, nd, 0, N, vec, BUILD_STORE), 8 * N,
      "8 B/Sa written, nothing read")
timed("K2+K3 normalised build, one launch (a6+a7)", lambda: eng.build_vector(rows, N, True, out=vec), 8 * N,
      "max pass over the planned spans + grid barrier + scaled store pass (b200vec_build_norm)")
v, _ = eng.build_vector(rows, N, True, out=vec)

nf = (N - 1024) // 512 + 1
spec = eng.alloc_spectrogram(1024, nf, frame_major=True)
timed("K4 wf1024 plain STFT 1024/512 of an existing vector (a9)",
      lambda: eng.stft(v, "hann", 1024, 512, 1024, FS, out=spec), 8 * N + 4 * 1024 * nf, "8 B/Sa read + 8 B/Sa written")
spec_c = eng.alloc_spectrogram(1024, nf, frame_major=False)
timed("K4 stft1024 tile kernel, C-order rows (a9)",
      lambda: eng.stft(v, "hann", 1024, 512, 1024, FS, out=spec_c), 8 * N + 4 * 1024 * nf, "same bytes, [1024, frames] memory")
del spec_c

def preview():
    with contextlib.redirect_stdout(sys.stderr):          # the reference's own progress prints (utils.py:186-200)
        return api.create_spectrogram_device(v, FS, time_resolution_us=1, engine=eng)


f, t, S = preview()
prev_bytes = 32 * (N // 56) + 4 * S.numel()
timed("K4 generic, create_spectrogram preview (a8: decimate x56, hann 512/412, nfft 1024)",
      preview, prev_bytes,
      f"32 B sector per decimated sample + {tuple(S.shape)} float32 out; includes the all-zero flag read-back")

timed("K5 normalize_spectrogram on the preview (a10)", lambda: eng.normalize_spectrogram(S, 5, 95, 25),
      8 * S.numel(), "4 B read + 4 B written per element (dB pass); the radix-select passes re-read")
Sf = spec.t().contiguous() if spec.stride(0) == 1 else spec
timed("K5 normalize_spectrogram on the full-rate spectrogram (a10)", lambda: eng.normalize_spectrogram(Sf, 5, 95, 25),
      8 * Sf.numel(), "112 M elements; 3 exact order statistics on |S|>0 and 6 on the dB image")
del Sf

q = torch.empty(2 * N, dtype=torch.int16, device=eng.device)
timed("K6 wv_quantize (a12)", lambda: eng.wv_quantize(v, out=q), 12 * N, "8 B read + 4 B written per sample")
pl = eng.split_planes(v)
timed("K7 split_planes (a11)", lambda: eng.split_planes(v), 16 * N, "8 B read + 8 B written per sample (allocates its output)")
del q, pl

pkt = v[:4_000_000].contiguous()
timed("K8+K9 detect_packet_bounds, 4 M samples (f4)", lambda: api.detect_packet_bounds(pkt, FS), 24 * pkt.numel(),
      "8 in + 8 out (box energy), 8 in (threshold); includes the median's select and host read-backs")
db, _, _ = eng.normalize_spectrogram(S, 5, 95, 25)
timed("K10 median_filter_rows2 on the preview (f4)", lambda: api.median_filter_rows2(db, engine=eng), 8 * db.numel(),
      "4 B read + 4 B written per element")

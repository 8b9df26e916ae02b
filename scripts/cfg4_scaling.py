"""SURVEY.md §8d cfg4 at FULL size, strong scaling: 10 s @ 200 MSa/s (2e9 samples, 16 GB vector + 16 GB
full-rate spectrogram in total), six packets made at 56 MSa/s and resampled to 200 MSa/s on the device
(K0), dense periods 2.0/1.5/1.0/0.8/0.5/0.4 ms, time-sharded over the ranks of this torchrun launch.
One fused launch per rank and step (in-kernel pass A + peer-memory MAX at N > 1).
Checks (size-independent): two 2 M-sample windows of the slab equal the two-pass placement kernels
bit for bit; Parseval per frame on a few frames.  Prints one JSON line on rank 0."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, scipy.signal, torch, torch.distributed as dist
from vector_b200 import dist as vdist, generate, synth
from vector_b200.engine import get_engine
from vector_b200._lib import BUILD_SCALE

world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
eng = get_engine(local)
FS = 200e6
SECONDS = float(os.environ.get("CFG4_SECONDS", "10"))
N = int(SECONDS * FS)
cfgs, pk = synth.scenario(synth.CFG4_ROWS, 56e6)
for c in cfgs:
    c["sample_rate"] = 56e6
info = generate.prepare_bank(eng, cfgs, pk, FS, resample=True)
rows = generate.placement_rows(cfgs, info, N, FS)
plan = vdist.plan_slab(N, world, rank, 1024, 512, 1, align=512)
n_own, nf = plan.hi - plan.lo, plan.frame_hi - plan.frame_lo
vec = torch.empty(n_own, dtype=torch.complex64, device=eng.device)
spec = eng.alloc_spectrogram(1024, nf, frame_major=True)
peer = vdist.PeerMax() if world > 1 else None


def step():
    return vdist.build_stft_slab(eng, rows, plan, True, None, fs=FS, out_vec=vec, out_spec=spec, peer=peer)


for _ in range(2):
    _, _, m = step()
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
reps = 5
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(reps):
    step()
b.record(); torch.cuda.synchronize()
t = torch.tensor([a.elapsed_time(b) / reps], dtype=torch.float64, device=eng.device)
if world > 1:
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
ms = t.item()

# ---- checks
ok = True
descs, nd = eng.make_descs(rows)
for off in (0, max(0, n_own - 2_000_000)):
    lo, hi = plan.lo + off, min(plan.lo + off + 2_000_000, plan.hi)
    ref = torch.empty(hi - lo, dtype=torch.complex64, device=eng.device)
    eng.build_pass(descs, nd, lo, hi, ref, BUILD_SCALE, maxsq=m)
    ok = ok and torch.equal(ref, vec[off:off + (hi - lo)])
w = torch.from_numpy(scipy.signal.get_window("hann", 1024)).to(eng.device)
scale = 1.0 / (FS * float((w * w).sum()))
for f in (0, nf // 3, nf - 1):
    seg = vec[f * 512: f * 512 + 1024]
    if seg.numel() < 1024:
        continue
    want = float(((seg.to(torch.complex128) * w).abs() ** 2).sum()) * 1024 * scale
    got = float(spec[:, f].double().sum())
    ok = ok and abs(got - want) <= 1e-5 * max(want, 1e-30)
flag = torch.tensor([1 if ok else 0], device=eng.device)
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    bytes_rank = 8 * n_own + 4 * 1024 * nf
    print(json.dumps({"workload": f"cfg4: {SECONDS:g} s @ 200 MSa/s, 6 resampled packets, dense periods, strong scaling",
                      "n_gpus": world, "total_samples": N, "instances": [r[2] for r in rows],
                      "ms_per_step": ms, "GSa_per_s": N / ms / 1e6, "per_gpu_alg_GBps": bytes_rank / ms / 1e6,
                      "max_abs": float(np.sqrt(m.item())), "checks_ok": bool(flag.item())}), flush=True)
if world > 1:
    dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)

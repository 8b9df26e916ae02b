#!/usr/bin/env python
"""bench.py — headline benchmark of the vector-build hot path (BASELINE.json metric:
"output GSa/s for 6-pkt 1 s@56 MSa/s vector build + STFT; % HBM peak @1/2/4/8 GPU").

One step = one pass of the hot path over one batch of synthetic input:
    pass A (max |v|^2 over one common period of the packet trains + head/tail, no stores)
    -> [all-reduce MAX of one double when N > 1] ->
    ONE fused kernel: recompute, scale, store complex64 AND the full-rate STFT (hann 1024/512,
    nfft 1024, 'density', fftshifted float32 [1024, frames] handed out as the transpose view of
    frame-major memory) — every warp builds, transforms and stores its own frames
on SURVEY.md §8d cfg2 (6 packets, periods 100/50/20/10/5/2 ms, shifts +2/-3/+1/-8/+6/-12 MHz,
normalisation on).  At N GPUs the vector is N x 1000 ms long and time-sharded, 56 M samples
(+ halo) per GPU: per-GPU work is fixed -> "scaling": "weak".  value = all samples / max-over-
ranks device time; `e2e` is the same work through the host-buffer API with the H2D of the
packets and the D2H of vector + spectrogram inside the timed region.

`--impl reference` times the reference's own CPU implementation (numpy/scipy, via the oracle
port that is pinned bit-exact to it) on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np

FS = 56e6
N_PER_GPU = 56_000_000
METRIC = "output GSa/s, 6-packet 1000 ms @56 MSa/s normalised vector build + full-rate STFT"
WORKLOAD = ("cfg2+cfg3(full): 6 synthetic packets (tone/chirp/BPSK/noise/multitone/tone), periods "
            "100/50/20/10/5/2 ms, shifts +2/-3/+1/-8/+6/-12 MHz, 1000 ms per GPU @56 MSa/s, max-abs "
            "normalisation, then hann 1024/512 nfft-1024 two-sided density STFT")


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback (of fallback)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def workload(n_gpus):
    from vector_b200 import synth
    cfgs, packets = synth.scenario(synth.CFG2_ROWS, FS)
    return cfgs, packets, N_PER_GPU * n_gpus


# ----------------------------------------------------------------------------- reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_port as O
    cfgs, packets, _ = workload(1)
    # bounded sample: 1000 ms of vector (the per-GPU share, ~2.5 s of CPU per step); shortened to
    # 200 ms when K+W would push the run past a few minutes
    sample_ms = 1000 if (args.steps + args.warmup) <= 40 else 200
    n = N_PER_GPU * sample_ms // 1000

    def step():
        vec = O.build_vector(packets, cfgs, n, True)[0]
        O.full_rate_spectrogram(vec, FS)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = (time.perf_counter() - t0) / args.steps
    val = n / dt / 1e9
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "GSa/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "dtype_note": "complex64 samples, float32 arithmetic; packet phase in fp64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": f"{sample_ms} ms of vector ({n} samples) per step"},
            "cpu_baseline": {"value": val, "unit": "GSa/s", "cores": 1, "kind": "port",
                             "host_cores": os.cpu_count(),
                             "sample": "oracle port (numpy %s / scipy %s): build_vector + scipy.signal.spectrogram "
                                       "1024/512 on %d samples per step; numpy ufuncs and pocketfft are single-threaded"
                                       % (np.__version__, __import__("scipy").__version__, n)},
            "e2e": {"value": val, "unit": "GSa/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from vector_b200 import dist as vdist, generate, hostmath as H
    from vector_b200.engine import get_engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly one JSON line: libraries that print there (the NCCL version banner) go to stderr
    sys.stdout.flush()
    json_out = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")      # stdout carries exactly one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    eng = get_engine(local)
    cfgs, packets, total = workload(world)
    reduce_max = vdist.make_max_reducer()
    peer = None
    if world > 1 and not args.nccl_max and not args.rows:
        try:                                   # symmetric 64-byte blocks for the in-kernel MAX exchange over NVLink
            peer = vdist.PeerMax()
        except Exception as exc:
            print(f"[bench] symmetric memory unavailable ({type(exc).__name__}: {exc}); NCCL all-reduce form",
                  file=sys.stderr)
            peer = None

    info = generate.prepare_bank(eng, cfgs, packets, FS)
    rows = generate.placement_rows(cfgs, info, total, FS)
    plan = vdist.plan_slab(total, world, rank, 1024, 512, 1, align=512)     # slabs cut on hop boundaries
    n_own = plan.hi - plan.lo
    n_frames = plan.frame_hi - plan.frame_lo
    vec = torch.empty(n_own, dtype=torch.complex64, device=eng.device)
    spec = eng.alloc_spectrogram(1024, n_frames, frame_major=not args.rows)   # [1024, frames] view of [frames, 1024] memory
    descs, nd = eng.make_descs(rows, eng.free_flags(rows, total))   # hint bit for the in-kernel max pass
    max_ranges = eng.max_ranges(rows, total, world, rank)      # this rank's share of one common period + head/tail
    from vector_b200._lib import BUILD_MAXONLY, BUILD_SCALE

    ev = lambda: torch.cuda.Event(enable_timing=True)
    marks = []

    one_launch = (world == 1 and not args.rows and not args.two_launch) or peer is not None

    def step(record=None):
        e = [ev() for _ in range(3)] if record is not None else None
        if one_launch:
            # single GPU: pass A (max |v|^2 over one common period + head/tail) runs INSIDE the fused kernel
            if e: e[0].record(); e[1].record()
            eng.build_stft1024(descs, nd, plan.lo, plan.hi, vec, BUILD_SCALE, plan.frame_lo, n_frames, spec,
                               window="hann", fs=FS, scaling="density", norm_ranges=max_ranges, peer=peer)
        else:
            eng._maxsq.zero_()
            if e: e[0].record()
            eng.max_pass(descs, nd, max_ranges)                                    # pass A: max |v|^2, no stores
            if reduce_max is not None:
                reduce_max(eng._maxsq)                                              # the one collective: 1 double, MAX
            if e: e[1].record()
            eng.build_stft1024(descs, nd, plan.lo, plan.hi, vec, BUILD_SCALE, plan.frame_lo, n_frames, spec,
                               window="hann", fs=FS, scaling="density")            # place + scale + store + STFT
        if e:
            e[2].record()
            record.append(e)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()
    # The step is 1 launch at N=1 and memset + 2 launches + 1 all-reduce + 1 launch at N>1: capture it in a
    # CUDA graph so that the timed loop replays device work back to back (no Python / launch gaps).
    graph = None
    launches_per_step = None
    if not args.no_graph and (world == 1 or peer is not None):   # the NCCL form stays eager (an all-reduce inside a captured step hung on this stack)
        try:
            l0 = eng.launches
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                step(); step()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            l0 = eng.launches
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                step()
            launches_per_step = eng.launches - l0
            g.replay()
            torch.cuda.synchronize()
            graph = g
        except Exception as exc:                                   # capture unsupported: time the eager loop
            print(f"[bench] CUDA graph capture failed ({type(exc).__name__}: {exc}); eager loop", file=sys.stderr)
            graph = None
            torch.cuda.synchronize()
    barrier()
    # Clocks: the timed region is a few milliseconds, far below nvidia-smi's sampling period, so the SAME step
    # keeps the GPU busy for ~0.5 s before it and ~0.3 s after it (fixed iteration counts: identical on every
    # rank) while rank 0 samples clocks and throttle reasons every 100 ms.
    def run_one():
        if graph is not None:
            graph.replay()
        else:
            step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    for _ in range(args.load_steps):
        run_one()
    barrier()
    launches0 = eng.launches
    t_start, t_end = ev(), ev()
    t_start.record()
    for _ in range(args.steps):
        if graph is not None:
            graph.replay()
        else:
            step(marks)
    t_end.record()
    barrier()
    ms_total = t_start.elapsed_time(t_end)
    for _ in range(args.load_steps * 3 // 5):
        run_one()
    barrier()
    launches = launches_per_step * args.steps if graph is not None else eng.launches - launches0
    if graph is not None:                                          # per-kernel split: a few eager steps after the timed region
        for _ in range(min(args.steps, 10)):
            step(marks)
        barrier()
    clocks = sampler.stop() if rank == 0 else None
    if clocks is not None:
        clocks["window"] = (f"{args.load_steps} untimed steps before + the timed region + {args.load_steps * 3 // 5} after, "
                            "GPU continuously busy with the same step; nvidia-smi every 100 ms")
    ms_t = torch.tensor([ms_total], dtype=torch.float64, device=eng.device)
    if world > 1:
        dist.all_reduce(ms_t, op=dist.ReduceOp.MAX)
    ms_step = ms_t.item() / args.steps
    k_ms = np.array([[m[i].elapsed_time(m[i + 1]) for i in range(2)] for m in marks]).mean(axis=0)
    k_src = "CUDA events around each launch in the timed loop"
    if graph is not None and one_launch:
        # the step IS one launch of the dominant kernel: its average duration over the timed region is the
        # event-timed region / K (launch gaps included, so this is the conservative figure)
        k_ms = np.array([0.0, ms_step])
        k_src = "timed region / K: the step is exactly one launch of this kernel (CUDA graph replay)"

    # ---- e2e through the host-buffer API: H2D packets, D2H vector + spectrogram, every step
    pinned_pk = [(torch.from_numpy(np.ascontiguousarray(y)).pin_memory(), pre) for y, pre in packets]
    out_vec = torch.empty(n_own, dtype=torch.complex64).pin_memory()
    spec_mem = spec.t() if spec.stride(0) == 1 else spec               # the contiguous side of the buffer
    out_spec = torch.empty(spec_mem.shape, dtype=torch.float32).pin_memory()

    def e2e_step():
        pk = [(t.numpy(), pre) for t, pre in pinned_pk]
        generate.prepare_bank(eng, cfgs, pk, FS)                          # H2D of every packet + shift
        v, s, _ = vdist.build_stft_slab(eng, rows, plan, True, reduce_max if peer is None else None, fs=FS,
                                        out_vec=vec, out_spec=spec, peer=peer)
        out_vec.copy_(v, non_blocking=True)
        out_spec.copy_(s.t() if s.stride(0) == 1 else s, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(out_spec.flatten()[512]) + float(out_vec[0].real)   # read the result on the host

    e2e_steps = max(2, min(args.steps, 5))
    e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = torch.tensor([(time.perf_counter() - t0) * 1e3 / e2e_steps], dtype=torch.float64, device=eng.device)
    if world > 1:
        dist.all_reduce(e2e_ms, op=dist.ReduceOp.MAX)
    h2d = sum(t.numel() * 8 for t, _ in pinned_pk)
    d2h = out_vec.numel() * 8 + out_spec.numel() * 4

    if rank == 0:
        peak, peak_src = peaks()
        fused_bytes = 8 * n_own + 4 * spec.numel()
        kernels = {
            "place_max(passA)": {"ms": k_ms[0], "alg_bytes": 0, "in_fused_kernel": bool(one_launch),
                                 "samples_evaluated": int(sum(hi - lo for lo, hi in max_ranges)),
                                 "note": "no HBM output; L2-resident bank reads only; the periodic packet trains repeat "
                                         "bit-exactly, so one common period + head/tail of the vector gives the exact max" + (
                                     "; includes the 1-double all-reduce" if world > 1 else "")},
            "build_stft1024(fused passB+STFT)": {"ms": k_ms[1], "alg_bytes": fused_bytes,
                                                 "note": "8 B/Sa vector store + 8 B/Sa spectrogram store, nothing read" + (
                                                     "; includes pass A (max over one common period + head/tail) and its grid barrier" if one_launch else "")},
        }
        for k in kernels.values():
            k["GBps"] = k["alg_bytes"] / (k["ms"] * 1e-3) / 1e9 if k["ms"] > 0 else None
        traffic = None
        try:                                                   # per-launch DRAM bytes of the dominant kernel, from the committed ncu capture
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as fh:
                t = json.load(fh)["wf1024_kernel"]
            if not args.rows and world == 1:
                traffic = t["dram_bytes_read"] + t["dram_bytes_write"]
        except Exception:
            traffic = None
        dom = "build_stft1024(fused passB+STFT)"
        kname = "tile1024_kernel<1,16,0> (C-order rows)" if args.rows else "wf1024_kernel (warp per frame, frame-major memory)"
        ach = kernels[dom]["GBps"]
        line = {
            "metric": METRIC, "value": total / (ms_step * 1e-3) / 1e9, "unit": "GSa/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "dtype_note": "complex64 samples, float32 arithmetic; packet phase in fp64", "data": "synthetic",
            "config": {"workload": WORKLOAD, "samples_per_gpu": N_PER_GPU, "total_samples": total,
                       "parallelism": (f"time-slab x{world}, ONE launch per GPU: MAX of one double exchanged inside the kernel over "
                                       f"NVLink peer memory (system-scope atomics, torch symmetric memory)" if peer is not None else
                                       f"time-slab x{world}, pass A launch + 1 scalar NCCL all-reduce(MAX) + fused launch")
                       if world > 1 else "single GPU",
                       "l2": "no flush: each step writes 896 MB per GPU (vector + spectrogram), far above the 126 MB L2",
                       "timed_loop": "CUDA graph replay of one step, K times" if graph is not None else "eager launches"},
            "roofline": {"bound": "hbm", "kernel": dom, "cuda_kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "frac_of_nominal_8000": ach / 8000.0, "traffic": traffic,
                         "traffic_source": "profiles/ncu_traffic.json (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, one launch)" if traffic else None,
                         "alg_bytes_per_launch": fused_bytes,
                         "peak_source": peak_src, "kernel_time_source": k_src,
                         "step_frac": fused_bytes / (ms_step * 1e-3) / 1e9 / peak},
            "kernels": kernels,
            "e2e": {"value": total / (e2e_ms.item() * 1e-3) / 1e9, "unit": "GSa/s", "ms_per_step": e2e_ms.item(),
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                    "api": "generate.prepare_bank + dist.build_stft_slab, pinned host buffers"},
            "gpu_launches": launches, "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline()
        print(json.dumps(line), file=json_out, flush=True)
    if world > 1:
        dist.destroy_process_group()


def cpu_baseline():
    """The oracle port (the reference's numpy/scipy calls) on the host cores: a bounded sample
    of the same workload — one 1000 ms cfg2 build + its full-rate STFT, best of 2."""
    from oracle import ref_port as O
    cfgs, packets, _ = workload(1)
    best = None
    parts = None
    for _ in range(2):
        t0 = time.perf_counter()
        vec = O.build_vector(packets, cfgs, N_PER_GPU, True)[0]
        t1 = time.perf_counter()
        O.full_rate_spectrogram(vec, FS)
        t2 = time.perf_counter()
        if best is None or t2 - t0 < best:
            best, parts = t2 - t0, (t1 - t0, t2 - t1)
    return {"value": N_PER_GPU / best / 1e9, "unit": "GSa/s", "cores": 1, "host_cores": os.cpu_count(),
            "kind": "port", "build_s": parts[0], "stft_s": parts[1],
            "sample": "full cfg2 workload once (56 M samples): oracle build_vector + scipy.signal.spectrogram 1024/512, "
                      "best of 2; numpy %s scipy %s, single-threaded like the reference" % (
                          np.__version__, __import__("scipy").__version__)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--load-steps", type=int, default=2500,
                    help="untimed steps run around the timed region so that nvidia-smi can sample clocks under load")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of CUDA-graph replays")
    ap.add_argument("--nccl-max", action="store_true",
                    help="A/B at N>1: pass A launch + NCCL all-reduce(MAX) + fused launch instead of the in-kernel peer-memory MAX")
    ap.add_argument("--two-launch", action="store_true",
                    help="A/B at N=1: pass A as its own launch (the multi-GPU form) instead of inside the fused kernel")
    ap.add_argument("--rows", action="store_true",
                    help="A/B: C-order [1024, frames] spectrogram memory (16-frame tile kernel) instead of "
                         "frame-major memory (warp-per-frame kernel)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

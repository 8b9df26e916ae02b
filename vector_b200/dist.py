"""Multi-GPU form of the path: contiguous time-slab sharding, one process per GPU.

The composite vector is a pure function of (packet bank, integer descriptors, sample index),
so rank g fills samples [lo_g, hi_g) with no data-path communication; the packet bank (a few
MB) is replicated.  The only exchange is the normalisation scalar: one double (max |v|^2)
all-reduced with MAX between pass A and pass B — NCCL over NVLink/NVSwitch on GPUs, gloo in
the CPU tests.  STFT frames belong to the rank that holds their first sample; the
<= (nperseg-1)*stride trailing halo is recomputed locally by the placement kernel rather than
exchanged.  The reference has no distributed code at all; this module is the
"long vector" answer to what it handles by decimation (utils.py:184-200).
"""
from __future__ import annotations

from dataclasses import dataclass

from . import hostmath as H


@dataclass(frozen=True)
class SlabPlan:
    rank: int
    world: int
    total: int
    lo: int               # first owned sample
    hi: int               # one past the last owned sample
    frame_lo: int         # first owned STFT frame
    frame_hi: int
    halo_hi: int          # one past the last sample the owned frames read (>= hi)
    n_frames_total: int


def plan_slab(total, world, rank, nperseg=1024, hop=512, stride=1, align=16):
    lo, hi = H.slab_bounds(total, world, rank, align)
    n_dec = (total + stride - 1) // stride
    nf = 0 if n_dec < nperseg else (n_dec - nperseg) // hop + 1
    f_lo, f_hi, halo = H.frame_range_for_slab(lo, hi, total, stride, nperseg, hop, nf)
    return SlabPlan(rank, world, total, lo, hi, f_lo, f_hi, halo, nf)


def make_max_reducer(group=None):
    """all-reduce(MAX) of the one-element max|v|^2 tensor; None when not distributed."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return None

    def reduce_max(t):
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return reduce_max


class PeerMax:
    """The 64-byte symmetric blocks of the in-kernel MAX exchange (b200vec_build_stft1024_fm_norm_peer):
    one per rank, allocated with torch symmetric memory so that every rank holds a mapping of every
    other rank's block (NVLink peer memory).  One epoch per fused step, the same on every rank."""

    def __init__(self, group=None, blocks=None):
        import torch
        if blocks is not None:                     # tests: explicit (ptrs, world, rank), e.g. one local block
            self.ptrs, self.world, self.rank = blocks
            self.epoch = 0
            return
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        grp = group if group is not None else dist.group.WORLD
        self.world, self.rank = dist.get_world_size(grp), dist.get_rank(grp)
        if self.world > 8:
            raise ValueError("PeerMax supports up to 8 ranks (one NVSwitch domain)")
        dev = torch.device("cuda", torch.cuda.current_device())
        self.buf = symm.empty(16, dtype=torch.int32, device=dev)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, grp)
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.epoch = 0
        torch.cuda.synchronize()
        dist.barrier(grp)                          # every block is zero before any kernel touches a peer

    def next_epoch(self):
        """Host-side step count (informational).  The kernel is passed epoch 0 = "count on the device"
        (u32 at byte 24 of the home block), so a captured step can be replayed from a CUDA graph."""
        self.epoch += 1
        return 0


def build_slab(engine, rows, plan, normalize=True, reduce_max=None, out=None):
    """Owned slab plus halo of the (optionally normalised) vector on this rank's GPU.
    The max pass covers this rank's share of hostmath.max_eval_ranges (one common period of the
    packet trains plus the head and tail of the vector — enough for the exact global max), the
    store pass also fills the halo.  Returns (tensor [halo_hi - lo], maxsq or None)."""
    import torch
    from ._lib import BUILD_MAXONLY, BUILD_SCALE, BUILD_STORE
    n = plan.halo_hi - plan.lo
    if out is None:
        out = torch.empty(max(n, 0), dtype=torch.complex64, device=engine.device)
    descs, nd = engine.make_descs(rows)
    if not normalize:
        engine.build_pass(descs, nd, plan.lo, plan.halo_hi, out, BUILD_STORE)
        return out, None
    engine._maxsq.zero_()
    engine.max_pass(descs, nd, engine.max_ranges(rows, plan.total, plan.world, plan.rank))
    if reduce_max is not None:
        reduce_max(engine._maxsq)
    engine.build_pass(descs, nd, plan.lo, plan.halo_hi, out, BUILD_SCALE)
    return out, engine._maxsq


def stft_slab(engine, vec, plan, window="hann", nperseg=1024, noverlap=512, nfft=1024, fs=56e6,
              scaling="density", stride=1, out=None):
    """This rank's columns [frame_lo, frame_hi) of the global spectrogram, computed from its
    slab+halo tensor.  Columns concatenate in rank order."""
    return engine.stft(vec, window, nperseg, noverlap, nfft, fs, scaling=scaling, stride=stride,
                       in_origin=plan.lo, frame_lo=plan.frame_lo,
                       n_frames=plan.frame_hi - plan.frame_lo, total_len=plan.total, out=out)


def build_stft_slab(engine, rows, plan, normalize=True, reduce_max=None, fs=56e6, window="hann",
                    scaling="density", out_vec=None, out_spec=None, peer=None):
    """Owned slab of the vector AND this rank's spectrogram columns (hann-1024/512 analysis STFT of
    vector_analyzer/spectrogram_analysis.py:35-49) with the fused kernel: pass A (max, no stores)
    -> [all-reduce MAX] -> one kernel that places, scales, stores the vector and transforms it
    out of shared memory.  `plan` must be cut on hop boundaries: plan_slab(..., align=512).
    Returns (vector [hi - lo], spectrogram [1024, frame_hi - frame_lo], maxsq or None)."""
    import torch
    from ._lib import BUILD_MAXONLY, BUILD_SCALE, BUILD_STORE
    nf = plan.frame_hi - plan.frame_lo
    frame_lo = plan.frame_lo if nf > 0 else plan.lo // 512          # a slab past the last frame has none
    if plan.lo != frame_lo * 512:
        raise ValueError("build_stft_slab needs slabs cut on hop boundaries: plan_slab(..., align=512)")
    n = plan.hi - plan.lo
    if out_vec is None:
        out_vec = torch.empty(max(n, 0), dtype=torch.complex64, device=engine.device)
    if out_spec is None:
        out_spec = engine.alloc_spectrogram(1024, nf, frame_major=True)
    frame_major = out_spec.dim() == 2 and out_spec.stride(0) == 1 and out_spec.stride(1) >= 1024
    in_kernel_max = normalize and frame_major and (peer is not None or (plan.world == 1 and reduce_max is None))
    descs, nd = engine.make_descs(rows, engine.free_flags(rows, plan.total) if in_kernel_max else None)
    mode, norm_ranges = BUILD_STORE, None
    if normalize and peer is not None and frame_major:
        # multi-GPU, ONE launch per rank: this rank's share of pass A and the MAX exchange over NVLink
        # peer memory both run inside the fused kernel (no NCCL call in the step)
        norm_ranges = engine.max_ranges(rows, plan.total, plan.world, plan.rank)
        if len(norm_ranges) > 4:
            raise ValueError("too many max ranges for one launch")
    elif normalize and plan.world == 1 and reduce_max is None and frame_major:
        # single GPU: pass A runs inside the fused kernel (grid barrier), ONE launch for the step
        norm_ranges = engine.max_ranges(rows, plan.total, 1, 0)
        if len(norm_ranges) > 4:
            norm_ranges = [(0, plan.total)]
    elif normalize:
        engine._maxsq.zero_()
        engine.max_pass(descs, nd, engine.max_ranges(rows, plan.total, plan.world, plan.rank))
        if reduce_max is not None:
            reduce_max(engine._maxsq)
        mode = BUILD_SCALE
    engine.build_stft1024(descs, nd, plan.lo, plan.hi, out_vec, mode, frame_lo, nf, out_spec,
                          window=window, fs=fs, scaling=scaling, norm_ranges=norm_ranges,
                          peer=peer if norm_ranges is not None else None)
    return out_vec, out_spec, (engine._maxsq if normalize else None)

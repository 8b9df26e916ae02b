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


TILE_SLOT = 62                 # bank slot the periodic tile is attached to (the vector build uses 0..31)
SPEC_TILE_BYTES = 64 << 20     # largest spectrogram period (frames * 4 KB) that is tiled: it has to stay in the 126 MB L2
REDUNDANT_MAX_SAMPLES = 16_000_000


def build_stft_slab(engine, rows, plan, normalize=True, reduce_max=None, fs=56e6, window="hann",
                    scaling="density", out_vec=None, out_spec=None, peer=None, tile="auto",
                    redundant_max=REDUNDANT_MAX_SAMPLES, spec_tile=True):
    """Owned slab of the vector AND this rank's spectrogram columns (hann-1024/512 analysis STFT of
    vector_analyzer/spectrogram_analysis.py:35-49) with the fused kernel: one launch that finds the
    max, places, scales, stores the vector and transforms it out of registers.  `plan` must be cut on
    hop boundaries: plan_slab(..., align=512).
    Returns (vector [hi - lo], spectrogram [1024, frame_hi - frame_lo], maxsq or None).

    Normalisation across ranks (frame-major memory):
      * the ranges that hold the global max (one common period of the packet trains + head and tail,
        hostmath.max_eval_ranges) are short (<= redundant_max samples): EVERY rank evaluates all of them
        inside its own launch — no exchange, no skew between GPUs, the same ONE launch as on a single GPU;
      * otherwise each rank evaluates its share and the maxima are MAX-combined inside the kernel over
        NVLink peer memory (`peer`, a PeerMax), or by `reduce_max` (NCCL / gloo all-reduce) between a
        pass-A launch and the fused launch.
    Periodic tiling (tile="auto" | True | False, hostmath.tile_plan): when the packet trains overlap
    densely and share a common period T that fits the L2, the steady part of the slab repeats its first T
    samples bit for bit; it is then placed from the vector itself (one L2 read per sample instead of one
    per overlapping instance) by a second launch that names out_vec[x1 - T, x1) as an attached bank slot.
    spec_tile: the spectrogram of that part repeats every lcm(T, 512) / 512 frames; when one such period fits
    the L2 only its first period is transformed and the rest are copies (hostmath.tile_plan)."""
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
    rows = list(rows)
    lengths = [engine.bank_len(r[3]) for r in rows]
    full_ranges = H.max_eval_ranges(rows, lengths, plan.total) if normalize else []
    if len(full_ranges) > 4:
        full_ranges = [(0, plan.total)]
    redundant = (normalize and frame_major and plan.world > 1 and
                 sum(hi - lo for lo, hi in full_ranges) <= redundant_max)
    in_kernel_max = normalize and frame_major and (redundant or peer is not None or
                                                   (plan.world == 1 and reduce_max is None))
    if normalize and plan.world > 1 and not redundant and peer is None and reduce_max is None:
        raise ValueError("a sharded normalised build needs a way to combine the per-rank maxima: pass `peer` "
                         "(in-kernel exchange), `reduce_max` (all-reduce), or let every rank evaluate the max "
                         "ranges itself (redundant_max)")
    descs, nd = engine.make_descs(rows, engine.free_flags(rows, plan.total) if in_kernel_max else None)

    tp = None
    if tile and frame_major and nf > 0:
        kw = {} if tile == "auto" else dict(min_density=0.0, min_gain=0.25)
        tp = H.tile_plan(rows, lengths, plan.total, plan.lo, plan.hi, **kw)
        if tp is not None and (tp["x2"] // 512 > plan.frame_hi or tp["x1"] >= tp["x2"]):
            tp = None
    # the launches: (lo, hi, frame_lo, frame_hi, descs, n_descs, use the slab's max logic?)
    x1, x2 = (tp["x1"], tp["x2"]) if tp else (plan.hi, plan.hi)
    f1 = min(plan.frame_hi, x1 // 512) if tp else plan.frame_hi

    def vec_at(a):
        return out_vec[a - plan.lo:]

    def spec_at(f):
        return out_spec[:, f - frame_lo:]

    # ---- launch 1: [lo, x1) from the packet bank, with the normalisation
    mode, norm_ranges, use_peer = BUILD_STORE, None, None
    if in_kernel_max:
        if redundant or (plan.world == 1 and peer is None):
            norm_ranges = full_ranges
        else:
            norm_ranges = H.split_ranges(full_ranges, plan.world, plan.rank)
            use_peer = peer
            if len(norm_ranges) > 4:
                raise ValueError("too many max ranges for one launch")
    elif normalize:
        engine._maxsq.zero_()
        engine.max_pass(descs, nd, engine.max_ranges(rows, plan.total, plan.world, plan.rank))
        if reduce_max is not None:
            reduce_max(engine._maxsq)
        mode = BUILD_SCALE
    engine.build_stft1024(descs, nd, plan.lo, x1, out_vec, mode, frame_lo, f1 - frame_lo, out_spec,
                          window=window, fs=fs, scaling=scaling, norm_ranges=norm_ranges, peer=use_peer)
    if tp:
        # ---- launch 2: [x1, x2) = the T samples before x1, repeated (already scaled: STORE)
        t = tp["T"]
        engine.bank_attach(TILE_SLOT, vec_at(x1 - t)[:t])
        tdescs, tnd = engine.make_descs([(x1, t, tp["count"], TILE_SLOT)])
        pf = tp["spec_period_frames"]
        xs = x1 + pf * 512                                   # end of the first spectrogram period of the tiled part
        if spec_tile and pf * 4096 <= SPEC_TILE_BYTES and xs + 512 * max(pf // 4, 1) <= x2:
            # The spectrogram of the periodic part repeats every pf frames: transform ONE period, then the vector
            # comes from the vector tile (span kernel, no transforms) and the frames are copies of that period
            # (b200vec_replicate_frames) — bit for bit what transforming them would give.
            engine.build_stft1024(tdescs, tnd, x1, xs, vec_at(x1), BUILD_STORE, x1 // 512, pf,
                                  spec_at(x1 // 512), window=window, fs=fs, scaling=scaling)
            engine.build_pass(tdescs, tnd, xs, x2, vec_at(xs), BUILD_STORE)
            engine.replicate_frames(out_spec, x1 // 512 - frame_lo, pf, xs // 512 - frame_lo, x2 // 512 - xs // 512)
        else:
            engine.build_stft1024(tdescs, tnd, x1, x2, vec_at(x1), BUILD_STORE, x1 // 512, x2 // 512 - x1 // 512,
                                  spec_at(x1 // 512), window=window, fs=fs, scaling=scaling)
        # ---- launch 3: the tail [x2, hi) from the packet bank again
        if x2 < plan.hi or x2 // 512 < plan.frame_hi:
            engine.build_stft1024(descs, nd, x2, plan.hi, vec_at(x2), BUILD_SCALE if normalize else BUILD_STORE,
                                  x2 // 512, plan.frame_hi - x2 // 512, spec_at(x2 // 512),
                                  window=window, fs=fs, scaling=scaling)
    return out_vec, out_spec, (engine._maxsq if normalize else None)

"""The N > 1 host path on CPU: two gloo ranks run the slab planning, the max-range split and the
one collective of the path (all-reduce MAX of one double, dist.make_max_reducer) exactly as the
GPU ranks do; the per-rank maxima come from the oracle (no GPU here).  Checks that
  * the ranks' slabs and spectrogram frame ranges tile the vector / the frame axis,
  * MAX over the ranks' shares of hostmath.max_eval_ranges equals max|v|^2 of the whole vector,
  * every rank ends up with the same scale factor the single-process build uses."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import ref_port as O
        from vector_b200 import dist as vdist, hostmath as H
        cfgs = O.cfg1_configs()
        pk = O.packets_for(cfgs)
        want, _, counts, max_abs = O.build_vector(pk, cfgs, n, False)
        rows, lengths = [], []
        for i, (cfg, (y, pre)) in enumerate(zip(cfgs, pk)):          # unshifted lengths == shifted lengths
            period = H.period_samples(cfg["period"])
            start = H.start_offset(cfg["start_time"], pre)
            rows.append((start, period, H.instance_count(start, period, len(y), n), i))
            lengths.append(len(y))
        assert [r[2] for r in rows] == counts
        plan = vdist.plan_slab(n, world, rank, 1024, 512, 1, align=512)
        # this rank's share of pass A, evaluated by the oracle on the very ranges the kernel would get
        ranges = H.split_ranges(H.max_eval_ranges(rows, lengths, n), world, rank)
        local = 0.0
        for lo, hi in ranges:
            seg = want[lo:hi].astype(np.complex128)
            if seg.size:
                local = max(local, float(np.max(seg.real ** 2 + seg.imag ** 2)))
        m = torch.tensor([local], dtype=torch.float64)
        reduce_max = vdist.make_max_reducer()
        assert reduce_max is not None
        reduce_max(m)                                                 # the one collective of the path
        full = want.astype(np.complex128)
        assert m.item() == float(np.max(full.real ** 2 + full.imag ** 2))
        scl = np.float32(1.0) / np.float32(np.sqrt(m.item()))
        # gather every rank's plan on rank 0 and check the tiling
        info = [None] * world
        dist.all_gather_object(info, (plan.lo, plan.hi, plan.frame_lo, plan.frame_hi, float(scl), sum(h - l for l, h in ranges)))
        if rank == 0:
            assert info[0][0] == 0 and info[-1][1] == n
            n_frames = (n - 1024) // 512 + 1
            assert info[0][2] == 0 and info[-1][3] == n_frames
            for a, b in zip(info, info[1:]):
                assert a[1] == b[0] and a[3] == b[2] and a[1] % 512 == 0
            assert len({i[4] for i in info}) == 1
            assert sum(i[5] for i in info) == sum(h - l for l, h in H.max_eval_ranges(rows, lengths, n))
            with open(os.path.join(out_dir, "ok"), "w") as fh:
                fh.write("ok")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_rank_gloo_max_exchange_and_slab_tiling(tmp_path, world):
    n = 1_400_000 + 77
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    assert (tmp_path / "ok").read_text() == "ok"


def test_make_max_reducer_is_none_without_process_group():
    from vector_b200 import dist as vdist
    assert vdist.make_max_reducer() is None

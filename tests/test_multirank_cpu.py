"""CPU, world_size 2 over gloo: the N > 1 host logic.  Each rank asks the product library which image
rows it owns (rtb200_band_map: the same band formula the kernels and the gather use), renders exactly
those rows with the oracle, and rank 0 assembles the gathered bands into the frame -- which must be
bit-identical to the single-process frame.  Also checks the max-over-ranks / sum-over-ranks
reductions bench.py relies on."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port_no, H, W, band, out_path):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from __graft_entry__ import load_package
    from oracle.portbind import PortOracle
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port_no}", rank=rank, world_size=world)
    rt = load_package()
    port = PortOracle(rt.abi)
    ps = rt.PackedScene.load(os.path.join(GOLDEN, "packed", "scene7.npz"))
    fr = rt.make_frame(ps.frame, W, H, max_bounces=3, band_rows=band)
    rows = rt.band_map(H, band, world, rank)
    local = np.empty((len(rows), W, 3), np.float32)
    # render the owned rows band by band (contiguous runs of global rows)
    i = 0
    while i < len(rows):
        j = i
        while j + 1 < len(rows) and rows[j + 1] == rows[j] + 1:
            j += 1
        blk, _ = port.render_region(ps, fr, 0, int(rows[i]), W, j - i + 1, threads=2)
        local[i:j + 1] = blk
        i = j + 1
    c = port.ray_counts()                                     # of the last band only; enough for the reduction check
    # gather the compact band buffers and the row maps on rank 0 (gloo)
    bufs = [None] * world if rank == 0 else None
    maps = [None] * world if rank == 0 else None
    dist.gather_object(local, bufs, dst=0)
    dist.gather_object(rows, maps, dst=0)
    t = torch.tensor([float(rank + 1), float(len(rows))], dtype=torch.float64)
    mx = t.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = t.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    assert mx[0].item() == world and sm[1].item() == H        # max over ranks / sum over ranks
    if rank == 0:
        full = np.full((H, W, 3), np.nan, np.float32)
        owner = np.full(H, -1, np.int32)
        for r in range(world):
            assert (owner[maps[r]] == -1).all()               # no row rendered twice
            owner[maps[r]] = r
            full[maps[r]] = bufs[r]
        assert (owner >= 0).all()                             # every row rendered
        np.save(out_path, full)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("H,band", [(70, 16), (48, 0), (33, 8)])
def test_two_ranks_assemble_the_single_process_frame(rt, port, tmp_path, H, band):
    import torch.multiprocessing as mp
    W, world = 40, 2
    out = str(tmp_path / "frame.npy")
    mp.spawn(_worker, args=(world, _free_port(), H, W, band, out), nprocs=world, join=True)
    got = np.load(out)
    ps = rt.PackedScene.load(os.path.join(GOLDEN, "packed", "scene7.npz"))
    want, _ = port.render(ps, rt.make_frame(ps.frame, W, H, max_bounces=3), threads=2)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_band_map_partitions_every_height(rt):
    for n in (1, 2, 3, 4, 8):
        for H in (1, 15, 16, 17, 100, 1080, 2161):
            for band in (0, 1, 16, 64):
                seen = np.zeros(H, np.int32)
                for r in range(n):
                    rows = rt.band_map(H, band, n, r)
                    assert (np.diff(rows) > 0).all()          # compact storage keeps image order
                    seen[rows] += 1
                assert (seen == 1).all()
    b = 16
    assert list(rt.band_map(70, b, 2, 1)) == list(range(16, 32)) + list(range(48, 64))

"""CPU tests of the multi-GPU host logic (libstevi_b200/sharding.py): partition plans and the collectives that carry the
N > 1 path, run with world_size 2 over gloo (no GPU, no compute calls into the CUDA library)."""
import os
import socket

import numpy as np
import pytest

from libstevi_b200 import sharding as S


def test_shard_pairs_partition():
    for n, world in ((1024, 8), (10, 4), (3, 8), (0, 2)):
        owned = [S.shard_pairs(n, world, r) for r in range(world)]
        assert sorted(k for o in owned for k in o) == list(range(n))
        assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1


@pytest.mark.parametrize("height,world,align", [(1080, 8, 1), (3072, 8, 2), (3071, 4, 2), (7, 8, 2), (720, 2, 1)])
def test_row_bands_cover_the_frame(height, world, align):
    bands = S.row_bands(height, world, align)
    assert len(bands) == world and bands[0][0] == 0 and bands[-1][1] == height
    for (a0, a1), (b0, b1) in zip(bands, bands[1:]):
        assert a1 == b0 and a0 <= a1
    assert all(g0 % align == 0 for g0, g1 in bands if g1 > g0)


def test_halo_rows_clip_at_the_frame():
    assert S.halo_rows(0, 10, 4, 100) == (0, 14)
    assert S.halo_rows(90, 100, 4, 100) == (86, 100)
    assert S.halo_rows(40, 50, 3, 100) == (37, 53)


@pytest.mark.parametrize("height,depth,vr", [(3072, 3, 3), (97, 2, 2), (48, 1, 2), (1081, 3, 4)])
def test_band_plan_is_self_consistent(height, depth, vr):
    hs = S.pyramid_heights(height, depth)
    assert hs[-1] == height and all(hs[k] == (hs[k + 1] + 1) // 2 for k in range(depth))
    for g0, g1 in S.row_bands(height, 4, 2):
        if g1 <= g0:
            continue
        plan = S.band_plan(height, depth, [vr] * (depth + 1), g0, g1)
        assert (plan[depth].out_lo, plan[depth].out_hi) == (g0, g1)
        for k, p in enumerate(plan):
            assert 0 <= p.img_lo <= p.out_lo < p.out_hi <= p.img_hi <= p.height == hs[k]
            if k == 0:  # the coarsest cost volume starts at the kernel's 32-row block boundary (bit-exact rounding)
                assert p.img_lo <= max(0, p.out_lo // 32 * 32 - vr)
            # own windows are inside the band unless they leave the frame
            assert p.img_lo <= max(0, p.out_lo - vr) and p.img_hi >= min(p.height, p.out_hi + vr)
            if k < depth:  # pooling the finer band gives exactly this band, anchored at even global rows
                f = plan[k + 1]
                assert f.img_lo % 2 == 0 and p.img_lo == f.img_lo // 2 and p.img_hi == (f.img_hi + 1) // 2
                # every guide row the finer level interpolates from is produced
                v_lo = f.out_lo * (p.height - 1) // (f.height - 1)
                v_hi = min(p.height - 1, -(-(f.out_hi - 1) * (p.height - 1) // (f.height - 1)) + 1)
                assert p.out_lo <= v_lo and v_hi < p.out_hi


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        H, W = 37, 11
        full = torch.arange(H * W, dtype=torch.int32).reshape(H, W)
        bands = S.row_bands(H, world, 2)
        g0, g1 = bands[rank]
        got = S.gather_rows(full[g0:g1].clone(), bands, dist)
        assert torch.equal(got, full)
        vol = torch.arange(H * W * 3, dtype=torch.float32).reshape(H, W, 3)
        assert torch.equal(S.gather_rows(vol[g0:g1].clone(), bands, dist), vol)
        # equal bands take the in-place all_gather_into_tensor path
        He = 40
        eq_bands = S.row_bands(He, world, 2)
        assert len({b[1] - b[0] for b in eq_bands}) == 1
        full_e = torch.arange(He * W * 3, dtype=torch.float32).reshape(He, W, 3)
        e0, e1 = eq_bands[rank]
        assert torch.equal(S.gather_rows(full_e[e0:e1].clone(), eq_bands, dist), full_e)
        slowest = S.max_over_ranks(1.5 + rank, dist)
        assert slowest == 1.5 + world - 1
        pairs = S.shard_pairs(9, world, rank)
        counts = torch.tensor([len(pairs)])
        dist.all_reduce(counts)
        assert int(counts) == 9
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_world_size_2_gloo(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))

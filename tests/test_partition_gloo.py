"""The N>1 host logic on CPU: world_size-2 (and 3) gloo process groups exercise the
sample-budget partition and the single sum-reduce exactly as bench.py drives them on
GPUs (there with NCCL).  The per-rank "render" is the CPU oracle restricted to the
rank's slice — possible because the oracle's per-row RNG makes slices of the sample
loop non-reproducible, so here each rank renders its slice as an independent image with
`samples = slice length` and the GLOBAL divisor applied afterwards."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from raytracingmin_b200.partition import sample_slice, plan_partition, reduce_partials  # noqa: E402


def test_sample_slice_properties():
    for S in (1, 2, 7, 25, 64, 100):
        for G in (1, 2, 3, 4, 8):
            if S < G:
                with pytest.raises(ValueError):
                    sample_slice(S, 0, G)
                continue
            sl = [sample_slice(S, g, G) for g in range(G)]
            assert sl[0][0] == 0 and sl[-1][1] == S
            assert all(a[1] == b[0] for a, b in zip(sl, sl[1:]))
            lens = [b - a for a, b in sl]
            assert min(lens) >= 1 and max(lens) - min(lens) <= 1
    assert sample_slice(64, 3, 8) == (24, 32)
    with pytest.raises(ValueError):
        sample_slice(8, 8, 8)


def test_plan_partition_covers_the_budget_once():
    for S, ss in ((64, 4), (25, 4), (100, 1), (7, 3), (8, 2)):
        for G in (1, 2, 3, 4, 8):
            if G > 1 and (ss * ss) % G != 0 and S < G:
                continue
            cover = np.zeros((ss * ss, S), int)
            for g in range(G):
                pl = plan_partition(S, ss, g, G)
                cover[pl["sub_begin"]:pl["sub_end"], pl["sample_begin"]:pl["sample_end"]] += 1
            assert (cover == 1).all(), (S, ss, G)
    # C4: 16 sub-pixels over 8 GPUs -> grid slices, full sample loops
    assert plan_partition(64, 4, 3, 8) == dict(sub_begin=6, sub_end=8, sample_begin=0, sample_end=64)
    # ss = 1 cannot slice the grid -> samples
    assert plan_partition(256, 1, 1, 2) == dict(sub_begin=0, sub_end=1, sample_begin=128, sample_end=256)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, S, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle_lib as O
        import raytracingmin_b200 as rtm
        sc = rtm.load_scene(os.path.join(ROOT, "scenes", "settingData.head.json"), "abc").with_dims(48, 26, S, 1)
        s0, s1 = sample_slice(S, rank, world)
        # the rank's partial buffer: its slice rendered with the GLOBAL divisor (Renderer.cpp:240)
        part, rays, samples = O.oracle_render(sc.with_dims(samples=s1 - s0), O.LEVEL_ABC, 1000 + rank)
        part = part * (s1 - s0) / S
        accum = torch.zeros((sc.height, sc.width, 4), dtype=torch.float32)
        accum[..., :3] = torch.from_numpy(part.astype(np.float32))
        cnt = torch.tensor([rays, samples], dtype=torch.float64)
        reduce_partials(accum, dst=0)
        dist.all_reduce(cnt)
        t = torch.tensor([float(rank + 1)])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)      # the max-over-ranks timing reduction of bench.py
        assert t.item() == world
        if rank == 0:
            np.save(os.path.join(out_dir, "sum.npy"), accum.numpy())
            np.save(os.path.join(out_dir, "cnt.npy"), cnt.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slices_reduce_to_full_image_gloo(tmp_path, world):
    S = 24   # slice length >= 8 > emission 5: the per-sample clamp (Renderer.cpp:241) stays inactive in every slice
    mp.spawn(_worker, args=(world, _free_port(), S, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "sum.npy")
    cnt = np.load(tmp_path / "cnt.npy")
    import oracle_lib as O
    import raytracingmin_b200 as rtm
    sc = rtm.load_scene(os.path.join(ROOT, "scenes", "settingData.head.json"), "abc").with_dims(48, 26, 2048, 1)
    ref, rays, samples = O.oracle_render(sc, O.LEVEL_ABC, 5)
    assert cnt[1] == 48 * 26 * S                       # every sample rendered exactly once
    assert (got[..., 3] == 0).all()
    # statistical agreement of the reduced image with a converged render (24 spp vs 2048 spp)
    assert abs(got[..., :3].mean() - ref.mean()) < 0.1 * ref.mean()
    # and with the deterministic part exactly: pixels that only see the light / nothing carry no noise
    quiet = (ref.max(axis=2) == 0)
    assert (got[..., :3][quiet] == 0).all()

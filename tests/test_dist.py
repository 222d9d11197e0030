"""Host-side multi-GPU logic on CPU: the shard plan partitions the wavefront exactly, and the film merge (all-reduce
sum) over a world_size-2 gloo group reproduces the single-process film."""
import os
import socket

import numpy as np
import pytest

from mitsuba3_experiments_b200.dist import shard_plan


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("rows,width,spp", [(720, 1280, 64), (23, 41, 12), (5, 7, 1)])
def test_shard_plan_partitions_wavefront(world, rows, width, spp):
    n = rows * width * spp
    ranges = sorted(r for rank in range(world) for r in shard_plan(n, spp, width, world, rank))
    assert ranges[0][0] == 0 and ranges[-1][1] == n
    for (a0, a1), (b0, b1) in zip(ranges[:-1], ranges[1:]):
        assert a1 == b0 and a1 > a0
    for a, b in ranges:
        assert a % (width * spp) == 0 and b % (width * spp) == 0       # whole image rows
    per_rank = [sum(b - a for a, b in shard_plan(n, spp, width, world, r)) for r in range(world)]
    assert sum(per_rank) == n
    if rows >= 8 * world:
        assert max(per_rank) - min(per_rank) <= n // world * 0.2


def _worker(rank, world, port, out_dir):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    from mitsuba3_experiments_b200 import dist as b2dist
    r, w, _ = b2dist.init_from_env("gloo")
    assert (r, w) == (rank, world)
    rows, width, spp = 16, 8, 4
    n = rows * width * spp
    film = torch.zeros((rows, width, 4))
    # stand-in for the render: every lane deposits its (deterministic) value into its own pixel
    for a, b in shard_plan(n, spp, width, world, rank, bands_per_rank=2):
        lanes = torch.arange(a, b)
        pix = lanes // spp
        vals = torch.stack([torch.sin(lanes * 0.37), torch.cos(lanes * 0.11), lanes.float() / n, torch.ones(len(lanes))], 1).double()
        film.view(-1, 4).index_add_(0, pix, vals.float())
    b2dist.merge_sum(film)
    np.save(os.path.join(out_dir, f"film{rank}.npy"), film.numpy())
    torch.distributed.destroy_process_group()


def test_film_merge_gloo_world2(tmp_path):
    import torch
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    f0, f1 = np.load(tmp_path / "film0.npy"), np.load(tmp_path / "film1.npy")
    assert np.array_equal(f0, f1)
    rows, width, spp = 16, 8, 4
    n = rows * width * spp
    lanes = torch.arange(n)
    ref = torch.zeros((rows * width, 4))
    ref.index_add_(0, lanes // spp, torch.stack([torch.sin(lanes * 0.37), torch.cos(lanes * 0.11), lanes.float() / n, torch.ones(n)], 1))
    np.testing.assert_allclose(f0.reshape(-1, 4), ref.numpy(), rtol=1e-5, atol=1e-5)


def test_sample_shard_partitions_samples():
    from mitsuba3_experiments_b200.dist import sample_shard
    for world in (2, 4, 8):
        parts = [sample_shard(64, world, r) for r in range(world)]
        assert [p[0] for p in parts] == [r * (64 // world) for r in range(world)] and all(p[1] == 64 // world for p in parts)
    assert sample_shard(64, 1, 0) is None and sample_shard(6, 4, 1) is None      # single rank / uneven split -> image bands

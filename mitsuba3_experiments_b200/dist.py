"""Multi-GPU sharding: one process per GPU (torchrun), lanes split by image band x all samples, films and
parameter gradients merged with one NCCL all-reduce over NVLink.

The reference has no multi-device layer at all; this is greenfield.  Lane ``i``'s PCG32 stream and pixel
are pure functions of ``(seed, i)`` (sampler.cpp:136-144, integrator.cpp:293-305), so any partition of the lane range
reproduces the single-GPU image up to fp32 summation order.  Bands are interleaved across ranks (rank r takes bands
r, r + world, ...) so that expensive image regions are spread evenly; the only collective is the film / gradient sum.
"""
from __future__ import annotations

import os


def shard_plan(wavefront: int, spp_per_pass: int, width: int, world: int, rank: int, bands_per_rank: int = 8):
    """Contiguous lane ranges [(begin, end), ...] of this rank.  Bands are whole image rows (so film tiles of
    different ranks only meet at band borders); their union over ranks is exactly [0, wavefront)."""
    assert wavefront % (spp_per_pass * width) == 0
    rows = wavefront // (spp_per_pass * width)
    n_bands = max(1, min(rows, world * bands_per_rank))
    out = []
    for b in range(rank, n_bands, world):
        r0, r1 = rows * b // n_bands, rows * (b + 1) // n_bands
        if r1 > r0:
            out.append((r0 * width * spp_per_pass, r1 * width * spp_per_pass))
    return out


def sample_shard(spp_per_pass: int, world: int, rank: int):
    """(sample_begin, sample_count) of this rank: every rank renders every pixel with its own slice of the samples of a
    pass, so the split is perfectly balanced whatever the image content (b2_render_args.sample_begin / sample_count).
    Returns None when the samples do not divide evenly -- use shard_plan (image bands) then."""
    if world <= 1 or spp_per_pass % world:
        return None
    n = spp_per_pass // world
    return rank * n, n


def init_from_env(backend: str | None = None):
    """Initialise torch.distributed from torchrun's environment (RANK / WORLD_SIZE / LOCAL_RANK / MASTER_*).
    Returns (rank, world, local_rank); a plain single-process run returns (0, 1, 0) without initialising anything."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local


def merge_sum(tensor):
    """In-place sum over ranks (film RGBW or a parameter gradient).  No-op for a single process."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
    return tensor


def render_sharded(integrator, scene, seed=0, spp=0, rank=0, world=1, develop=True, bands_per_rank=8):
    """Render this rank's bands into a private film, all-reduce, develop.  Every rank returns the full image."""
    ds = integrator._scene(scene)
    spp = spp or ds.desc.spp
    spp_pp = integrator.samples_per_pass or spp
    n = ds.wavefront_size(spp, integrator.samples_per_pass)
    spp_pp = n // (ds.lane_grid[0] * ds.lane_grid[1])
    film = ds.new_film()
    for a, b in shard_plan(n, spp_pp, ds.lane_grid[1], world, rank, bands_per_rank):
        integrator.render(ds, seed=seed, spp=spp, develop=False, evaluate=False, film=film, lane_begin=a, lane_end=b)
    merge_sum(film)
    return ds.develop(film) if develop else film


def render_sharded_samples(integrator, scene, seed=0, spp=0, rank=0, world=1, develop=True, film=None):
    """The balanced split: every rank renders every pixel with its slice of the samples (``sample_shard``), one film
    all-reduce.  Falls back to image bands when the samples per pass do not divide by the world size."""
    ds = integrator._scene(scene)
    spp = spp or ds.desc.spp
    n = ds.wavefront_size(spp, integrator.samples_per_pass)
    spp_pp = n // (ds.lane_grid[0] * ds.lane_grid[1])
    sh = sample_shard(spp_pp, world, rank)
    if world > 1 and sh is None:
        return render_sharded(integrator, scene, seed, spp, rank, world, develop)
    film = ds.new_film() if film is None else film
    kw = dict(sample_begin=sh[0], sample_count=sh[1]) if sh else {}
    integrator.render(ds, seed=seed, spp=spp, develop=False, evaluate=False, film=film, **kw)
    merge_sum(film)
    return ds.develop(film) if develop else film


def render_backward_sharded(integrator, scene, params, grad_in, seed=0, spp=0, rank=0, world=1, bands_per_rank=8, split="samples"):
    """PRB adjoint of this rank's shard; gradients are all-reduced into ``params.grad`` (RBIntegrator.render_backward,
    common.py:623-781, over N devices).  The weight film of the WHOLE sample set is recomputed on every rank (positions
    depend on (seed, lane) only; it is cheap and keeps the pass free of a second collective).  ``split``: "samples" =
    every rank replays its slice of every pixel's samples (balanced), "bands" = image bands."""
    ds = integrator._scene(scene)
    spp = spp or ds.desc.spp
    n = ds.wavefront_size(spp, 0)
    sh = sample_shard(spp, world, rank) if split == "samples" else None
    if world <= 1:
        integrator.render_backward(ds, params, grad_in, seed=seed, spp=spp)
    elif sh is not None:
        integrator.render_backward(ds, params, grad_in, seed=seed, spp=spp, sample_begin=sh[0], sample_count=sh[1])
    else:
        for i, (a, b) in enumerate(shard_plan(n, spp, ds.lane_grid[1], world, rank, bands_per_rank)):
            integrator.render_backward(ds, params, grad_in, seed=seed, spp=spp, lane_begin=a, lane_end=b, reuse_weights=i > 0)
    for g in params.grad.values():
        merge_sum(g)

"""Multi-GPU correctness on hardware (needs >= 2 CUDA devices; `gpurun --gpus 2 -- python -m pytest tests/test_gpu_multi.py -m gpu`):
the NCCL-merged film of a 2-rank sample-slice render equals the 1-GPU render, and the all-reduced PRB gradients of two ranks
equal the 1-rank gradients (RBIntegrator.render_backward over N devices).  The CPU twin (gloo) is tests/test_dist.py.
"""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def _need_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 CUDA devices")


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch
    import torch.distributed as tdist
    from mitsuba3_experiments_b200 import dist as b2dist
    from mitsuba3_experiments_b200.integrator import PathIntegrator, PRBIntegrator, traverse
    from mitsuba3_experiments_b200.runtime import DeviceScene
    from mitsuba3_experiments_b200.scenes import cornell_box
    from scenes_for_tests import materials_scene
    r, w, local = b2dist.init_from_env("nccl")
    assert (r, w) == (rank, world)
    torch.cuda.set_device(local)
    # ---- (a) film merge: sample slices and image bands
    ds = DeviceScene(materials_scene(64, 48, 16, envmap=True, filter_type="gaussian"), device=local)
    integ = PathIntegrator({"max_depth": 6, "rr_depth": 2})
    img_s = b2dist.render_sharded_samples(integ, ds, seed=9, spp=16, rank=rank, world=world)
    img_b = b2dist.render_sharded(integ, ds, seed=9, spp=16, rank=rank, world=world, bands_per_rank=3)
    # ---- (b) PRB gradients, both splits
    rng = np.random.RandomState(0)
    d = cornell_box(48, 48, spp=16)
    d["white"] = {"type": "diffuse", "reflectance": {"type": "bitmap", "data": rng.uniform(0.2, 0.9, (24, 32, 3)).astype(np.float32), "raw": True}}
    dp = DeviceScene(d, device=local)
    prb = PRBIntegrator({"max_depth": 6, "rr_depth": 3})
    grad_in = torch.from_numpy(np.random.RandomState(1).normal(size=(48, 48, 3)).astype(np.float32) / (48 * 48 * 3)).to(dp.device)
    grads = {}
    for split in ("samples", "bands"):
        params = traverse(dp); params.keep(["white.reflectance.data", "red.reflectance.value"])
        b2dist.render_backward_sharded(prb, dp, params, grad_in, seed=7, spp=16, rank=rank, world=world, bands_per_rank=2, split=split)
        grads[split] = {k: v.cpu().numpy() for k, v in params.grad.items()}
    torch.cuda.synchronize()
    if rank == 0:
        one, _ = ds.render(seed=9, spp=16, max_depth=6, rr_depth=2, variant=1)
        params = traverse(dp); params.keep(["white.reflectance.data", "red.reflectance.value"])
        prb.render_backward(dp, params, grad_in, seed=7, spp=16)
        torch.cuda.synchronize()
        np.savez(os.path.join(out_dir, "out.npz"), img_s=img_s.cpu().numpy(), img_b=img_b.cpu().numpy(), one=one.cpu().numpy(),
                 **{f"g1_{k}": v.cpu().numpy() for k, v in params.grad.items()},
                 **{f"g_{s}_{k}": v for s, g in grads.items() for k, v in g.items()})
    tdist.barrier()
    tdist.destroy_process_group()


def test_two_rank_nccl_render_and_prb_gradients_equal_one_rank(tmp_path):
    _need_two_gpus()
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 400)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    o = np.load(os.path.join(str(tmp_path), "out.npz"))
    np.testing.assert_allclose(o["img_s"], o["one"], rtol=2e-5, atol=1e-6)
    np.testing.assert_allclose(o["img_b"], o["one"], rtol=2e-5, atol=1e-6)
    for k in ("white.reflectance.data", "red.reflectance.value"):
        ref = o[f"g1_{k}"]
        scale = np.abs(ref).max()
        for s in ("samples", "bands"):
            g = o[f"g_{s}_{k}"]
            assert np.abs(g - ref).max() <= 1e-5 * scale + 1e-12, (k, s, np.abs(g - ref).max(), scale)
            assert abs(g.sum() - ref.sum()) <= 1e-5 * np.abs(ref).sum(), (k, s)

#!/usr/bin/env python
"""Inverse rendering with the PRB integrator: recover the WoodFloor albedo texture of the Cornell box (BASELINE config 3).

The reference flow (mitsuba3/src/python/python/util.py:395-526, ad/optimizers.py; tutorials/inverse_rendering):

    image = mi.render(scene, params, seed=it, spp=spp)      # primal, detached inside the custom op
    loss  = mse(image, target);  dr.backward(loss)          # -> integrator.render_backward(grad_in, seed_grad, spp_grad)
    opt.step();  params.update(opt)

Here ``render`` is the same custom op on torch.autograd (``seed_grad = sample_tea_32(seed, 1)[0]``, util.py:504-509), the
optimiser is ``torch.optim.Adam`` on the fp32 texture tensor, and every kernel runs through the C-ABI on the GPU.

    python examples/optimize_woodfloor.py [--res 128] [--spp 64] [--iters 80]
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def optimize(res=128, spp=64, iters=80, lr=0.02, eval_spp=4096, device=0, verbose=False):
    """Returns a dict with the clean image loss and the texel error before and after the optimisation.  The clean loss compares two
    independent eval_spp renders (target and current), so its floor is twice the Monte-Carlo variance at eval_spp: 4 096 spp
    keeps that floor ~40x below the initial loss."""
    import torch
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, load_dict, render, traverse
    from mitsuba3_experiments_b200.scenes import cornell_prb_scene, woodfloor_texture

    truth = torch.from_numpy(woodfloor_texture())
    ds = load_dict(cornell_prb_scene(res, res, spp), device=device)
    integ = PRBIntegrator({"max_depth": 6, "rr_depth": 5})
    key = "white.reflectance.data"
    params = traverse(ds); params.keep([key])
    truth = truth.to(ds.device)
    target = integ.render(ds, seed=12345, spp=eval_spp)                 # reference image of the true texture, another seed

    params[key] = (0.5 * truth).clone()                                 # start: the texture at half its albedo
    params.update()

    def clean_loss():
        return float(((integ.render(ds, seed=999, spp=eval_spp) - target) ** 2).mean())

    def texel_err():
        return float((params[key].detach() - truth).abs().mean())

    out = {"loss_before": clean_loss(), "texel_err_before": texel_err(), "history": []}
    params[key].requires_grad_(True)
    opt = torch.optim.Adam([params[key]], lr=lr)
    for it in range(iters):
        opt.zero_grad()
        image = render(ds, params, integrator=integ, seed=it, spp=spp)   # differentiable w.r.t. params[key]
        loss = ((image - target) ** 2).mean()
        loss.backward()                                                  # -> render_backward with seed_grad = tea(seed, 1)
        opt.step()
        with torch.no_grad():
            params[key].clamp_(0.0, 1.0)                                 # reflectances stay physical (tutorial: dr.clip)
        out["history"].append(float(loss.detach()))
        if verbose:
            print(f"iter {it:3d}  loss {float(loss.detach()):.6f}  texel err {texel_err():.4f}")
    params[key].requires_grad_(False)
    params.update()
    out.update(loss_after=clean_loss(), texel_err_after=texel_err())
    return out


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--res", type=int, default=128)
    ap.add_argument("--spp", type=int, default=64)
    ap.add_argument("--iters", type=int, default=80)
    ap.add_argument("--lr", type=float, default=0.02)
    a = ap.parse_args()
    r = optimize(a.res, a.spp, a.iters, a.lr, verbose=True)
    print({k: v for k, v in r.items() if k != "history"})

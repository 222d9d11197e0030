"""Profiling driver (GPU box): PRB fwd + adjoint iterations on the C3 workload (Cornell 512x512, 32 spp, textured albedo)."""
import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import numpy as np, torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
from mitsuba3_experiments_b200.runtime import DeviceScene
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
d = scenes.cornell_prb_scene(512, 512, 32)
ds = DeviceScene(d)
prb = PRBIntegrator({"max_depth": 8, "rr_depth": 5})
params = traverse(ds); params.keep(["white.reflectance.data"])
target = prb.render(ds, seed=1, spp=32) * 0.5
for i in range(n):
    torch.cuda.synchronize(); t = time.perf_counter()
    img = prb.render(ds, seed=i, spp=32, evaluate=False)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    grad_in = 2.0 * (img - target) / img.numel()
    params.zero_grad()
    prb.render_backward(ds, params, grad_in, seed=1000 + i, spp=32)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"iter {i}: fwd {1e3*(t1-t):.2f} ms  bwd {1e3*(t2-t1):.2f} ms")

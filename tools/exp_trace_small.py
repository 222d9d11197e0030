"""Experiment (GPU box): b2_trace on random rays inside a small scene, wavefront trace kernel vs per-thread traversal."""
import sys, os, time, ctypes as C
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import numpy as np, torch
from mitsuba3_experiments_b200 import scenes, capi
from mitsuba3_experiments_b200.runtime import DeviceScene
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8_000_000
ds = DeviceScene(scenes.cornell_box(64, 64))
rng = np.random.RandomState(1)
o = rng.uniform(-0.9, 0.9, size=(n, 3)); d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
rays = torch.from_numpy(np.concatenate([o, d, np.full((n, 1), 3.4e38), np.zeros((n, 1))], axis=1).astype(np.float32)).to(ds.device)
hits = torch.empty((n, 4), dtype=torch.float32, device=ds.device)
for flags, name in ((0, "wavefront closest"), (2, "per-thread closest"), (1, "wavefront any"), (3, "per-thread any")):
    for i in range(4):
        torch.cuda.synchronize(); t = time.perf_counter()
        capi.check(ds.lib.b2_trace(ds.handle, rays.data_ptr(), n, flags, hits.data_ptr(), 0))
        torch.cuda.synchronize(); dt = time.perf_counter() - t
    print(f"{name}: {dt*1e3:.3f} ms  {n/dt/1e9:.2f} Grays/s")

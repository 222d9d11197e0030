"""Experiment (GPU box): cost of the bounce-0 trace alone -- bedroom stand-in, max_depth = 1 (one trace launch per render)."""
import sys, os, time, json
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.runtime import DeviceScene
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
palette = sys.argv[2] if len(sys.argv) > 2 else None
d = scenes.bedroom_standin(1280, 720, spp, **({"palette": palette} if palette else {}))
ds = DeviceScene(d)
film = ds.new_film()
for md in (1, 2, 8):
    kw = dict(variant=1, max_depth=md, rr_depth=2)
    for i in range(3):
        film.zero_(); torch.cuda.synchronize(); t = time.perf_counter()
        ds.render_film(film=film, seed=0, spp=spp, **kw); torch.cuda.synchronize()
        dt = time.perf_counter() - t
    _, c = ds.render_film(film=film, counters=True, seed=0, spp=spp, **kw)
    ds.set_kernel_timing(True) if hasattr(ds, "set_kernel_timing") else None
    print(f"max_depth {md}: {dt*1e3:.2f} ms; per closest ray: nodes {c['node_fetches']/max(1,c['closest_rays']+c['shadow_rays']):.2f} tris {c['tri_fetches']/max(1,c['closest_rays']+c['shadow_rays']):.2f}", json.dumps(c))

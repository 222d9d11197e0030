"""Profiling driver (GPU box): a few launches of the render kernel on a named workload; used under ncu."""
import sys, os, time, json
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.runtime import DeviceScene
wl = sys.argv[1] if len(sys.argv) > 1 else "bedroom"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
spp = int(sys.argv[3]) if len(sys.argv) > 3 else 0
kernel = int(sys.argv[4]) if len(sys.argv) > 4 else 0   # 0 wavefront, 1 megakernel
chunk = int(sys.argv[5]) if len(sys.argv) > 5 else 0
if wl == "bedroom":
    spp = spp or 64; d = scenes.bedroom_standin(1280, 720, spp); kw = dict(variant=1, max_depth=8, rr_depth=2)
elif wl == "bedroom_xml":
    spp = spp or 64; d = scenes.bedroom_standin(1280, 720, spp, palette="scene_xml"); kw = dict(variant=1, max_depth=8, rr_depth=2)
elif wl == "cornell":
    spp = spp or 16; d = scenes.cornell_box(256, 256, spp); kw = dict(variant=0, max_depth=8, rr_depth=4)
elif wl == "tris":
    spp = spp or 8; d = scenes.random_triangles(10_000_000, 1920, 1080, spp); kw = dict(variant=0, max_depth=8, rr_depth=4)
kw.update(kernel=kernel, chunk_lanes=chunk, env_warp=int(os.environ.get('B2_ENV_WARP', '0')))
ds = DeviceScene(d)
print("build", ds.build_stats.as_dict())
film = ds.new_film()
for i in range(n):
    film.zero_(); torch.cuda.synchronize(); t = time.perf_counter()
    ds.render_film(film=film, seed=0, spp=spp, **kw); torch.cuda.synchronize()
    dt = time.perf_counter() - t
    print(f"launch {i}: {dt*1e3:.2f} ms  {film.shape[0]*film.shape[1]*spp/dt/1e6:.1f} Msamples/s")
_, c = ds.render_film(film=film, counters=True, seed=0, spp=spp, **kw)
print(json.dumps(c))

"""A/B helper: per-kind kernel times of one render of a named workload, with the package found first on sys.path (argv[1])."""
import sys, os, time
root = sys.argv[1]; sys.path.insert(0, root)
import torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.runtime import DeviceScene
wl, spp = sys.argv[2], int(sys.argv[3])
if wl == "soup": d = scenes.random_triangles(10_000_000, 1920, 1080, spp); kw = dict(variant=0, max_depth=8, rr_depth=4)
elif wl == "bedroom": d = scenes.bedroom_standin(1280, 720, spp); kw = dict(variant=1, max_depth=8, rr_depth=2)
else: d = scenes.cornell_box(256, 256, spp); kw = dict(variant=0, max_depth=8, rr_depth=4)
ds = DeviceScene(d); film = ds.new_film()
for i in range(2): ds.render_film(film=film, seed=0, spp=spp, **kw)
torch.cuda.synchronize(); ds.set_kernel_timing(True); ds.kernel_times()
t = time.perf_counter(); ds.render_film(film=film, seed=0, spp=spp, **kw); torch.cuda.synchronize(); dt = time.perf_counter() - t
print(os.path.basename(root.rstrip("/")) or "cur", wl, spp, "wall %.2f ms" % (dt * 1e3), {k: round(v[0], 2) for k, v in ds.kernel_times().items()}, ds.build_stats.as_dict()["builder"])

import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.runtime import DeviceScene
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 8
ds = DeviceScene(scenes.bedroom_standin(1280, 720, spp))
kw = dict(variant=1, max_depth=8, rr_depth=2)
film = ds.new_film()
for i in range(3): ds.render_film(film=film, seed=0, spp=spp, **kw)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ds.render_film(film=film, seed=0, spp=spp, **kw); e1.record(); torch.cuda.synchronize()
print("device time of one render (no per-kernel events):", e0.elapsed_time(e1))
ds.set_kernel_timing(True); ds.kernel_times()
t = time.perf_counter(); ds.render_film(film=film, seed=0, spp=spp, **kw); t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
kt = ds.kernel_times()
print("host enqueue ms", (t1 - t) * 1e3, "wall ms", (t2 - t) * 1e3, "sum of kernel ms", sum(v[0] for v in kt.values()), kt)

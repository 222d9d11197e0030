"""Where does the host-buffer call (b2_render_host) spend its wall-clock?  (GPU box)"""
import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.runtime import DeviceScene
spp = 64
d = scenes.bedroom_standin(1280, 720, spp); kw = dict(variant=1, max_depth=8, rr_depth=2)
ds = DeviceScene(d); H, W = ds.film_shape
film = ds.new_film()
film_h = torch.zeros((H, W, 4), dtype=torch.float32).pin_memory(); img_h = torch.zeros((H, W, 3), dtype=torch.float32).pin_memory()
def wall(f, n=4):
    ts = []
    for _ in range(n):
        torch.cuda.synchronize(); t = time.perf_counter(); f(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t) * 1e3)
    return " ".join(f"{x:.1f}" for x in ts)
def dev():
    film.zero_(); ds.render_film(film=film, seed=0, spp=spp, **kw); return ds.develop(film)
def dev_copy():
    film.copy_(film_h, non_blocking=True); ds.render_film(film=film, seed=0, spp=spp, **kw); img = ds.develop(film)
    img_h.copy_(img, non_blocking=True); film_h.copy_(film, non_blocking=True)
print("device path wall ms:", wall(dev))
print("device path + pinned copies wall ms:", wall(dev_copy))
print("render_host wall ms:", wall(lambda: ds.render_host(film_h.numpy(), img_h.numpy(), seed=0, spp=spp, **kw)))
print("render_host (no image) wall ms:", wall(lambda: ds.render_host(film_h.numpy(), None, seed=0, spp=spp, **kw)))

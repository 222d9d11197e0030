"""Experiment (one GPU as a proxy for rank r of 8): the bedroom at a fixed 64 spp split over 8 ranks by sample slices (8 spp of every
pixel) or by interleaved image bands (all 64 spp of 1/8 of the rows): time of each rank's share on this GPU."""
import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from mitsuba3_experiments_b200 import scenes, dist as b2dist
from mitsuba3_experiments_b200.runtime import DeviceScene
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
spp = 64
ds = DeviceScene(scenes.bedroom_standin(1280, 720, spp)); film = ds.new_film()
kw = dict(variant=1, max_depth=8, rr_depth=2)
def timed(fn):
    for _ in range(2): fn()
    torch.cuda.synchronize(); t = time.perf_counter()
    for _ in range(3): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t) / 3 * 1e3
sl = []
for r in (0, world // 2, world - 1):
    sh = b2dist.sample_shard(spp, world, r)
    sl.append(timed(lambda: ds.render_film(film=film, seed=0, spp=spp, sample_begin=sh[0], sample_count=sh[1], **kw)))
print(f"sample slices, {world} ranks: rank 0 / mid / last = " + " / ".join(f"{x:.2f}" for x in sl) + " ms")
H, W = ds.film_shape
for bpr in (1, 2, 4, 8, 16):
    ts = []
    for r in range(world):
        plan = b2dist.shard_plan(W * H * spp, spp, W, world, r, bands_per_rank=bpr)
        def job():
            for a, b in plan: ds.render_film(film=film, seed=0, spp=spp, lane_begin=a, lane_end=b, **kw)
        ts.append(timed(job))
    print(f"image bands, {bpr} per rank: max {max(ts):.2f} ms, mean {sum(ts)/len(ts):.2f} ms, min {min(ts):.2f} ms")

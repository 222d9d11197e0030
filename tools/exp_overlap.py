"""Experiment: two half renders (32 spp each) on two CUDA streams vs one 64-spp render -- do the issue-bound trace kernel of one
half and the latency-bound shade kernels of the other overlap on the SMs?  (B2_TRACE_BLOCKS limits the persistent trace grid.)"""
import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R)
import torch
from mitsuba3_experiments_b200 import scenes
from mitsuba3_experiments_b200.runtime import DeviceScene
SPP = int(sys.argv[1]) if len(sys.argv) > 1 else 64
d = scenes.bedroom_standin(1280, 720, SPP)
kw = dict(variant=1, max_depth=8, rr_depth=2)
a, b = DeviceScene(d), DeviceScene(d)
fa, fb = a.new_film(), b.new_film()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def single():
    fa.zero_(); torch.cuda.synchronize(); t = time.perf_counter()
    a.render_film(film=fa, seed=0, spp=SPP, **kw); torch.cuda.synchronize(); return time.perf_counter() - t
def dual(stagger=0):
    fa.zero_(); fb.zero_(); torch.cuda.synchronize(); t = time.perf_counter()
    with torch.cuda.stream(s1):
        a.render_film(film=fa, seed=0, spp=SPP, sample_begin=0, sample_count=SPP // 2, **kw)
    with torch.cuda.stream(s2):
        b.render_film(film=fb, seed=0, spp=SPP, sample_begin=SPP // 2, sample_count=SPP // 2, **kw)
    torch.cuda.synchronize(); return time.perf_counter() - t
for i in range(3): print("single %.2f ms" % (single() * 1e3))
for i in range(4): print("dual   %.2f ms" % (dual() * 1e3))

"""Experiment (GPU box): crossover of the per-thread trace kernel (k_wf_trace_simple) against the vote-scheduled one by tree size."""
import sys, os, time
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch
from scenes_for_tests import triangles_scene
from mitsuba3_experiments_b200.runtime import DeviceScene
for n in (50, 200, 800, 3000, 12000, 50000):
    ds = DeviceScene(triangles_scene(n, 512, 512, 16)); film = ds.new_film()
    out = []
    for thr in ("1000000000", "0"):
        os.environ["B2_SIMPLE_TRACE_NODES"] = thr
        for i in range(3): ds.render_film(film=film, seed=0, spp=16, variant=1, max_depth=8, rr_depth=4)
        torch.cuda.synchronize(); ds.set_kernel_timing(True); ds.kernel_times()
        for i in range(3): ds.render_film(film=film, seed=0, spp=16, variant=1, max_depth=8, rr_depth=4)
        torch.cuda.synchronize(); out.append(ds.kernel_times()["trace"][0] / 3); ds.set_kernel_timing(False)
    print(f"tris {n:6d} nodes {ds.build_stats.as_dict()['n_nodes8']:6d}: trace simple {out[0]:.3f} ms, vote-scheduled {out[1]:.3f} ms")

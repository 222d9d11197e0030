import sys, os
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
from scenes_for_tests import materials_scene
from mitsuba3_experiments_b200.runtime import DeviceScene
from mitsuba3_experiments_b200.scenes import cornell_box
from oracle.oracle_py import OracleScene
W, H = 256, 192
d = cornell_box(W, H, 1); m = materials_scene(W, H, 1)
d["glass"] = m["glass"]; d["small-box"]["bsdf"] = {"type": "ref", "id": "glass"}
d["sensor"]["film"]["rfilter"] = {"type": "box"}
ds = DeviceScene(d); orc = OracleScene(ds.desc)
def bad(**kw):
    img, c = ds.render(counters=True, **kw); ref, rc = orc.render(**kw); img = img.cpu().numpy()
    e = np.abs(img - ref).max(axis=2) / np.maximum(np.abs(ref).max(axis=2), 2e-2)
    return np.argwhere(e > 1e-3), img, ref, c, rc
for rr in (3, 100):
    for md in (4, 5, 6, 7, 10):
        b, img, ref, c, rc = bad(seed=3, spp=1, max_depth=md, rr_depth=rr, variant=1)
        print("rr", rr, "md", md, "bad", len(b), [(int(y * W + x)) for y, x in b[:8]], "closest", c["closest_rays"], rc["closest_rays"], "shadow", c["shadow_rays"], rc["shadow_rays"])
b, img, ref, c, rc = bad(seed=3, spp=1, max_depth=7, rr_depth=3, variant=1)
for (y, x) in b[:4]:
    lane = y * W + x
    for md in (4, 5, 6, 7):
        kw = dict(seed=3, spp=1, max_depth=md, rr_depth=3, variant=1)
        film, _ = ds.render_film(lane_begin=lane, lane_end=lane + 1, **kw)
        rec, L = orc.debug_lane(lane, **kw)
        print(" lane", lane, "md", md, "gpu", film[y, x].cpu().numpy()[:3], "oracle", L, "nrays", len(rec))

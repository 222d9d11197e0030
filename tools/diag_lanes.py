"""Diagnostic (GPU box): per-lane comparison at 1 spp / box filter; for mismatching lanes replay the oracle's rays on the GPU."""
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
for variant in (1,):
  for md in (2, 3, 4, 7):
    kw = dict(seed=3, spp=1, max_depth=md, rr_depth=3, variant=variant)
    img, _ = ds.render(**kw); ref, _ = orc.render(**kw); img = img.cpu().numpy()
    e = np.abs(img - ref).max(axis=2) / np.maximum(np.abs(ref).max(axis=2), 2e-2)
    bad = np.argwhere(e > 1e-3)
    print("variant", variant, "max_depth", md, "bad lanes", len(bad), "of", W * H)
    for (y, x) in bad[:6]:
        lane = y * W + x
        rec, L = orc.debug_lane(lane, **kw)
        rays = np.concatenate([rec[:, :7], np.zeros((len(rec), 1), np.float32)], axis=1)
        g = ds.trace(rays)
        face = rec[:, 10].view(np.uint32)
        print(" lane", lane, "oracle L", L, "gpu L", img[y, x], "ref", ref[y, x])
        for i in range(len(rec)):
            flag = "" if (g["prim"][i] == face[i] and g["t"][i] == rec[i, 7]) else "   <<<< TRACE DIFFERS"
            print("   ray", i, "o", rec[i, :3], "d", rec[i, 3:6], "maxt", rec[i, 6], "| oracle t,face", rec[i, 7], face[i], "| gpu", g["t"][i], g["prim"][i], flag)

"""Diagnostic (GPU box): which scene element breaks per-sample parity with the oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from scenes_for_tests import materials_scene
from mitsuba3_experiments_b200.runtime import DeviceScene
from mitsuba3_experiments_b200.scenes import cornell_box
from oracle.oracle_py import OracleScene

def frac(d, **kw):
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    img, c = ds.render(counters=True, **kw); ref, rc = orc.render(**kw)
    img = img.cpu().numpy()
    e = np.abs(img - ref) / np.maximum(np.abs(ref), 2e-2)
    return round(float((e <= 1e-3).mean()), 5), round(float(np.minimum(e, 1).mean()), 6), {k: c[k] - rc[k] for k in ("closest_rays", "shadow_rays", "tex_evals", "nee_env", "env_escapes")}

kw = dict(seed=3, spp=16, max_depth=7, rr_depth=3, variant=0)
full = materials_scene(64, 48, 16)
print("full", frac(full, **kw))
print("no env", frac(materials_scene(64, 48, 16, envmap=False), **kw))
print("no meshlight", frac(materials_scene(64, 48, 16, mesh_light=False), **kw))
for name, target in [("metal", "large-box"), ("beck", "red-wall"), ("glass", "small-box"), ("tex", "back"), ("tex16", "floor")]:
    d = cornell_box(64, 48, 16)
    m = materials_scene(64, 48, 16)
    d[name] = m[name]; d[target]["bsdf"] = {"type": "ref", "id": name}
    print("only", name, frac(d, **kw))
d = cornell_box(64, 48, 16); m = materials_scene(64, 48, 16); del d["ceiling"]; d["env"] = m["env"]
print("only env", frac(d, **kw))
d = cornell_box(64, 48, 16); d["tetra-light"] = m["tetra-light"]
print("only meshlight", frac(d, **kw))

"""b2_trace_noalloc (include/b2path.h): the wavefront trace kernel with its lane state in the library's static device array.
Collected last on purpose.  Through the C-ABI, two calls in a row, against b2_trace on the
same rays under the tolerance of the other traversal tests: (t, u, v, prim) bit for bit.
"""
import numpy as np
import pytest

from scenes_for_tests import triangles_scene

pytestmark = pytest.mark.gpu

HIT = np.dtype([("t", "f4"), ("u", "f4"), ("v", "f4"), ("prim", "u4")])


def _rays(n, seed, extent=1.5):
    rng = np.random.RandomState(seed)
    o = rng.uniform(-extent, extent, size=(n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    maxt = np.where(rng.uniform(size=n) < 0.3, rng.uniform(0.2, 2.0, size=n), 3.4e38)
    return np.concatenate([o, d, maxt[:, None], np.zeros((n, 1))], axis=1).astype(np.float32)


def test_trace_noalloc_equals_trace_on_two_calls_in_a_row():
    import torch
    from mitsuba3_experiments_b200 import capi
    from mitsuba3_experiments_b200.runtime import DeviceScene
    ds = DeviceScene(triangles_scene(30000))
    nmax = capi.TRACE_NOALLOC_MAX_RAYS
    # second call: fewer rays than the first, so the array still holds records of the first call beyond its own
    sets = [_rays(nmax, 21), _rays(10001, 22)]
    want = [(ds.trace(r), ds.trace(r, any_hit=True)) for r in sets]
    dev = [torch.from_numpy(r).to(ds.device) for r in sets]
    out = [torch.full((r.shape[0], 4), -7.0, dtype=torch.float32, device=ds.device) for r in sets]
    out_any = [torch.full((r.shape[0], 4), -7.0, dtype=torch.float32, device=ds.device) for r in sets]
    torch.cuda.synchronize(ds.device)
    free = []
    for k in range(2):                                   # two calls in a row, nothing in between
        free.append(torch.cuda.mem_get_info(ds.device)[0])
        ds.trace_noalloc(dev[k], out[k])
    torch.cuda.synchronize(ds.device)
    free.append(torch.cuda.mem_get_info(ds.device)[0])
    assert free[2] >= free[1], f"the second call changed free device memory: {free}"       # (the first may load the module)
    for k in range(2):
        ds.trace_noalloc(dev[k], out_any[k], any_hit=True)
    torch.cuda.synchronize(ds.device)
    for k in range(2):
        g, (c, ca) = out[k].cpu().numpy().view(HIT).reshape(-1), want[k]
        assert np.array_equal(g["prim"], c["prim"]), k
        hit = c["prim"] != 0xffffffff
        assert 0.3 < hit.mean() < 1.0
        for f in ("t", "u", "v"):
            assert np.array_equal(g[f][hit], c[f][hit]), (k, f)
        ga = out_any[k].cpu().numpy().view(HIT).reshape(-1)
        assert np.array_equal(ga["prim"], ca["prim"]) and np.array_equal(ga["t"], ca["t"]), k
    # a second scene on the same device shares the array; an empty call is a no-op
    from mitsuba3_experiments_b200.scenes import cornell_box
    ds2 = DeviceScene(cornell_box(16, 16))
    r2 = _rays(5000, 23, 0.95)
    o2 = torch.empty((5000, 4), dtype=torch.float32, device=ds2.device)
    ds2.trace_noalloc(torch.from_numpy(r2).to(ds2.device), o2)
    ds2.trace_noalloc(torch.empty((0, 8), dtype=torch.float32, device=ds2.device), torch.empty((0, 4), dtype=torch.float32, device=ds2.device))
    torch.cuda.synchronize(ds2.device)
    g2, c2 = o2.cpu().numpy().view(HIT).reshape(-1), ds2.trace(r2)
    assert np.array_equal(g2["prim"], c2["prim"]) and np.array_equal(g2["t"][c2["prim"] != 0xffffffff], c2["t"][c2["prim"] != 0xffffffff])


def test_trace_noalloc_rejects_what_does_not_fit():
    import torch
    from mitsuba3_experiments_b200 import capi
    from mitsuba3_experiments_b200.runtime import DeviceScene
    from mitsuba3_experiments_b200.scenes import cornell_box
    ds = DeviceScene(cornell_box(16, 16))
    n = capi.TRACE_NOALLOC_MAX_RAYS + 1
    rays = torch.zeros((n, 8), dtype=torch.float32, device=ds.device)
    hits = torch.full((n, 4), -7.0, dtype=torch.float32, device=ds.device)
    with pytest.raises(capi.B2Error, match="exceeds B2_TRACE_NOALLOC_MAX_RAYS"):
        ds.trace_noalloc(rays, hits)
    torch.cuda.synchronize(ds.device)
    assert bool((hits == -7.0).all())                     # nothing was launched
    lib = capi.load_library()
    assert lib.b2_trace_noalloc(ds.handle, rays.data_ptr(), 16, 2, hits.data_ptr(), 0) == -2       # megakernel flag: not offered here
    assert b"any_hit" in lib.b2_last_error()

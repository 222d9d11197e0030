"""GPU parity tests: the CUDA path (through the C-ABI) against the CPU oracle on identical seeds.

Tolerances: traversal (t, u, v, prim) bit exact -- the triangle test is the same IEEE arithmetic;
images: |b2 - oracle| / max(|oracle|, 2e-2) <= 1e-3 on >= 99.9 % of values (Cornell) / >= 99 % (glossy, envmap,
triangle soup: more epsilon-sensitive paths) and mean error <= 1e-4 / 2e-3; event counters within 0.1 %;
PRB gradients: aggregate relative error <= 1e-3, per-texel <= 1e-2 on texels above the 10th |grad| percentile.
"""
import numpy as np
import pytest

from scenes_for_tests import materials_scene, triangles_scene

pytestmark = pytest.mark.gpu


def _mods():
    import torch
    from mitsuba3_experiments_b200.runtime import DeviceScene
    from mitsuba3_experiments_b200.scene import SceneDesc
    from mitsuba3_experiments_b200.scenes import cornell_box
    from oracle.oracle_py import OracleScene
    return torch, DeviceScene, SceneDesc, cornell_box, OracleScene


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 2e-2)


def check_image(a, b, frac, mean_tol):
    assert a.shape == b.shape and np.isfinite(a).all()
    e = rel_err(a, b)
    ok = float((e <= 1e-3).mean())
    assert ok >= frac, f"only {ok:.5f} of values within 1e-3 (need {frac})"
    assert float(np.minimum(e, 1.0).mean()) <= mean_tol, f"mean rel err {e.mean():.3e}"


def random_rays(n, seed, extent=1.5):
    rng = np.random.RandomState(seed)
    o = rng.uniform(-extent, extent, size=(n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    maxt = np.where(rng.uniform(size=n) < 0.3, rng.uniform(0.2, 2.0, size=n), 3.4e38)
    return np.concatenate([o, d, maxt[:, None], np.zeros((n, 1))], axis=1).astype(np.float32)


BUILDERS = [pytest.param(0, id="lbvh"), pytest.param(1, id="ploc")]


def test_sampler_stream_bit_exact():
    """a1 (integer work, bit exact): the draws a render kernel makes for a lane -- TEA(seed, lane) -> PCG32 -> next_float32
    (sampler.cpp:129-148, random.h:77-90, drjit random.h:60-79,211-213) -- dumped by the b2_debug_sampler hook equal the oracle's
    stream (itself pinned to the reference's headers and known answers) bit for bit, also for lanes near 2^32 and for sample
    shards; the sample position is pixel + the first two draws (integrator.cpp:293-310)."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from oracle.oracle_py import lib
    ds = DeviceScene(cornell_box(64, 48, spp=12))
    out = np.zeros(16, np.float32)
    for seed, lane0, spp in ((0, 0, 12), (7, 64 * 48 * 12 - 40, 12), (123456789, 1000, 16), (0xffffffff, 5, 1)):
        g, pos = ds.debug_sampler(lane0, 40, 16, seed=seed, spp=spp)
        for i in range(40):
            lib().orc_sampler_floats(seed, lane0 + i, 16, out.ctypes.data_as(__import__("ctypes").c_void_p))
            assert np.array_equal(g[i].view(np.uint32), out.view(np.uint32)), (seed, lane0 + i)
            pix = (lane0 + i) // spp
            assert pos[i, 0] == np.float32(pix % 64) + out[0] and pos[i, 1] == np.float32(pix // 64) + out[1]
    big = DeviceScene(cornell_box(4096, 4096, spp=255))           # 4096^2 x 255 lanes: the last lanes of a pass just below 2^32
    lane0 = 4096 * 4096 * 255 - 8
    g, _ = big.debug_sampler(lane0, 8, 16, seed=3, spp=255)
    for i in range(8):
        lib().orc_sampler_floats(3, lane0 + i, 16, out.ctypes.data_as(__import__("ctypes").c_void_p))
        assert np.array_equal(g[i].view(np.uint32), out.view(np.uint32))


@pytest.mark.parametrize("kernel", [pytest.param(0, id="wavefront"), pytest.param(1, id="megakernel")])
@pytest.mark.parametrize("builder", BUILDERS)
@pytest.mark.parametrize("scene_name", ["cornell", "soup"])
def test_trace_bit_exact(scene_name, builder, kernel, monkeypatch):
    """b2_trace through the production trace kernel (k_wf_trace: staged rays, vote-scheduled node / triangle steps) and through
    the megakernel's per-thread traversal: (t, u, v, prim) equal the oracle's exact BVH bit for bit."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    d = cornell_box(16, 16) if scene_name == "cornell" else triangles_scene(30000)
    ds = DeviceScene(d, builder=builder); orc = OracleScene(ds.desc)
    rays = random_rays(200000, 11, 0.95 if scene_name == "cornell" else 1.5)
    g, c = ds.trace(rays, kernel=kernel), orc.trace(rays)
    assert np.array_equal(g["prim"], c["prim"])
    hit = c["prim"] != 0xffffffff
    assert hit.mean() > 0.3
    for k in ("t", "u", "v"):
        assert np.array_equal(g[k][hit], c[k][hit]), k
    ga, ca = ds.trace(rays, any_hit=True, kernel=kernel), orc.trace(rays, any_hit=True)
    assert np.array_equal(ga["prim"], ca["prim"])
    if kernel == 0:
        # the wavefront path picks its trace kernel by tree size (k_wf_trace_simple up to 32 wide nodes, k_wf_trace above): force the
        # other one on the same rays -- identical records
        monkeypatch.setenv("B2_SIMPLE_TRACE_NODES", "0" if ds.build_stats.as_dict()["n_nodes8"] <= 32 else "1000000000")
        g2, ga2 = ds.trace(rays, kernel=kernel), ds.trace(rays, any_hit=True, kernel=kernel)
        monkeypatch.delenv("B2_SIMPLE_TRACE_NODES")
        for k in ("prim", "t", "u", "v"):
            assert np.array_equal(g2[k].view(np.uint32), g[k].view(np.uint32)), k
        assert np.array_equal(ga2["prim"], ga["prim"]) and np.array_equal(ga2["t"].view(np.uint32), ga["t"].view(np.uint32))


@pytest.mark.parametrize("scene_name", ["cornell", "materials", "soup"])
def test_primary_packet_trace_bit_exact(scene_name):
    """Bounce 0 of every render goes through k_wf_trace_packet (from 8 samples per pixel on): one warp-uniform traversal per 32 or 64 consecutive lanes, children
    culled against the interval of the packet's rays, per-ray Moeller-Trumbore.  The closest hit of a ray is the minimum of
    (t, face) over the triangles it intersects, whatever the traversal, so (t, u, v, prim) must equal the per-ray trace kernel's
    bit for bit: for whole images (packets around the image centre contain rays of both signs on an axis), for 1 .. 64 samples
    per pixel (one sample: a packet spans 64 pixels and the row ends), ragged lane ranges and sample shards."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    d = {"cornell": lambda: cornell_box(40, 32), "materials": lambda: materials_scene(56, 40, 4, envmap=True), "soup": lambda: triangles_scene(30000, 48, 36)}[scene_name]()
    ds = DeviceScene(d)
    w, h = {"cornell": (40, 32), "materials": (56, 40), "soup": (48, 36)}[scene_name]
    some_hit = False
    for spp, kw in ((1, {}), (3, {}), (16, {}), (64, {}), (64, dict(sample_begin=24, sample_count=24)), (5, dict(sample_begin=2, sample_count=3))):
        n = w * h * kw.get("sample_count", spp)
        for lane0, cnt in ((0, n), (37, min(n - 37, 1000 + 13)), (n - 65, 65), (n // 2 + 5, 1)):
            b = ds.debug_primary_hits(lane0, cnt, kernel=0, seed=5, spp=spp, **kw)
            for kernel in (1, 2):
                a = ds.debug_primary_hits(lane0, cnt, kernel=kernel, seed=5, spp=spp, **kw)
                for f in ("prim", "t", "u", "v"):
                    assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), (scene_name, spp, kw, lane0, cnt, kernel, f)
            some_hit |= bool((a["prim"] != 0xffffffff).any())
    assert some_hit


@pytest.mark.parametrize("builder", BUILDERS)
def test_bvh_stats_and_edge_cases(builder):
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from functools import partial
    DeviceScene = partial(DeviceScene, builder=builder)
    ds = DeviceScene(cornell_box(8, 8))
    st = ds.build_stats.as_dict()
    assert st["n_tris"] == 36 and st["n_nodes8"] >= 1 and st["node_bytes"] == 80 * st["n_nodes8"] and st["tri_bytes"] == 48 * 36
    # degenerate inputs: a single triangle, and many exactly coincident triangles (identical Morton codes)
    one = {"type": "scene", "sensor": {"type": "perspective", "fov": 45, "film": {"type": "hdrfilm", "width": 4, "height": 4}},
           "m": {"type": "mesh", "positions": [[-1, -1, 0], [1, -1, 0], [0, 1, 0]], "faces": [[0, 1, 2]]}}
    ds1 = DeviceScene(one); o1 = OracleScene(ds1.desc)
    rays = random_rays(5000, 5, 1.0)
    assert np.array_equal(ds1.trace(rays)["prim"], o1.trace(rays)["prim"])
    pos = np.tile(np.array([[-1, -1, 0], [1, -1, 0], [0, 1, 0]], np.float32), (50, 1))
    dup = dict(one); dup["m"] = {"type": "mesh", "positions": pos, "faces": np.arange(150).reshape(-1, 3)}
    ds2 = DeviceScene(dup); o2 = OracleScene(ds2.desc)
    g, c = ds2.trace(rays), o2.trace(rays)
    assert np.array_equal(g["prim"], c["prim"])        # ties resolve to the lowest face index on both sides
    assert (c["prim"][c["prim"] != 0xffffffff] == 0).all()


KERNELS = [pytest.param(0, id="wavefront"), pytest.param(1, id="megakernel")]


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("variant,max_depth,rr_depth", [(0, 8, 4), (1, 8, 2), (2, 6, 5), (0, 1, 4), (1, 1, 2), (0, 2, 1)])
def test_cornell_render_parity(variant, max_depth, rr_depth, kernel):
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    ds = DeviceScene(cornell_box(96, 96, spp=16)); orc = OracleScene(ds.desc)
    kw = dict(seed=0, spp=16, max_depth=max_depth, rr_depth=rr_depth, variant=variant)
    img, cnt = ds.render(counters=True, kernel=kernel, all_shadow_rays=True, **kw)    # counter parity: trace what the reference traces
    ref, rc = orc.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.999, 1e-4)
    for k in ("samples", "closest_rays", "shadow_rays", "vertices", "nee_area", "splat_taps"):
        assert abs(cnt[k] - rc[k]) <= 1e-3 * max(rc[k], 1), (k, cnt[k], rc[k])
    assert cnt["samples"] == 96 * 96 * 16 and (cnt["launches"] == 1 if kernel == 1 else cnt["launches"] >= 3)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("filt,envmap", [("gaussian", True), ("tent", False), ("box", True)])
def test_materials_render_parity(variant, filt, envmap, kernel):
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    ds = DeviceScene(materials_scene(64, 48, 16, envmap=envmap, filter_type=filt)); orc = OracleScene(ds.desc)
    kw = dict(seed=3, spp=16, max_depth=7, rr_depth=3, variant=variant)
    img, cnt = ds.render(counters=True, kernel=kernel, all_shadow_rays=True, **kw)    # counter parity: trace what the reference traces
    ref, rc = orc.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.99, 2e-3)
    for k in ("closest_rays", "shadow_rays", "vertices", "tex_evals", "nee_area", "nee_env", "env_escapes", "splat_taps"):
        assert abs(cnt[k] - rc[k]) <= 2e-3 * max(rc[k], 1) + 2, (k, cnt[k], rc[k])


@pytest.mark.parametrize("kernel", KERNELS)
def test_triangle_soup_render_parity(kernel):
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    ds = DeviceScene(triangles_scene(20000, 64, 48, 8)); orc = OracleScene(ds.desc)
    kw = dict(seed=1, spp=8, max_depth=6, rr_depth=3, variant=1)
    img, _ = ds.render(kernel=kernel, **kw)
    ref, _ = orc.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.99, 2e-3)


@pytest.mark.parametrize("variant", [0, 1, 2])
def test_wavefront_equals_megakernel(variant, monkeypatch):
    """Both kernel organisations run the same per-lane arithmetic: images agree up to the film's fp32 summation order,
    event counters agree exactly."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    ds = DeviceScene(materials_scene(64, 48, 16, envmap=True, filter_type="gaussian"))
    kw = dict(seed=9, spp=16, max_depth=6, rr_depth=2, variant=variant)
    a, ca = ds.render(counters=True, kernel=0, all_shadow_rays=True, **kw)
    b, cb = ds.render(counters=True, kernel=1, **kw)
    np.testing.assert_allclose(a.cpu().numpy(), b.cpu().numpy(), rtol=2e-5, atol=1e-6)
    # bounce 0 through the per-ray trace kernel instead of the packet kernel: same image; its traversal counters are the ones
    # comparable with the megakernel's (the packet kernel counts one node fetch per warp and the union of the warp's triangles)
    monkeypatch.setenv("B2_PRIMARY_PACKET", "0")
    a0, ca = ds.render(counters=True, kernel=0, all_shadow_rays=True, **kw)
    monkeypatch.delenv("B2_PRIMARY_PACKET")
    np.testing.assert_allclose(a0.cpu().numpy(), a.cpu().numpy(), rtol=2e-5, atol=1e-6)
    # default product path: shadow rays of emitter samples whose value is exactly zero are not traced -- same image
    e, ce = ds.render(counters=True, kernel=0, **kw)
    np.testing.assert_allclose(e.cpu().numpy(), a.cpu().numpy(), rtol=2e-5, atol=1e-6)
    assert ce["shadow_rays"] < ca["shadow_rays"] and ce["closest_rays"] == ca["closest_rays"] and ce["vertices"] == ca["vertices"]
    for k in ca:
        if k in ("node_fetches", "tri_fetches"):      # traversal work: the wavefront parks leaf triangles, so the
            assert abs(ca[k] - cb[k]) <= 0.15 * cb[k], (k, ca[k], cb[k])   # closest-hit bound shrinks a little later
        elif k != "launches":
            assert ca[k] == cb[k], (k, ca[k], cb[k])
    # unbounded depth (max_depth = -1 -> 1 << 20): the wavefront polls the live-lane count instead of a fixed trip count
    c, _ = ds.render(kernel=0, **dict(kw, max_depth=1 << 20))
    d, _ = ds.render(kernel=1, **dict(kw, max_depth=1 << 20))
    np.testing.assert_allclose(c.cpu().numpy(), d.cpu().numpy(), rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("kernel", KERNELS)
def test_envmap_alias_mode_parity_with_the_oracle(kernel):
    """The alias-table envmap mode against the oracle's restatement of it (oracle_scene.h alias_sample; the density is pinned by the
    chi^2 test of tests/test_oracle_golden.py): same seeds => same image, same event counters."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200 import capi
    ds = DeviceScene(materials_scene(64, 48, 16, envmap=True, filter_type="gaussian")); orc = OracleScene(ds.desc)
    kw = dict(seed=4, spp=16, max_depth=6, rr_depth=3, variant=1, env_warp=capi.ENV_WARP_ALIAS)
    img, cnt = ds.render(counters=True, kernel=kernel, all_shadow_rays=True, **kw)
    ref, rc = orc.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.99, 2e-3)
    for k in ("closest_rays", "shadow_rays", "vertices", "nee_env", "env_escapes"):
        assert abs(cnt[k] - rc[k]) <= 2e-3 * max(rc[k], 1) + 2, (k, cnt[k], rc[k])
    hier, _ = ds.render(kernel=kernel, **dict(kw, env_warp=capi.ENV_WARP_HIERARCHICAL))
    assert not np.array_equal(hier.cpu().numpy(), img.cpu().numpy())      # a different sample -> direction map


def test_envmap_alias_table_matches_hierarchical_in_expectation():
    """Fast envmap mode: a different (u1, u2) -> direction map with the SAME density, so only the converged image agrees
    (SURVEY.md section 0 fact 9).  1024 spp: mean relative difference of the two estimates <= 1.5 %, no bias in the mean."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200 import capi
    ds = DeviceScene(materials_scene(48, 36, 16, envmap=True, filter_type="box"))
    kw = dict(seed=2, spp=1024, max_depth=5, rr_depth=3, variant=1)
    a, _ = ds.render(env_warp=capi.ENV_WARP_HIERARCHICAL, **kw)
    b, _ = ds.render(env_warp=capi.ENV_WARP_ALIAS, **kw)
    a, b = a.cpu().numpy(), b.cpu().numpy()
    assert not np.array_equal(a, b)
    assert abs(a.mean() - b.mean()) <= 5e-3 * a.mean(), (a.mean(), b.mean())
    rel = np.abs(a - b).mean() / a.mean()
    assert rel <= 1.5e-2, rel


@pytest.mark.parametrize("kernel", KERNELS)
def test_film_edge_cases_crop_multipass_shards(kernel):
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    d = cornell_box(80, 60, spp=12)           # non power-of-two spp -> division path of integrator.cpp:293-300
    d["sensor"]["film"].update(crop_offset_x=17, crop_offset_y=9, crop_width=41, crop_height=23)
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    kw = dict(seed=5, spp=12, max_depth=5, rr_depth=2, variant=0)
    ref, _ = orc.render(**kw)
    kw["kernel"] = kernel
    img, _ = ds.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.999, 1e-4)
    # multi-pass: 3 passes x 4 spp continue each lane's stream (integrator.cpp:319-328)
    ref2, _ = orc.render(spp_per_pass=4, **kw)
    img2, _ = ds.render(spp_per_pass=4, **kw)
    check_image(img2.cpu().numpy(), ref2, 0.999, 1e-4)
    # lane shards accumulate into the same film: the union equals the full render up to fp32 summation order
    n = ds.wavefront_size(12)
    film = ds.new_film()
    cuts = [0, n // 3 + 7, n // 2, n]
    for a, b in zip(cuts[:-1], cuts[1:]):
        ds.render_film(film=film, lane_begin=a, lane_end=b, **kw)
    np.testing.assert_allclose(ds.develop(film).cpu().numpy(), img.cpu().numpy(), rtol=2e-5, atol=1e-6)
    # chunk size must not change the result beyond summation order
    img3, _ = ds.render(chunk_lanes=300, **kw)
    np.testing.assert_allclose(img3.cpu().numpy(), img.cpu().numpy(), rtol=2e-5, atol=1e-6)


def test_render_host_and_errors():
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200 import capi
    ds = DeviceScene(cornell_box(32, 32, spp=4))
    film = np.zeros((32, 32, 4), np.float32); image = np.zeros((32, 32, 3), np.float32)
    ds.render_host(film, image, seed=0, spp=4)
    img, _ = ds.render(seed=0, spp=4)
    np.testing.assert_allclose(image, img.cpu().numpy(), rtol=2e-5, atol=1e-6)
    assert np.isfinite(film).all() and (film[..., 3] > 0).all()
    # the host film is accumulated into: its upload runs on a second stream and only the splat launches wait for it
    film2 = film.copy(); ds.render_host(film2, image, seed=0, spp=4)
    np.testing.assert_allclose(film2, 2.0 * film, rtol=2e-5, atol=1e-6)
    big = DeviceScene(cornell_box(768, 768, spp=1))                       # a 9.4 MB film against a ~0.5 ms render
    base = np.random.RandomState(3).uniform(1.0, 2.0, (768, 768, 4)).astype(np.float32)
    zero = np.zeros_like(base); acc = base.copy()
    big.render_host(zero, None, seed=5, spp=1); big.render_host(acc, None, seed=5, spp=1)
    np.testing.assert_allclose(acc - base, zero, rtol=0, atol=2e-5 * float(np.abs(zero).max() + 2.0))
    for kern in (capi.KERNEL_MEGAKERNEL,):
        acc2 = base.copy(); big.render_host(acc2, None, seed=5, spp=1, kernel=kern)
        np.testing.assert_allclose(acc2 - base, zero, rtol=0, atol=2e-5 * float(np.abs(zero).max() + 2.0))
    with pytest.raises(capi.B2Error):
        ds.render(seed=0, spp=6, spp_per_pass=4)             # sample_count must be a multiple of spp_per_pass
    with pytest.raises(capi.B2Error):
        ds.set_param(0, "data", torch.zeros(3))               # constant reflectance has no .data


@pytest.mark.parametrize("kernel", ["wavefront", "megakernel"])
@pytest.mark.parametrize("textured", [False, True])
def test_prb_backward_parity(textured, kernel):
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    d = cornell_box(48, 48, spp=16)
    if textured:
        rng = np.random.RandomState(0)
        d["white"] = {"type": "diffuse", "reflectance": {"type": "bitmap", "data": rng.uniform(0.2, 0.9, (24, 32, 3)).astype(np.float32), "raw": True}}
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    rng = np.random.RandomState(1)
    grad_in = rng.normal(size=(48, 48, 3)).astype(np.float32) / (48 * 48 * 3)
    params = traverse(ds)
    keys = [k for k in params if k.startswith("white.") or k.startswith("red.")]
    params.keep(keys)
    prb = PRBIntegrator({"max_depth": 6, "rr_depth": 3, "kernel": kernel})
    prb.render_backward(ds, params, torch.from_numpy(grad_in), seed=7, spp=16, counters=True)
    torch.cuda.synchronize()
    targets = []
    for k in keys:
        i, what = params._index[k]
        targets.append((i, 1 if what == "data" else 0, np.zeros(tuple(params[k].shape), np.float32)))
    rc = orc.render_backward(grad_in, targets, seed=7, spp=16, max_depth=6, rr_depth=3)
    for k, (_, kind, ref) in zip(keys, targets):
        g = params.grad[k].cpu().numpy()
        assert np.isfinite(g).all()
        assert abs(g.sum() - ref.sum()) <= 1e-3 * np.abs(ref).sum(), (k, g.sum(), ref.sum())
        if kind == 1:
            thr = np.percentile(np.abs(ref), 10)
            m = np.abs(ref) > thr
            e = np.abs(g - ref)[m] / np.abs(ref)[m]
            assert (e <= 1e-2).mean() >= 0.99, (k, np.sort(e)[-5:])
        else:
            np.testing.assert_allclose(g, ref, rtol=1e-3)
    cnt = prb.last_counters
    if kernel == "wavefront":        # the replay reuses the visibility bits AND the closest hits recorded by the primal pass:
        cnt["shadow_rays"] *= 2      # it traces nothing, so the wavefront counts half the rays of the reference's two passes
        cnt["closest_rays"] *= 2
    for k in ("samples", "closest_rays", "shadow_rays", "grad_atomics"):
        assert abs(cnt[k] - rc[k]) <= 1e-3 * max(rc[k], 1), (k, cnt[k], rc[k])


def test_prb_gradient_matches_finite_differences():
    """The reference's own methodology (test_ad_integrators.py:966-994): d(mean image)/d(albedo) vs central FD."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    ds = DeviceScene(cornell_box(32, 32, spp=64))
    prb = PRBIntegrator({"max_depth": 5, "rr_depth": 3})
    params = traverse(ds); params.keep(["green.reflectance.value"])
    base = params["green.reflectance.value"].clone()
    grad_in = torch.ones((32, 32, 3), device=ds.device) / (32 * 32 * 3)
    prb.render_backward(ds, params, grad_in, seed=11, spp=256)
    g = params.grad["green.reflectance.value"].cpu().numpy()
    eps, fd = 1e-2, np.zeros(3)
    for c in range(3):
        vals = []
        for s in (+1, -1):
            v = base.clone(); v[c] += s * eps
            params.update({"green.reflectance.value": v})
            vals.append(float(prb.render(ds, seed=11, spp=256).mean()))
        fd[c] = (vals[0] - vals[1]) / (2 * eps)
    params.update({"green.reflectance.value": base})
    assert np.all(np.abs(g - fd) <= 0.05 * np.abs(fd) + 1e-5), (g, fd)


def test_autograd_render_op():
    """``render()`` (= mi.render, util.py:395-526) through torch.autograd: the gradient that reaches the parameter equals a
    direct ``render_backward`` with grad_in = dLoss/dImage at ``seed_grad = sample_tea_32(seed, 1)[0]`` (util.py:504-509)."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, render, sample_tea_32, traverse
    ds = DeviceScene(cornell_box(24, 24, spp=8))
    params = traverse(ds); params.keep(["red.reflectance.value"])
    params["red.reflectance.value"].requires_grad_(True)
    integ = PRBIntegrator({"max_depth": 4})
    img = render(ds, params, integrator=integ, seed=3, spp=8)
    loss = (img ** 2).mean()
    loss.backward()
    g = params["red.reflectance.value"].grad
    assert g is not None and torch.isfinite(g).all() and (g.abs() > 0).any()
    # the same by hand
    p2 = traverse(ds); p2.keep(["red.reflectance.value"])
    primal = integ.render(ds, seed=3, spp=8)
    assert torch.allclose(primal, img.detach(), rtol=2e-5, atol=1e-7)      # the film is summed with atomics: order varies
    integ.render_backward(ds, p2, 2.0 * primal / primal.numel(), seed=sample_tea_32(3, 1)[0], spp=8)
    torch.cuda.synchronize()
    np.testing.assert_allclose(g.cpu().numpy(), p2.grad["red.reflectance.value"].cpu().numpy(), rtol=1e-5)
    # and a different seed_grad gives a different (independent) estimate
    p3 = traverse(ds); p3.keep(["red.reflectance.value"])
    integ.render_backward(ds, p3, 2.0 * primal / primal.numel(), seed=3, spp=8)
    torch.cuda.synchronize()
    assert not np.allclose(g.cpu().numpy(), p3.grad["red.reflectance.value"].cpu().numpy(), rtol=1e-6)


def test_prb_sample_shards_sum_to_the_full_gradient():
    """The balanced multi-GPU split of the adjoint: render_backward on sample slices (b2_render_args.sample_begin / sample_count)
    uses the weight film of the WHOLE sample set, so the shards' gradients add up to the single-call gradient."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    d = cornell_box(48, 48, spp=16)
    rng = np.random.RandomState(0)
    d["white"] = {"type": "diffuse", "reflectance": {"type": "bitmap", "data": rng.uniform(0.2, 0.9, (24, 32, 3)).astype(np.float32), "raw": True}}
    ds = DeviceScene(d)
    grad_in = torch.from_numpy(np.random.RandomState(1).normal(size=(48, 48, 3)).astype(np.float32) / (48 * 48 * 3))
    keys = ["white.reflectance.data", "green.reflectance.value"]
    prb = PRBIntegrator({"max_depth": 6, "rr_depth": 3})
    full = traverse(ds); full.keep(keys)
    prb.render_backward(ds, full, grad_in, seed=7, spp=16)
    parts = traverse(ds); parts.keep(keys)
    for r in range(4):
        prb.render_backward(ds, parts, grad_in, seed=7, spp=16, sample_begin=4 * r, sample_count=4)
    torch.cuda.synchronize()
    for k in keys:
        a, b = parts.grad[k].cpu().numpy(), full.grad[k].cpu().numpy()
        assert np.abs(a - b).max() <= 1e-5 * np.abs(b).max(), (k, np.abs(a - b).max(), np.abs(b).max())


@pytest.mark.parametrize("kernel", KERNELS)
def test_max_depth_one_edge_cases(kernel):
    """path.py:235-239 masks the first intersection AND the primary emission with depth < max_depth: Path(max_depth = 1) is black,
    also under an environment map and on a multi-class scene (no class bytes exist when the first trace is skipped).
    path-mis.py:129-133 on its terminating iteration still samples the BSDF for valid_ray: an opacity mask reached at the last
    depth without an environment keeps the sample with probability `opacity` only."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import mirror_scene
    ds = DeviceScene(materials_scene(32, 24, 8, envmap=True, filter_type="box")); orc = OracleScene(ds.desc)
    kw = dict(seed=1, spp=8, max_depth=1, rr_depth=3, variant=0)
    img, cnt = ds.render(counters=True, kernel=kernel, **kw)
    ref, rc = orc.render(**kw)
    assert float(np.abs(img.cpu().numpy()).max()) == 0.0 and float(np.abs(ref).max()) == 0.0
    assert cnt["closest_rays"] == rc["closest_rays"] == 0 and cnt["samples"] == 32 * 24 * 8
    # mask at the final depth, no environment emitter: an emissive masked rectangle fills the view
    d = mirror_scene(16, 16, 64, opacity=0.3); del d["env"]
    d["plane"]["emitter"] = {"type": "area", "radiance": {"type": "rgb", "value": [2.0, 3.0, 4.0]}}
    d["m"] = {"type": "mask", "opacity": 0.3, "nested": {"type": "twosided", "nested": {"type": "diffuse", "reflectance": {"type": "rgb", "value": [0.5, 0.5, 0.5]}}}}
    dm = DeviceScene(d); om = OracleScene(dm.desc)
    kw = dict(seed=2, spp=64, max_depth=1, rr_depth=3, variant=1)
    img, _ = dm.render(kernel=kernel, **kw)
    ref, _ = om.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.999, 1e-4)
    m = img.cpu().numpy().mean(axis=(0, 1))
    assert np.allclose(m, 0.3 * np.array([2.0, 3.0, 4.0]), rtol=0.08), m       # E[valid_ray] = opacity


@pytest.mark.parametrize("kernel", KERNELS)
def test_smooth_conductor_known_answer_and_parity(kernel):
    """smooth conductor.  CUDA == oracle on the mirror scene, and both == the closed form
    specular_reflectance * fresnel_conductor(cos_i) * radiance (tests/test_oracle_golden.py::mirror_expected)."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import mirror_scene
    from test_oracle_golden import mirror_expected
    from oracle.oracle_py import lib
    d = mirror_scene(16, 16, 64)
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    for variant in (0, 1):
        kw = dict(seed=0, spp=64, max_depth=4, rr_depth=8, variant=variant)
        img, cnt = ds.render(counters=True, kernel=kernel, **kw)
        ref, rc = orc.render(**kw)
        check_image(img.cpu().numpy(), ref, 0.999, 1e-4)
        assert cnt["closest_rays"] == rc["closest_rays"] and cnt["shadow_rays"] == 0
        exp = mirror_expected(d, lib())
        assert np.abs(img.cpu().numpy() - exp).max() / exp.max() < 5e-3


@pytest.mark.parametrize("kernel", KERNELS)
def test_opacity_mask_parity(kernel):
    """OpacityMask (mask.cpp:123-215).  Same seeds => CUDA == oracle per pixel (the null / nested choice is
    one comparison of the same PCG32 draw); a fully transparent mask must leave only the environment."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import mirror_scene
    for op in (0.4, 0.0, 1.0):
        d = mirror_scene(16, 16, 32, opacity=op)
        ds = DeviceScene(d); orc = OracleScene(ds.desc)
        for variant in (0, 1, 2):
            kw = dict(seed=5, spp=32, max_depth=5, rr_depth=8, variant=variant)
            img, cnt = ds.render(counters=True, kernel=kernel, **kw)
            ref, rc = orc.render(**kw)
            check_image(img.cpu().numpy(), ref, 0.999, 1e-4)
            assert cnt["closest_rays"] == rc["closest_rays"]
        if op == 0.0:
            assert np.allclose(img.cpu().numpy(), np.asarray(d["env"]["radiance"]["value"])[None, None, :], rtol=1e-5)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("dist,visible", [("beckmann", False), ("ggx", True)])
def test_roughdielectric_parity(kernel, dist, visible):
    """RoughDielectric (roughdielectric.cpp:242-605), both sampling modes: CUDA == oracle on a Cornell box
    whose small box is frosted glass (NEE through the glossy lobes, transmission with eta tracking for Russian roulette)."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    d = cornell_box(64, 64, 16)
    d["frost"] = {"type": "roughdielectric", "distribution": dist, "alpha_u": 0.12, "alpha_v": 0.3, "int_ior": "bk7", "ext_ior": "air",
                  "sample_visible": visible, "specular_reflectance": [1.0, 0.9, 0.8]}
    d["small-box"]["bsdf"] = {"type": "ref", "id": "frost"}
    d["small-box"]["to_world"] = d["small-box"]["to_world"].__class__().translate([0.335, -0.62, 0.38]).rotate([0, 1, 0], -17).scale(0.3)
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    for variant in (0, 1, 2):
        kw = dict(seed=2, spp=16, max_depth=7, rr_depth=3, variant=variant)
        img, cnt = ds.render(counters=True, kernel=kernel, all_shadow_rays=True, **kw)
        ref, rc = orc.render(**kw)
        check_image(img.cpu().numpy(), ref, 0.99, 2e-3)
        for k in ("closest_rays", "shadow_rays", "vertices"):
            assert abs(cnt[k] - rc[k]) <= 2e-3 * max(rc[k], 1) + 2, (k, cnt[k], rc[k])


@pytest.mark.parametrize("kernel", KERNELS)
def test_roughplastic_parity(kernel):
    """RoughPlastic (roughplastic.cpp:216-430) as the bedroom uses it (scene.xml:108-143): GGX and Beckmann,
    nonlinear, textured and constant diffuse_reflectance.  The device-side table precomputation (tables.cu) is covered through
    the image: CUDA == oracle, whose own tables are pinned to the reference's eval_transmittance / eval_reflectance."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import value_noise_texture
    d = cornell_box(64, 64, 16)
    d["wood"] = {"type": "twosided", "nested": {"type": "roughplastic", "distribution": "ggx", "alpha": 0.15, "int_ior": 1.5, "ext_ior": 1.0,
                                                 "nonlinear": True, "diffuse_reflectance": {"type": "bitmap", "data": value_noise_texture(), "raw": True}}}
    d["panel"] = {"type": "roughplastic", "distribution": "beckmann", "alpha": 0.15, "int_ior": 1.5, "ext_ior": 1.0, "nonlinear": False,
                  "diffuse_reflectance": [0.3, 0.5, 0.2], "specular_reflectance": [0.9, 0.9, 1.0]}
    d["floor"]["bsdf"] = {"type": "ref", "id": "wood"}
    d["large-box"]["bsdf"] = {"type": "ref", "id": "panel"}
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    assert 0.0 < ds.desc.bsdfs[ds.desc.bsdf_ids["panel"]].spec_sampling_weight < 1.0
    for variant in (0, 1, 2):
        kw = dict(seed=4, spp=16, max_depth=6, rr_depth=3, variant=variant)
        img, cnt = ds.render(counters=True, kernel=kernel, all_shadow_rays=True, **kw)
        ref, rc = orc.render(**kw)
        check_image(img.cpu().numpy(), ref, 0.99, 2e-3)
        for k in ("closest_rays", "shadow_rays", "vertices", "tex_evals"):
            assert abs(cnt[k] - rc[k]) <= 2e-3 * max(rc[k], 1) + 2, (k, cnt[k], rc[k])


def test_xml_scene_parity(tmp_path):
    """a scene.xml in the bedroom's style (defaults, refs, OBJ mesh, bitmap file, the rank-1 material set)
    loaded through load_file renders like the oracle fed the same dictionary."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import load_file, make_integrator
    from test_host import _write_mini_scene
    ds = load_file(_write_mini_scene(tmp_path), spp=16, resx=48, resy=32)
    orc = OracleScene(ds.desc)
    integ = make_integrator(None, ds)                      # the <integrator> of the file: path, max_depth 5
    assert integ.max_depth == 5
    kw = dict(seed=1, spp=16, max_depth=5, rr_depth=3, variant=1)
    img, _ = ds.render(**kw)
    ref, _ = orc.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.99, 2e-3)


@pytest.mark.parametrize("kernel", KERNELS)
def test_sample_shards_sum_to_the_full_render(kernel):
    """Multi-GPU split by sample slices (b2_render_args.sample_begin / sample_count): shard lane i draws the stream of the
    global lane pixel * spp + sample, so the shards' films add up to the full render (fp32 summation order aside) and the
    event counters add up exactly; spp = 12 also covers the non-power-of-two lane -> pixel path."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    for spp, parts in ((16, 4), (12, 3)):
        ds = DeviceScene(materials_scene(48, 32, spp, envmap=True, filter_type="tent"))
        kw = dict(seed=7, spp=spp, max_depth=6, rr_depth=2, variant=1, kernel=kernel)
        full, cf = ds.render_film(counters=True, **kw)
        acc = ds.new_film(); tot = None
        n = spp // parts
        for r in range(parts):
            _, c = ds.render_film(film=acc, counters=True, sample_begin=r * n, sample_count=n, **kw)
            tot = c if tot is None else {k: tot[k] + c[k] for k in c}
        a, b = acc.cpu().numpy(), full.cpu().numpy()
        assert np.allclose(a, b, rtol=2e-5, atol=1e-5 * np.abs(b).max())
        for k in ("samples", "closest_rays", "shadow_rays", "vertices", "tex_evals", "nee_area", "nee_env", "env_escapes", "splat_taps"):
            assert tot[k] == cf[k], (k, tot[k], cf[k])
    with pytest.raises(Exception):
        ds.render_film(sample_begin=8, sample_count=8, **kw)             # exceeds the 12 samples of a pass


def test_render_forward_parity_fd_and_adjoint_consistency():
    """b2_render_forward.  (1) CUDA == oracle per pixel for a constant and a bitmap tangent; (2) == central
    finite differences of the CUDA primal render; (3) <grad_in, J t> == <J^T grad_in, t> with J^T from render_backward on the same
    sample set -- forward and adjoint differentiate the same estimator."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    from mitsuba3_experiments_b200.scenes import cornell_prb_scene
    d = cornell_prb_scene(48, 48, 16)
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    prb = PRBIntegrator({"max_depth": 5, "rr_depth": 3})
    params = traverse(ds); params.keep(["white.reflectance.data", "red.reflectance.value"])
    rng = np.random.RandomState(0)
    t_tex = rng.uniform(-1, 1, size=tuple(params["white.reflectance.data"].shape)).astype(np.float32)
    t_red = np.array([0.5, -1.0, 0.25], np.float32)
    params.set_tangent("white.reflectance.data", t_tex); params.set_tangent("red.reflectance.value", t_red)
    g = prb.render_forward(ds, params, seed=3, spp=16).cpu().numpy()
    names = ds.desc.bsdf_ids
    ref, _ = orc.render_forward([(names["white"], 1, t_tex), (names["red"], 0, t_red)], seed=3, spp=16, max_depth=5, rr_depth=3)
    assert np.abs(g - ref).max() <= 1e-3 * np.abs(ref).max() and np.abs(g - ref).mean() <= 1e-5 * np.abs(ref).max()
    # (2) finite differences of the primal render along the same direction
    base_tex = params["white.reflectance.data"].clone(); base_red = params["red.reflectance.value"].clone()
    def primal(eps):
        params.update({"white.reflectance.data": base_tex + eps * torch.from_numpy(t_tex).to(base_tex.device),
                       "red.reflectance.value": base_red + eps * torch.from_numpy(t_red).to(base_red.device)})
        return prb.render(ds, seed=3, spp=16).cpu().numpy()
    fd = (primal(1e-2) - primal(-1e-2)) / 2e-2
    params.update({"white.reflectance.data": base_tex, "red.reflectance.value": base_red})
    assert np.abs(g - fd).mean() / np.abs(fd).mean() < 5e-3
    # (3) adjoint consistency
    grad_in = torch.from_numpy(rng.uniform(-1, 1, size=g.shape).astype(np.float32)).to(base_tex.device)
    params.zero_grad()
    prb.render_backward(ds, params, grad_in, seed=3, spp=16)
    torch.cuda.synchronize()
    lhs = float((grad_in.cpu().numpy().astype(np.float64) * g).sum())
    rhs = float((params.grad["white.reflectance.data"].cpu().numpy().astype(np.float64) * t_tex).sum()
                + (params.grad["red.reflectance.value"].cpu().numpy().astype(np.float64) * t_red).sum())
    assert abs(lhs - rhs) <= 2e-3 * max(abs(lhs), abs(rhs)), (lhs, rhs)


@pytest.mark.parametrize("filter_type", ["gaussian", "tent", "box"])
@pytest.mark.parametrize("kernel", [0, 1])
def test_sample_border_crop_window_parity(kernel, filter_type):
    """film.sample_border (integrator.cpp:134-136,306-309) -- the lanes enumerate the crop window plus a
    rim of border_size pixels, the film stays crop-sized.  Both launch modes against the oracle, every filter, and the band
    split over the wider lane grid."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import crop_window_scene
    ds = DeviceScene(crop_window_scene(filter_type)); orc = OracleScene(ds.desc)
    border = {"gaussian": 2, "tent": 1, "box": 0}[filter_type]
    assert ds.film_shape == (32, 32) and ds.lane_grid == (32 + 2 * border,) * 2
    kw = dict(seed=5, spp=16, max_depth=3, rr_depth=2, variant=1)
    img, cnt = ds.render(counters=True, kernel=kernel, **kw)
    ref, rcnt = orc.render(**kw)
    check_image(img.cpu().numpy(), ref, 0.999, 1e-4)
    assert cnt["samples"] == rcnt["samples"] == ds.lane_grid[0] * ds.lane_grid[1] * 16 and cnt["splat_taps"] == rcnt["splat_taps"]
    # lane ranges over the wider grid: two halves accumulate to the whole
    n = ds.wavefront_size(16); cut = (n // 2 // 16) * 16 + 16 * 7
    film = ds.new_film()
    ds.render_film(film=film, kernel=kernel, lane_begin=0, lane_end=cut, **kw)
    ds.render_film(film=film, kernel=kernel, lane_begin=cut, lane_end=n, **kw)
    check_image(ds.develop(film).cpu().numpy(), ref, 0.999, 1e-4)


@pytest.mark.parametrize("name", ["crop_window", "diffuse_albedo", "diffuse_albedo_g_i", "area_light_radiance",
                                  "directly_visible_area_light_radiance"])
def test_reference_prb_goldens_cuda(name):
    """The reference's own stored PRB images (tests/golden/ref_ad_integrators.npz <- resources/data/tests/integrators/
    test_<config>_image_{primal,fwd}_ref.exr) under the reference's own acceptance criteria and sample counts
    (test_ad_integrators.py:750-864), through the public integrator API: primal with the path and prb integrators,
    render_forward and render_backward for the configs that differentiate a reflectance."""
    import os
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, Path, traverse
    from scenes_for_tests import ad_config_scene
    refs = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_ad_integrators.npz"))
    d, max_depth, key = ad_config_scene(name)
    ds = DeviceScene(d)
    ref, fref = refs[f"{name}_primal"], refs[f"{name}_fwd"]
    for integ in (Path({"max_depth": max_depth}), PRBIntegrator({"max_depth": max_depth})):
        img = integ.render(ds, seed=0, spp=1024).cpu().numpy()
        err = np.abs(img - ref) / np.maximum(np.abs(ref), 2e-2)
        assert err.mean() < 0.01 and err.max() < 0.5, (err.mean(), err.max())     # reference thresholds: 0.05 / 0.5
    if key is None:
        return
    prb = PRBIntegrator({"max_depth": max_depth})
    params = traverse(ds); params.keep([key])
    params.set_tangent(key, np.ones(3, np.float32))             # theta is added to every channel (ConfigBase.update :95-102)
    fwd = prb.render_forward(ds, params, seed=0, spp=1024).cpu().numpy()
    err = np.abs(fwd - fref) / np.maximum(np.abs(fref), 2e-1)
    assert err.mean() < 0.05 and err.max() < 0.5, (err.mean(), err.max())
    grad_in = torch.full(fref.shape, 0.001, device="cuda:0")
    params.zero_grad()
    prb.render_backward(ds, params, grad_in, seed=0, spp=256)
    grad, grad_ref = float(params.grad[key].sum()) / fref.size, float(fref.mean()) * 0.001
    assert abs(grad - grad_ref) / max(abs(grad_ref), 1e-3) < 0.05, (grad, grad_ref)


@pytest.mark.parametrize("name", ["roughplastic_linear", "roughplastic_nonlinear", "roughdielectric", "infinite_mirrors"])
def test_reference_render_z_tests_bsdf_spheres_cuda(name):
    """The reference's render Z-tests (test_renders.py:158-226, stored mean + variance images of bsdf_spheres/
    test_<name>_path.xml, 868 spp) through the CUDA path, both launch modes; and CUDA == oracle per pixel on the same scene
    (rough plastic / rough dielectric + area light + envmap + point light)."""
    import os
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import bsdf_sphere_scene, render_z_test
    fx = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_bsdf_spheres.npz"))
    d = bsdf_sphere_scene(name, fx)
    ds = DeviceScene(d)
    ref = fx[f"{name}_mean"]; spp = int(2e6) // (ref.shape[0] * ref.shape[1])
    for kernel in (0, 1):
        img, _ = ds.render(seed=0, spp=spp, max_depth=d["integrator"]["max_depth"], variant=0, kernel=kernel)
        ok, frac = render_z_test(img.cpu().numpy(), spp, ref, fx[f"{name}_var"])
        assert ok and frac > 0.999, (kernel, frac)
    kw = dict(seed=2, spp=16, max_depth=4, rr_depth=2, variant=1)
    img, cnt = ds.render(counters=True, all_shadow_rays=True, **kw)
    ref, rcnt = OracleScene(ds.desc).render(**kw)
    check_image(img.cpu().numpy(), ref, 0.99, 2e-3)
    for k in ("closest_rays", "shadow_rays", "nee_area", "nee_env"):
        assert abs(cnt[k] - rcnt[k]) <= 1e-3 * max(rcnt[k], 1), (k, cnt[k], rcnt[k])


def test_mitsuba_front_door():
    """The mi.register_integrator door end to end on the GPU, through the recording fake of tests/fake_mitsuba.py: the registered
    plugin renders a live scene exactly like the dictionary door, render_backward accumulates into dr.grad of the scene's own
    parameter objects (twosided key nesting included), render_forward reads the tangents, and parameters changed through
    params.update() are picked up by the next render (no stale device copy)."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    import fake_mitsuba as fm
    from mitsuba3_experiments_b200.integrator import PathIntegrator, PRBIntegrator, traverse
    from mitsuba3_experiments_b200.mitsuba_bridge import register_with_mitsuba
    register_with_mitsuba(fm)
    # ---- primal: every BSDF / emitter kind of the materials scene
    d = materials_scene(48, 36, 8, envmap=True, filter_type="gaussian")
    for v in d.values():
        if isinstance(v, dict):
            for t in (v.get("reflectance"), (v.get("nested") or {}).get("reflectance")):
                if isinstance(t, dict):
                    t.pop("wrap_mode", None)         # not observable on a live scene (see mitsuba_bridge.py)
    scene = fm.Scene(d)
    integ = fm.load_dict({"type": "b2path_mis", "max_depth": 6, "rr_depth": 2})
    img = integ.render(scene, seed=3, spp=8)
    assert isinstance(img, fm.TensorXf)
    ref = PathIntegrator({"max_depth": 6, "rr_depth": 2}).render(DeviceScene(d), seed=3, spp=8)
    check_image(img.torch().cpu().numpy(), ref.cpu().numpy(), 0.995, 1e-3)
    # ---- adjoint, forward mode and parameter refresh on the Cornell box (twosided-wrapped diffuse BSDFs)
    c = cornell_box(32, 32, spp=8)
    for k in ("white", "red", "green"):
        c[k] = {"type": "twosided", "nested": c[k]}
    live = fm.Scene(c)
    prb = fm.load_dict({"type": "b2prb", "max_depth": 4, "rr_depth": 3})
    params = fm.traverse(live)
    key = "red.brdf_0.reflectance.value"
    assert key in params
    grad_in = torch.from_numpy(np.random.RandomState(3).normal(size=(32, 32, 3)).astype(np.float32) / (32 * 32 * 3)).cuda()
    prb.render_backward(live, params, grad_in, seed=5, spp=8)
    ds = DeviceScene(c)
    p = traverse(ds); p.keep(["red.reflectance.value", "white.reflectance.value", "green.reflectance.value"])
    PRBIntegrator({"max_depth": 4, "rr_depth": 3}).render_backward(ds, p, grad_in, seed=5, spp=8)
    torch.cuda.synchronize()
    np.testing.assert_allclose(fm._dr.grad_of(params[key]), p.grad["red.reflectance.value"].cpu().numpy(), rtol=1e-5)
    np.testing.assert_allclose(fm._dr.grad_of(params["white.brdf_0.reflectance.value"]), p.grad["white.reflectance.value"].cpu().numpy(), rtol=1e-5)
    fm._dr.set_grad(params[key], [1.0, 0.5, 0.25])
    fwd = prb.render_forward(live, params, seed=5, spp=8).torch()
    p2 = traverse(ds); p2.keep(["red.reflectance.value"]); p2.set_tangent("red.reflectance.value", [1.0, 0.5, 0.25])
    fwd_ref = PRBIntegrator({"max_depth": 4, "rr_depth": 3}).render_forward(ds, p2, seed=5, spp=8)
    np.testing.assert_allclose(fwd.cpu().numpy(), fwd_ref.cpu().numpy(), rtol=2e-5, atol=1e-7)
    before = prb.render(live, seed=1, spp=8).torch().clone()
    params.update({key: [0.1, 0.7, 0.2]})                               # mutates the scene's parameter object, as Mitsuba does
    after = prb.render(live, seed=1, spp=8).torch()
    c2 = cornell_box(32, 32, spp=8); c2["red"]["reflectance"]["value"] = [0.1, 0.7, 0.2]
    fresh = PRBIntegrator({"max_depth": 4, "rr_depth": 3}).render(DeviceScene(c2), seed=1, spp=8)
    assert not torch.allclose(before, after)
    np.testing.assert_allclose(after.cpu().numpy(), fresh.cpu().numpy(), rtol=2e-5, atol=1e-7)


def test_optimisation_loop_recovers_the_woodfloor_albedo():
    """BASELINE config 3 says *optimise*: examples/optimize_woodfloor.py runs the reference's inverse-rendering loop (mi.render ->
    loss -> backward -> Adam step, util.py:395-526) through the public autograd op.  Starting from half the true albedo, the
    clean image loss must fall at least 10x and the texel error must shrink."""
    import importlib.util
    import os
    spec = importlib.util.spec_from_file_location("optimize_woodfloor", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "examples", "optimize_woodfloor.py"))
    mod = importlib.util.module_from_spec(spec); spec.loader.exec_module(mod)
    r = mod.optimize(res=96, spp=64, iters=80, lr=0.02, eval_spp=4096)
    assert r["loss_after"] * 10.0 <= r["loss_before"], (r["loss_before"], r["loss_after"])
    assert r["texel_err_after"] < 0.9 * r["texel_err_before"], (r["texel_err_before"], r["texel_err_after"])
    assert r["history"][-1] < r["history"][0]


@pytest.mark.parametrize("which", ["roughplastic_nonlinear", "roughplastic_linear", "roughdielectric", "roughplastic_textured"])
def test_prb_rough_bsdf_gradient_targets_parity(which):
    """PRB gradient targets on the bedroom's own textured BSDFs: roughplastic.diffuse_reflectance (nonlinear: d/d rho of
    rho / (1 - rho r), roughplastic.cpp:315-367; constant and bitmap) and roughdielectric.specular_reflectance -- CUDA
    render_backward / render_forward against the oracle (which is pinned to finite differences, tests/test_oracle_golden.py)."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from mitsuba3_experiments_b200.integrator import PRBIntegrator, traverse
    from test_oracle_golden import rough_target_scene
    textured = which == "roughplastic_textured"
    d, name, slot = rough_target_scene("roughplastic_nonlinear" if textured else which, res=40, spp=16)
    if textured:
        d[name]["diffuse_reflectance"] = {"type": "bitmap", "data": np.random.RandomState(0).uniform(0.2, 0.9, (20, 28, 3)).astype(np.float32), "raw": True}
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    key = f"{name}.{slot}.{'data' if textured else 'value'}"
    params = traverse(ds); assert key in params, list(params)
    params.keep([key])
    grad_in = np.random.RandomState(1).uniform(0.5, 1.5, size=(40, 40, 3)).astype(np.float32) / (40 * 40 * 3)
    prb = PRBIntegrator({"max_depth": 5, "rr_depth": 3})
    prb.render_backward(ds, params, torch.from_numpy(grad_in), seed=7, spp=16)
    torch.cuda.synchronize()
    g = params.grad[key].cpu().numpy()
    i, what = params._index[key]
    ref = np.zeros(tuple(params[key].shape), np.float32)
    orc.render_backward(grad_in, [(i, 1 if textured else 0, ref)], seed=7, spp=16, max_depth=5, rr_depth=3)
    assert np.isfinite(g).all() and np.abs(ref).sum() > 0
    # (paths through the rough dielectric are epsilon-sensitive -- refraction decisions flip in a few lanes, as in the primal
    # parity test of that BSDF -- so its aggregate is compared at 2 %)
    agg = 2e-2 if which == "roughdielectric" else 1e-3
    assert abs(g.sum() - ref.sum()) <= agg * np.abs(ref).sum(), (g.sum(), ref.sum())
    if which == "roughdielectric":
        np.testing.assert_allclose(g, ref, rtol=3e-2)
    elif textured:
        m = np.abs(ref) > np.percentile(np.abs(ref), 10)
        e = np.abs(g - ref)[m] / np.abs(ref)[m]
        assert (e <= 1e-2).mean() >= 0.99, np.sort(e)[-5:]
    else:
        np.testing.assert_allclose(g, ref, rtol=2e-3)
    # forward mode on the same target
    tan = np.full(tuple(params[key].shape), 0.5, np.float32)
    params.set_tangent(key, tan)
    fwd = prb.render_forward(ds, params, seed=7, spp=16).cpu().numpy()
    fref, _ = orc.render_forward([(i, 1 if textured else 0, tan)], seed=7, spp=16, max_depth=5, rr_depth=3)
    check_image(fwd, fref, 0.98 if which == "roughdielectric" else 0.99, 2e-3)
    # a changed parameter reaches the device (and, for rough plastic, its lobe-sampling weight): the next render equals a fresh scene
    new = torch.clamp(params[key] * 0.5 + 0.1, 0, 1)
    params.update({key: new})
    img = prb.render(ds, seed=3, spp=16).cpu().numpy()
    d2, _, _ = rough_target_scene("roughplastic_nonlinear" if textured else which, res=40, spp=16)
    if textured:
        d2[name]["diffuse_reflectance"] = {"type": "bitmap", "data": new.cpu().numpy(), "raw": True}
    else:
        d2[name][slot]["value"] = new.cpu().numpy().tolist()
    fresh = PRBIntegrator({"max_depth": 5, "rr_depth": 3}).render(DeviceScene(d2), seed=3, spp=16).cpu().numpy()
    check_image(img, fresh, 0.999, 1e-4)


def test_coplanar_glass_box_in_expectation():
    """mi.cornell_box() rests its small box exactly ON the floor; with a glass BSDF a ray inside the box that reaches that plane
    resolves the coplanar triangles by the last ulp of t, so per-sample parity is undefined there (tests/scenes_for_tests.py lifts
    the box for the per-sample tests).  On the UNMODIFIED placement the CUDA render must still agree with the oracle in expectation:
    the reference's own render acceptance test (Z-test, test_renders.py:158-226) against the oracle's mean and variance."""
    torch, DeviceScene, SceneDesc, cornell_box, OracleScene = _mods()
    from scenes_for_tests import render_z_test
    d = cornell_box(48, 40, 64)
    d["glass"] = {"type": "dielectric", "int_ior": 1.5, "ext_ior": 1.0}
    d["small-box"]["bsdf"] = {"type": "ref", "id": "glass"}            # placement untouched: bottom face coplanar with the floor
    ds = DeviceScene(d); orc = OracleScene(ds.desc)
    kw = dict(max_depth=8, rr_depth=4, variant=1)
    # mean + per-sample variance from 256 independent 16-spp CUDA batches (4 096 spp), mean of the oracle at 1 024 spp
    batches = np.stack([ds.render(seed=1000 + k, spp=16, **kw)[0].cpu().numpy() for k in range(256)])
    mean_c, var = batches.mean(axis=0), batches.var(axis=0, ddof=1) * 16.0
    mean_o, _ = orc.render(seed=7, spp=1024, **kw)
    spp_eff = 1.0 / (1.0 / 4096.0 + 1.0 / 1024.0)                       # both means are noisy
    _, frac = render_z_test(mean_c, spp_eff, mean_o, var)
    # The reference accepts at 99.75 % of the pixels with a CONVERGED variance image; here the variance comes from the same finite
    # sample (caustic paths through the glass are heavy tailed), so the bar is 99 % of the values -- a bias on the coplanar faces
    # would reject the whole footprint of the box (> 5 % of the image).
    assert frac >= 0.99, frac

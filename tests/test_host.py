"""CPU-only tests of the host layer: the C-ABI library loads and exports every declared symbol (no compute calls),
header <-> ctypes struct layouts agree, scene ingestion, integrator properties and error behaviour."""
import ctypes as C
import os
import re
import subprocess
import sys
import tempfile

import numpy as np
import pytest

from mitsuba3_experiments_b200 import capi
from mitsuba3_experiments_b200.scene import SceneDesc, parse_fov, gaussian_coeffs
from mitsuba3_experiments_b200.scenes import cornell_box

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_all_symbols():
    if not os.path.exists(capi.LIB_PATH):
        capi.build_library()
    lib = capi.load_library()
    header = open(os.path.join(ROOT, "include", "b2path.h")).read()
    declared = set(re.findall(r"\b(b2_[a-z_]+)\s*\(", header))
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.b2_abi_version() == capi.B2_ABI_VERSION


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = capi.load_library()
    ctx = C.c_void_p()
    rc = lib.b2_init(0, C.byref(ctx))
    assert rc != 0 and b"no CUDA device" in lib.b2_last_error()
    from mitsuba3_experiments_b200.runtime import DeviceScene
    with pytest.raises(capi.B2Error):
        DeviceScene(cornell_box(8, 8))


def test_struct_layouts_match_header():
    """Compile a tiny C program printing sizeof/offsetof for every ABI struct and compare with ctypes."""
    fields = {
        "b2_texture": (capi.Texture, ["data", "width", "to_uv"]),
        "b2_bsdf": (capi.Bsdf, ["type", "reflectance", "k", "ior"]),
        "b2_mesh": (capi.Mesh, ["first_vertex", "bsdf", "flags"]),
        "b2_emitter": (capi.Emitter, ["type", "radiance", "to_world", "inv_area", "cdf_offset"]),
        "b2_envmap": (capi.Envmap, ["data", "scale", "bsphere_radius", "warp_data", "level_size", "alias_idx"]),
        "b2_sensor": (capi.Sensor, ["sample_to_camera", "near_clip", "crop_height", "rfilter_radius", "gaussian_coeff"]),
        "b2_scene_desc": (capi.SceneDesc, ["abi_version", "positions", "meshes", "n_area_cdf", "envmap", "sensor"]),
        "b2_render_args": (capi.RenderArgs, ["seed", "variant", "lane_begin", "lane_end", "chunk_lanes"]),
        "b2_counters": (capi.Counters, ["samples", "launches"]),
        "b2_build_stats": (capi.BuildStats, ["ms_total", "n_nodes8", "node_bytes"]),
        "b2_hit": (capi.Hit, ["t", "prim"]),
        "b2_grad_target": (capi.GradTarget, ["bsdf", "grad"]),
    }
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "b2path.h"', 'int main(void) {']
    for s, (_, fs) in fields.items():
        src.append(f'printf("{s} %zu\\n", sizeof({s}));')
        for f in fs:
            src.append(f'printf("{s}.{f} %zu\\n", offsetof({s}, {f}));')
    src.append("return 0; }")
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "t.c"), "w").write("\n".join(src))
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(td, "t"), os.path.join(td, "t.c")], check=True)
        out = subprocess.run([os.path.join(td, "t")], capture_output=True, text=True, check=True).stdout
    for line in out.strip().splitlines():
        name, val = line.split()
        if "." in name:
            s, f = name.split(".")
            assert getattr(fields[s][0], f).offset == int(val), name
        else:
            assert C.sizeof(fields[name][0]) == int(val), name


def test_cornell_box_ingestion():
    sd = SceneDesc(cornell_box(256, 256, 16))
    assert sd.c.n_faces == 36 and sd.c.n_vertices == 72 and sd.c.n_emitters == 1 and sd.c.n_bsdfs == 3
    assert sd.emitters[0].shape_kind == capi.SHAPE_RECTANGLE
    np.testing.assert_allclose(1.0 / sd.emitters[0].inv_area, 4 * 0.23 * 0.19, rtol=1e-5)
    # rectangle triangulation (rectangle.cpp:102-110): faces (1,2,0), (1,3,2)
    first = sd.meshes[0].first_face
    assert (sd.faces[first:first + 2] - sd.meshes[0].first_vertex).tolist() == [[1, 2, 0], [1, 3, 2]]
    assert sd.spp == 16 and sd.film_shape == (256, 256)
    assert sd.sensor.rfilter == capi.RFILTER_GAUSSIAN and abs(sd.sensor.rfilter_radius - 2.0) < 1e-6


def test_parse_fov_and_filters():
    assert parse_fov({"fov": 39.3077, "fov_axis": "smaller"}, 1.0) == pytest.approx(39.3077)
    assert parse_fov({"fov": 40, "fov_axis": "y"}, 16 / 9) == pytest.approx(65.8, abs=0.2)
    c, r = gaussian_coeffs(0.5)
    assert r == 2.0 and c.dtype == np.float32 and len(c) == 10
    with pytest.raises(ValueError):
        SceneDesc({"type": "scene", "sensor": {"type": "perspective", "fov": 45, "film": {"type": "hdrfilm", "rfilter": {"type": "gaussian", "stddev": 1.0}}}})
    with pytest.raises(ValueError):
        SceneDesc({"type": "scene", "sensor": {"type": "thinlens", "fov": 45}})
    with pytest.raises(ValueError):
        SceneDesc({"type": "scene", "s": {"type": "cylinder"}, "sensor": {"type": "perspective", "fov": 45}})
    # default focal length 50 mm over the 36 x 24 mm diagonal (sensor.cpp:163-177)
    assert parse_fov({}, 1.0) == pytest.approx(34.0221, abs=1e-3) and parse_fov({"focal_length": "50mm"}, 1.5) == pytest.approx(39.5978, abs=1e-3)
    with pytest.raises(ValueError):
        parse_fov({"fov": 45, "focal_length": "35mm"}, 1.0)


def test_tessellated_sphere_and_lane_grid():
    """sphere.cpp (center / radius / to_world, uv = (phi / 2 pi, theta / pi)) as a closed, outward-oriented UV sphere with
    exact vertex normals; film.sample_border widens the lane grid by rfilter.border_size (rfilter.cpp:25)."""
    film = {"type": "hdrfilm", "width": 8, "height": 6, "sample_border": True, "rfilter": {"type": "tent"}}
    d = {"type": "scene", "s": {"type": "sphere", "center": [1, 2, 3], "radius": 0.5, "tessellation": 16},
         "sensor": {"type": "perspective", "fov": 45, "film": film}}
    sd = SceneDesc(d)
    pos, nrm, faces = sd.positions.reshape(-1, 3), sd.normals.reshape(-1, 3), sd.faces.reshape(-1, 3)
    assert len(faces) == 2 * 16 * 32 - 2 * 32
    c = np.array([1, 2, 3], np.float32)
    assert np.allclose(np.linalg.norm(pos - c, axis=1), 0.5, atol=1e-6) and np.allclose(nrm, (pos - c) / 0.5, atol=1e-5)
    p0, p1, p2 = pos[faces[:, 0]], pos[faces[:, 1]], pos[faces[:, 2]]
    ng = np.cross(p1 - p0, p2 - p0)
    assert (np.einsum("ij,ij->i", ng, (p0 + p1 + p2) / 3 - c) > 0).all()              # outward orientation, no degenerate faces
    assert 0.5 * np.linalg.norm(ng, axis=1).sum() == pytest.approx(4 * np.pi * 0.25, rel=2e-2)
    assert sd.film_shape == (6, 8) and sd.lane_grid == (8, 10)
    film["rfilter"] = {"type": "box"}
    assert SceneDesc(d).lane_grid == (6, 8)


def test_integrator_props_and_registry():
    from mitsuba3_experiments_b200 import integrator as I
    assert (I.Path().max_depth, I.Path().rr_depth) == (8, 4)                       # path.py:23-24
    assert (I.PathIntegrator().max_depth, I.PathIntegrator().rr_depth) == (8, 2)   # path-mis.py:21-22
    assert (I.PRBIntegrator().max_depth, I.PRBIntegrator().rr_depth) == (6, 5)     # common.py:28-40
    assert I.Path({"max_depth": 3, "rr_depth": 2}).max_depth == 3
    with pytest.raises(RuntimeError):
        I.Path({"rr_depth": 0})
    with pytest.raises(RuntimeError):
        I.Path({"max_depth": -2})
    with pytest.raises(RuntimeError):
        I.Path().sample(None, None, None)
    assert isinstance(I.make_integrator({"type": "path_test", "max_depth": 5}), I.PathIntegrator)
    I.register_integrator("custom", lambda props: I.Path(props))
    assert isinstance(I.make_integrator("custom"), I.Path)
    assert "max_depth = 8" in I.Path().to_string()


def test_host_tea_matches_oracle(oracle_lib):
    from mitsuba3_experiments_b200.integrator import sample_tea_32
    for v0, v1 in [(0, 1), (1, 1), (12345, 1), (0xffffffff, 7)]:
        out = np.zeros(2, np.uint32)
        oracle_lib.orc_tea32(v0, v1, 4, out.ctypes.data_as(C.c_void_p))
        assert sample_tea_32(v0, v1) == (int(out[0]), int(out[1]))


# ------------------------------------------------------------------------------------------------ XML scene files
MINI_XML = """<scene version="3.0.0">
    <default name="integrator" value="path" />
    <default name="spp" value="8" />
    <default name="resx" value="24" />
    <default name="resy" value="16" />
    <default name="max_depth" value="5" />
    <integrator type="$integrator"><integer name="max_depth" value="$max_depth" /></integrator>
    <sensor type="perspective">
        <float name="fov" value="65" />
        <transform name="to_world"><matrix value="-1 0 0 0  0 1 0 1  0 0 -1 4  0 0 0 1" /></transform>
        <sampler type="independent"><integer name="sample_count" value="$spp" /></sampler>
        <film type="hdrfilm">
            <integer name="width" value="$resx" /><integer name="height" value="$resy" />
            <rfilter type="tent" />
        </film>
    </sensor>
    <bsdf type="twosided" id="WoodBSDF">
        <bsdf type="roughplastic">
            <float name="alpha" value="0.15" /><string name="distribution" value="ggx" />
            <float name="int_ior" value="1.5" /><float name="ext_ior" value="1" /><boolean name="nonlinear" value="true" />
            <texture name="diffuse_reflectance" type="bitmap"><string name="filename" value="textures/wood.png" /><string name="filter_type" value="bilinear" /></texture>
        </bsdf>
    </bsdf>
    <bsdf type="mask" id="CurtainsBSDF">
        <float name="opacity" value="0.531049" />
        <bsdf type="twosided"><bsdf type="diffuse"><rgb name="reflectance" value="1, 1, 1" /></bsdf></bsdf>
    </bsdf>
    <bsdf type="roughdielectric" id="LampGlassBSDF">
        <float name="alpha" value="0.1" /><string name="distribution" value="beckmann" /><float name="int_ior" value="1.5" /><float name="ext_ior" value="1" />
    </bsdf>
    <bsdf type="twosided" id="SteelBSDF"><bsdf type="conductor"><rgb name="eta" value="1.65, 0.88, 0.52" /><rgb name="k" value="9.2, 6.3, 4.8" /></bsdf></bsdf>
    <shape type="obj" id="Floor_0001">
        <string name="filename" value="models/quad.obj" />
        <transform name="to_world"><scale value="3" /><rotate x="1" angle="-90" /><translate y="0.0" /></transform>
        <ref id="WoodBSDF" />
    </shape>
    <shape type="rectangle" id="Curtain">
        <transform name="to_world"><scale x="0.8" y="0.8" z="1" /><translate x="-0.5" y="1" z="0.5" /></transform>
        <ref id="CurtainsBSDF" />
    </shape>
    <shape type="cube" id="Lamp">
        <transform name="to_world"><scale value="0.25" /><translate x="0.8" y="0.6" z="0.2" /></transform>
        <ref id="LampGlassBSDF" />
    </shape>
    <shape type="cube" id="Block">
        <transform name="to_world"><scale value="0.2" /><translate x="0.1" y="0.2" z="-0.6" /></transform>
        <ref id="SteelBSDF" />
    </shape>
    <shape type="rectangle">
        <transform name="to_world"><matrix value="0.5 0 0 0  0 0 -0.5 2.5  0 0.5 0 0  0 0 0 1" /></transform>
        <bsdf type="twosided"><bsdf type="diffuse"><rgb name="reflectance" value="0, 0, 0" /></bsdf></bsdf>
        <emitter type="area"><rgb name="radiance" value="16.4648, 16.4648, 16.4648" /></emitter>
    </shape>
</scene>
"""


def _write_mini_scene(tmp_path):
    from PIL import Image
    (tmp_path / "models").mkdir(); (tmp_path / "textures").mkdir()
    (tmp_path / "models" / "quad.obj").write_text(
        "v -1 -1 0\nv 1 -1 0\nv 1 1 0\nv -1 1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 0 1\nf 1/1/1 2/2/1 3/3/1 4/4/1\n")
    rng = np.random.RandomState(0)
    Image.fromarray((rng.uniform(40, 220, size=(8, 8, 3))).astype(np.uint8)).save(tmp_path / "textures" / "wood.png")
    (tmp_path / "scene.xml").write_text(MINI_XML)
    return str(tmp_path / "scene.xml")


def test_xml_loader_subset(tmp_path):
    """xml.cpp subset: <default> / $param with overrides, <ref>, nested plugins, property elements, <transform> composition."""
    from mitsuba3_experiments_b200 import capi
    from mitsuba3_experiments_b200.scene import SceneDesc, T
    from mitsuba3_experiments_b200.xml_loader import XMLError, load_file, load_string
    path = _write_mini_scene(tmp_path)
    d = load_file(path, spp=4, resx=32)
    assert d["type"] == "scene" and d["integrator"] == {"type": "path", "max_depth": 5}
    assert d["sensor"]["film"]["width"] == 32 and d["sensor"]["film"]["height"] == 16 and d["sensor"]["sampler"]["sample_count"] == 4
    assert d["sensor"]["film"]["rfilter"] == {"type": "tent"}
    assert d["Floor_0001"]["bsdf"] == {"type": "ref", "id": "WoodBSDF"} and d["CurtainsBSDF"]["opacity"] == pytest.approx(0.531049)
    # transform ops compose in document order: scale, then rotate, then translate
    want = T().translate([0.8, 0.6, 0.2]).scale(0.25).matrix
    assert np.allclose(d["Lamp"]["to_world"], want)
    want = T().rotate([1, 0, 0], -90).scale(3).matrix
    assert np.allclose(d["Floor_0001"]["to_world"], want)
    with pytest.raises(XMLError):
        load_file(path, no_such_parameter=1)
    with pytest.raises(XMLError):
        load_string('<scene version="3.0.0"><integrator type="$undefined"/></scene>')
    # the dictionary feeds the same scene ingestion as mi.cornell_box()-style dictionaries
    desc = SceneDesc(d, base_path=str(tmp_path))
    names = desc.bsdf_names
    wood = desc.bsdfs[names.index("WoodBSDF")]
    assert wood.type == capi.BSDF_ROUGHPLASTIC and wood.flags & capi.BSDF_FLAG_TWOSIDED and wood.reflectance_tex >= 0 and wood.nonlinear == 1
    cur = desc.bsdfs[names.index("CurtainsBSDF")]
    assert cur.type == capi.BSDF_DIFFUSE and cur.flags == (capi.BSDF_FLAG_MASK | capi.BSDF_FLAG_TWOSIDED) and cur.opacity == pytest.approx(0.531049)
    assert desc.bsdfs[names.index("LampGlassBSDF")].type == capi.BSDF_ROUGHDIELECTRIC
    assert desc.bsdfs[names.index("SteelBSDF")].type == capi.BSDF_CONDUCTOR
    assert desc.c.n_faces == 2 + 2 + 12 + 12 + 2 and desc.c.n_emitters == 1
    assert desc.integrator_props["type"] == "path" and desc.integrator_props["max_depth"] == 5


def test_xml_scene_renders_like_its_dictionary(tmp_path):
    """The XML route and the hand-written dictionary describe one scene: identical oracle images."""
    from mitsuba3_experiments_b200.scene import SceneDesc
    from mitsuba3_experiments_b200.xml_loader import load_file
    from oracle.oracle_py import OracleScene
    path = _write_mini_scene(tmp_path)
    d = load_file(path)
    img, cnt = OracleScene(SceneDesc(d, base_path=str(tmp_path))).render(seed=0, spp=8, max_depth=5, rr_depth=3, variant=1)
    assert np.isfinite(img).all() and img.mean() > 1e-3 and cnt["tex_evals"] > 0
    d2 = {k: v for k, v in d.items()}
    d2["Lamp"] = dict(d["Lamp"]); d2["Lamp"]["bsdf"] = dict(d["LampGlassBSDF"])          # inline instead of <ref>
    img2, _ = OracleScene(SceneDesc(d2, base_path=str(tmp_path))).render(seed=0, spp=8, max_depth=5, rr_depth=3, variant=1)
    assert np.array_equal(img, img2)


def test_xml_include_search_path_and_moment(tmp_path):
    """xml.cpp subset, continued: <include> splices a fragment (its <default>s included), <path> adds a search directory for
    relative filenames, the render tests' `moment` wrapper resolves to the nested integrator, wavelength-list spectra raise."""
    from mitsuba3_experiments_b200.scene import SceneDesc
    from mitsuba3_experiments_b200.xml_loader import XMLError, load_file, load_string
    (tmp_path / "common" / "meshes").mkdir(parents=True)
    (tmp_path / "common" / "meshes" / "quad.obj").write_text("v -1 -1 0\nv 1 -1 0\nv -1 1 0\nv 1 1 0\nf 2 3 1\nf 2 4 3\n")
    (tmp_path / "scenes" / "fragments").mkdir(parents=True)
    (tmp_path / "scenes" / "fragments" / "camera.xml").write_text("""<scene version='2.0.0'>
        <default name="spp" value="16"/>
        <integrator type="moment"><integrator type="path"><integer name="max_depth" value="3"/></integrator></integrator>
        <sensor type="perspective"><float name="fov" value="40"/>
            <transform name="to_world"><lookat origin="0, 0, 4" target="0, 0, 0" up="0, 1, 0"/></transform>
            <film type="hdrfilm"><integer name="width" value="8"/><integer name="height" value="8"/></film>
            <sampler type="independent"><integer name="sample_count" value="$spp"/></sampler></sensor>
        </scene>""")
    main = tmp_path / "scenes" / "main.xml"
    main.write_text("""<scene version='2.0.0'><path value="../common"/>
        <shape type="obj"><string name="filename" value="meshes/quad.obj"/><bsdf type="diffuse"/></shape>
        <emitter type="constant"/>
        <include filename="fragments/camera.xml"/></scene>""")
    d = load_file(str(main), spp=4)
    shape = next(v for v in d.values() if isinstance(v, dict) and v.get("type") == "obj")
    assert os.path.isabs(shape["filename"]) and shape["filename"].endswith(os.path.join("common", "meshes", "quad.obj"))
    sd = SceneDesc(d, base_path=str(tmp_path / "scenes"))
    assert sd.integrator_props == {"max_depth": 3, "type": "path"} and sd.spp == 4 and len(sd.faces) == 2
    with pytest.raises(XMLError):
        load_string("<scene version='2.0.0'><include filename='x.xml'/></scene>")
    with pytest.raises(XMLError):
        load_string("<scene version='2.0.0'><bsdf type='diffuse'><spectrum name='reflectance' value='400:0.5, 700:0.5'/></bsdf></scene>")


REF_SCENES = "/root/reference/mitsuba3/resources/data/tests/scenes"


@pytest.mark.skipif(not os.path.exists(REF_SCENES), reason="the reference tree is only present in the build container")
def test_reference_scene_file_passes_its_own_render_test():
    """End to end in the build container: the reference's infinite_mirrors/scene_path.xml through load_file (its <path>,
    the `moment` wrapper, meshes/sphere.obj, textures/museum.exr) and the oracle passes the reference's own Z-test."""
    from mitsuba3_experiments_b200.scene import SceneDesc
    from mitsuba3_experiments_b200.xml_loader import load_file
    from oracle.oracle_py import OracleScene
    from scenes_for_tests import render_z_test
    path = os.path.join(REF_SCENES, "infinite_mirrors", "scene_path.xml")
    sd = SceneDesc(load_file(path, spp=8), base_path=os.path.dirname(path))
    assert len(sd.faces) == 3 * 32512 and sd.integrator_props["max_depth"] == 8
    fx = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_bsdf_spheres.npz"))
    img, _ = OracleScene(sd).render(seed=0, spp=1953, max_depth=8, variant=0)
    ok, frac = render_z_test(img, 1953, fx["infinite_mirrors_mean"], fx["infinite_mirrors_var"])
    assert ok and frac > 0.999


def test_minimal_exr_reader_matches_openexr(tmp_path):
    """tests/golden/exr_min.py (the reader that extracted the reference's stored images) against OpenCV's bundled OpenEXR:
    PIZ, ZIP and uncompressed files, float32 and half, sizes that are not multiples of the 32-line PIZ chunk or of 2."""
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    cv2 = pytest.importorskip("cv2")
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
    import exr_min
    rng = np.random.RandomState(0)
    yy, xx = np.mgrid[0:71, 0:53]
    smooth = np.stack([np.sin(xx * 0.11) * np.cos(yy * 0.07), (xx + yy) / 120.0, np.exp(-((xx - 20) ** 2 + (yy - 30) ** 2) / 300.0)], -1)
    img = (smooth + 0.01 * rng.standard_normal(smooth.shape)).astype(np.float32)
    img[5:9, 7:30] = 0.0                                           # runs of equal values exercise the Huffman run-length symbol
    cases = [(cv2.IMWRITE_EXR_COMPRESSION_PIZ, cv2.IMWRITE_EXR_TYPE_FLOAT), (cv2.IMWRITE_EXR_COMPRESSION_PIZ, cv2.IMWRITE_EXR_TYPE_HALF),
             (cv2.IMWRITE_EXR_COMPRESSION_ZIP, cv2.IMWRITE_EXR_TYPE_FLOAT), (cv2.IMWRITE_EXR_COMPRESSION_NO, cv2.IMWRITE_EXR_TYPE_FLOAT)]
    for i, (comp, typ) in enumerate(cases):
        path = str(tmp_path / f"t{i}.exr")
        if not cv2.imwrite(path, img[:, :, ::-1], [cv2.IMWRITE_EXR_COMPRESSION, comp, cv2.IMWRITE_EXR_TYPE, typ]):
            pytest.skip("this OpenCV build cannot write EXR")
        want = cv2.imread(path, cv2.IMREAD_UNCHANGED)[:, :, ::-1]
        got = exr_min.read_rgb(path)
        assert got.shape == want.shape and np.array_equal(got, want), (comp, typ)


# ---------------------------------------------------------------------------------------------- mi.register_integrator front door
def _bridge_scenes():
    from scenes_for_tests import materials_scene, mirror_scene
    from mitsuba3_experiments_b200.scenes import bedroom_standin, cornell_box
    d = cornell_box(24, 24, 4)
    d["sensor"]["film"].update(sample_border=True, crop_offset_x=3, crop_offset_y=2, crop_width=16, crop_height=20)
    m = mirror_scene(16, 16, 4, opacity=0.4)
    m["lamp"] = {"type": "point", "position": [0.5, 2.0, 1.0], "intensity": {"type": "rgb", "value": [3.0, 2.0, 1.0]}}
    mats = materials_scene(24, 18, 4, envmap=True, filter_type="tent")
    for v in mats.values():        # a bitmap's wrap_mode / filter_type are not observable from Python (bitmap.cpp:396-399 exposes data + to_uv
        if isinstance(v, dict):    # only): the front door assumes the plugin defaults (repeat, bilinear), as the bedroom scene uses
            for t in (v.get("reflectance"), (v.get("nested") or {}).get("reflectance")):
                if isinstance(t, dict):
                    t.pop("wrap_mode", None)
    return {"cornell": d, "materials": mats, "mask_constant_point": m,
            "bedroom_xml": bedroom_standin(32, 18, 2, n_tris=40_000, tex_res=32, palette="scene_xml")}


@pytest.mark.parametrize("which", ["cornell", "materials", "mask_constant_point", "bedroom_xml"])
def test_mitsuba_bridge_scene_extraction_round_trip(which):
    """scene_to_dict on a live scene (the recording fake of tests/fake_mitsuba.py presents the dictionary the way Mitsuba 3.6
    does: Mesh buffers, per-plugin traverse keys, class_name(), to_string()) reproduces the scene: same BSDF table
    (twosided, mask, rough plastic / dielectric, conductor, textures), emitters (area, envmap, constant, point), sensor / film
    (crop, sample_border, filter) -- and the CPU oracle renders the identical image from both."""
    import fake_mitsuba as fm
    from mitsuba3_experiments_b200.mitsuba_bridge import scene_to_dict
    from oracle.oracle_py import OracleScene
    d = _bridge_scenes()[which]
    a = SceneDesc(d)
    b = SceneDesc(scene_to_dict(fm.Scene(d), fm))
    assert a.c.n_faces == b.c.n_faces and a.c.n_emitters == b.c.n_emitters and len(a.meshes) == len(b.meshes)
    def per_mesh(s):      # (a live scene only shows the BSDFs that shapes use: compare what every mesh is shaded with)
        out = {}
        for name, m in zip(s.mesh_names, s.meshes):
            x = s.bsdfs[m.bsdf]
            out[name] = ((x.type, x.flags, x.distribution, x.sample_visible, x.nonlinear, round(x.alpha_u, 6), round(x.ior, 5), round(x.opacity, 6),
                        tuple(np.round(x.reflectance[:], 6)), x.reflectance_tex >= 0, tuple(np.round(x.eta[:], 5)), tuple(np.round(x.k[:], 5))), m.n_faces, m.emitter >= 0)
        return out
    assert per_mesh(a) == per_mesh(b)
    assert sorted(x.type for x in a.emitters) == sorted(x.type for x in b.emitters)
    for f, _ in a.sensor._fields_:             # x_fov crosses as fp32 (perspective.cpp:157): the projection agrees to rounding
        va, vb = np.asarray(getattr(a.sensor, f)[:] if hasattr(getattr(a.sensor, f), "__len__") else getattr(a.sensor, f)), \
                 np.asarray(getattr(b.sensor, f)[:] if hasattr(getattr(b.sensor, f), "__len__") else getattr(b.sensor, f))
        np.testing.assert_allclose(vb, va, rtol=2e-6, atol=1e-7, err_msg=f)
    kw = dict(seed=3, spp=4, max_depth=5, rr_depth=3, variant=1)
    ia, _ = OracleScene(a).render(**kw)
    ib, _ = OracleScene(b).render(**kw)
    e = np.abs(ib - ia) / np.maximum(np.abs(ia), 2e-2)
    assert (e <= 1e-3).mean() >= 0.995, (e <= 1e-3).mean()      # (a 1-ulp camera ray difference moves a few edge samples)


def test_mitsuba_bridge_rejects_what_it_cannot_export():
    import fake_mitsuba as fm
    from mitsuba3_experiments_b200.mitsuba_bridge import bsdf_params_to_dict, scene_to_dict
    from mitsuba3_experiments_b200.scenes import cornell_box
    with pytest.raises(RuntimeError, match="outside the hot-path scope"):
        bsdf_params_to_dict({"base_color.value": np.ones(3)}, "Principled[]", "p")
    s = fm.Scene(cornell_box(8, 8, 1))

    class AnalyticSphere(fm.Object):            # a shape that is not a Mesh subclass: no *_buffer() methods (sphere.cpp)
        _cn = "Sphere"

        def is_mesh(self):
            return False

        def is_emitter(self):
            return False

    s._shapes.append(AnalyticSphere("ball"))
    with pytest.raises(RuntimeError, match="not a triangle mesh"):
        scene_to_dict(s, fm)


def test_mitsuba_bridge_registration_and_key_mapping():
    """register_with_mitsuba registers the three plugin names through mi.register_integrator (scene_v.cpp:166-171); the plugin
    manager's constructor call builds a SamplingIntegrator subclass that carries the reference's properties; parameter keys of a
    live scene ('<id>.brdf_0.reflectance.value' under twosided, 'nested_bsdf.' under mask) map onto ours."""
    import fake_mitsuba as fm
    from mitsuba3_experiments_b200.mitsuba_bridge import map_param_keys, register_with_mitsuba
    fm.calls.clear()
    classes = register_with_mitsuba(fm)
    assert [c for c in fm.calls if c[0] == "register_integrator"] == [("register_integrator", n) for n in ("b2path", "b2path_mis", "b2prb")]
    integ = fm.load_dict({"type": "b2prb", "max_depth": 5, "rr_depth": 3})
    assert isinstance(integ, fm.SamplingIntegrator) and type(integ) is classes["b2prb"]
    assert integ._impl.max_depth == 5 and integ._impl.rr_depth == 3 and "PRBIntegrator" in integ.to_string() and integ.aov_names() == []
    assert fm.load_dict({"type": "b2path_mis"})._impl.rr_depth == 2          # path-mis.py:22 default
    with pytest.raises(RuntimeError, match="render"):
        integ.sample(None, None, None)
    theirs = {"white.brdf_0.reflectance.value": 1, "curtain.nested_bsdf.brdf_0.reflectance.value": 2, "floor.bsdf.reflectance.data": 3,
              "wood.brdf_0.diffuse_reflectance.data": 4, "sensor.x_fov": 5}
    ours = ["white.reflectance.value", "curtain.reflectance.value", "floor.bsdf.reflectance.data", "wood.reflectance.data", "absent.reflectance.value"]
    assert map_param_keys(ours, theirs) == {"white.reflectance.value": "white.brdf_0.reflectance.value",
                                            "curtain.reflectance.value": "curtain.nested_bsdf.brdf_0.reflectance.value",
                                            "floor.bsdf.reflectance.data": "floor.bsdf.reflectance.data",
                                            "wood.reflectance.data": "wood.brdf_0.diffuse_reflectance.data"}

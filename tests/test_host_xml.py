"""Scene XML subset loader of the C++ host (host/vrl_xml.cpp, driven through `vrl_host xml-info` / `render-xml`).
CPU part: a synthetic scene exercising <default>/$param, <include>, transforms, obj / rectangle / cube / sphere shapes,
homogeneous + nonlinear media, bsdf / medium / emitter references, the perspective sensor (against scenes.make_camera)
and the integrator's property errors (xml.cpp:1050-1068 behaviour through VRLIntegrator's constructor); the reference's
own fixture scenes where /root/reference is mounted.  GPU part: `render-xml` end to end against the Python front-end."""
import json
import os
import subprocess

import numpy as np
import pytest

from mitsuba_nlvrl_b200 import abi, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host", "vrl_host")
REF_SCENES = "/root/reference/data/scenes/cbox"

MAIN_XML = """<?xml version="1.0" encoding="utf-8"?>
<!-- synthetic fixture -->
<scene version="2.0.0">
    <include filename="frag/bsdfs.xml"/>
    <default name="spp" value="2"/>
    <default name="res" value="64"/>
    <default name="resy" value="48"/>
    <integrator type="vrl">
        <integer name="target_vrls" value="300"/>
        <integer name="samples_per_query" value="1"/>
        <boolean name="use_light_cut" value="false"/>
        <float name="min_vrl_length" value="0.5"/>
        {extra}
    </integrator>
    <sensor type="perspective">
        <string name="fov_axis" value="smaller"/>
        <float name="near_clip" value="0.1"/>
        <float name="far_clip" value="100"/>
        <float name="fov" value="40"/>
        <transform name="to_world">
            <lookat origin="0, 1, -30" target="0, 0, 0" up="0, 1, 0"/>
        </transform>
        <sampler type="independent"><integer name="sample_count" value="$spp"/><integer name="seed" value="7"/></sampler>
        <film type="hdrfilm"><integer name="width" value="$res"/><integer name="height" value="$resy"/><rfilter type="box"/></film>
    </sensor>
    <medium id="fog" type="nonlinear">
        <float name="scale" value="0.5"/>
        <rgb name="sigma_t" value="0.2, 0.22, 0.24"/>
        <rgb name="albedo" value="0.9"/>
        <phase type="hg"><float name="g" value="0.3"/></phase>
        <integer name="res_x" value="1"/><integer name="res_y" value="64"/><integer name="res_z" value="1"/>
        <float name="top_ior" value="0.9"/><float name="bottom_ior" value="1.0"/>
    </medium>
    <emitter type="area" id="lamp"><rgb name="radiance" value="10, 9, 8"/></emitter>
    <shape type="cube">
        <bsdf type="null"/>
        <transform name="to_world"><scale x="8" y="6" z="8"/><translate y="1"/></transform>
        <ref id="fog" name="interior"/>
    </shape>
    <shape type="rectangle">
        <transform name="to_world"><rotate x="1" angle="90"/><scale value="2"/><translate x="0" y="6.5" z="0"/></transform>
        <ref id="white"/>
        <ref id="lamp"/>
        <ref id="fog" name="exterior"/>
    </shape>
    <shape type="obj">
        <string name="filename" value="meshes/tri.obj"/>
        <ref id="white"/>
        <transform name="to_world"><translate x="1" y="-2" z="3"/></transform>
    </shape>
    <shape type="sphere">
        <point name="center" x="2" y="0" z="1"/>
        <float name="radius" value="1.5"/>
        <bsdf type="diffuse"/>
    </shape>
</scene>
"""


@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    return HOST


def _write_scene(tmp, extra=""):
    os.makedirs(os.path.join(tmp, "frag"), exist_ok=True)
    os.makedirs(os.path.join(tmp, "meshes"), exist_ok=True)
    with open(os.path.join(tmp, "frag", "bsdfs.xml"), "w") as f:
        f.write('<scene version="2.0.0"><bsdf type="diffuse" id="white"><rgb name="reflectance" value="0.8"/></bsdf></scene>')
    with open(os.path.join(tmp, "meshes", "tri.obj"), "w") as f:
        f.write("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    path = os.path.join(tmp, "scene.xml")
    with open(path, "w") as f:
        f.write(MAIN_XML.format(extra=extra))
    return path


def _info(host, path, *args):
    r = subprocess.run([host, "xml-info", path, *args], capture_output=True, text=True)
    return r.returncode, (json.loads(r.stdout) if r.returncode == 0 else None), r.stderr


def test_synthetic_scene_loads(host, tmp_path):
    path = _write_scene(str(tmp_path))
    rc, info, err = _info(host, path)
    assert rc == 0, err
    assert info["triangles"] == 12 + 2 + 1 + (2 * 32 - 2) * 64 and info["null_triangles"] == 12 and info["interior_triangles"] == 12
    assert info["emitter_triangles"] == 2 and info["emitters"] == [[10.0, 9.0, 8.0]]
    assert np.allclose(info["bbox"], [-8, -5, -8, 8, 7, 8])             # the scaled, lifted cube
    m = info["media"][0]
    assert (m["id"], m["kind"], m["phase"]) == ("fog", abi.MEDIUM_NONLINEAR, abi.PHASE_HG)
    assert np.allclose([m["g"], m["scale"], m["top_ior"], m["bottom_ior"]], [0.3, 0.5, 0.9, 1.0])
    assert np.allclose(m["sigma_t"], [0.2, 0.22, 0.24]) and np.allclose(m["albedo"], [0.9] * 3)
    assert m["nl_res"] == [1, 64, 1] and np.allclose(m["nl_bbox"], [-8, -5, -8, 8, 7, 8])   # bbox of the shape it is interior to
    s = info["sensor"]
    assert (s["width"], s["height"], s["spp"], s["seed"], s["medium"]) == (64, 48, 2, 7, -1)
    cam = scenes.make_camera(64, 48, (0, 1, -30), (0, 0, 0), fov=40, fov_axis="smaller", near_clip=0.1, far_clip=100)
    assert np.allclose(s["sample_to_camera"], list(cam.sample_to_camera), rtol=2e-6, atol=1e-7)
    assert np.allclose(s["to_world"], list(cam.to_world), rtol=2e-6, atol=1e-7)
    assert "tessellated" in err                                         # the analytic sphere is reported
    # what the light tracer reads of the BSDFs: the obj triangle is "white" (0.8), the sphere a default diffuse (0.5), the
    # emitter rectangle and the null cube reflect nothing
    n_sphere = (2 * 32 - 2) * 64
    assert info["reflecting_triangles"] == 1 + n_sphere and info["twosided_triangles"] == 0
    assert np.allclose(info["mean_reflectance"], [(0.8 + 0.5 * n_sphere) / (1 + n_sphere)] * 3, rtol=1e-6)
    # <default> can be overridden from the command line (-Dname=value, mitsuba.cpp:137-146)
    rc, info, err = _info(host, path, "-Dres=128", "-Dspp=5")
    assert rc == 0 and info["sensor"]["width"] == 128 and info["sensor"]["spp"] == 5 and info["sensor"]["height"] == 48


def test_transform_order_and_rectangle(host, tmp_path):
    """<rotate x=1 angle=90> then <scale 2> then <translate y=6.5>: the unit rectangle (z = 0 plane) becomes a 4 x 4 quad in
    the plane y = 6.5 -- operations apply in document order (xml.cpp:806-905)."""
    path = _write_scene(str(tmp_path))
    txt = open(path).read()
    only = txt[:txt.index("<shape type=\"cube\">")] + txt[txt.index("<shape type=\"rectangle\">"):txt.index("<shape type=\"obj\">")] + "</scene>"
    only = only.replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
    p2 = os.path.join(str(tmp_path), "rect.xml")
    open(p2, "w").write(only)
    rc, info, err = _info(host, p2)
    assert rc == 0, err
    assert info["triangles"] == 2 and np.allclose(info["bbox"], [-2, 6.5, -2, 2, 6.5, 2], atol=1e-6)


def test_bsdf_reflectance_forms(host, tmp_path):
    """twosided adapter, rgb reflectance, wavelength table (band means), delta BSDF (no reflection), nested by reference."""
    xml = """<scene version="2.0.0">
    <sensor type="perspective"><float name="fov" value="40"/><film type="hdrfilm"/></sensor>
    <medium id="m" type="homogeneous"/>
    <bsdf type="diffuse" id="inner"><rgb name="reflectance" value="0.1, 0.2, 0.3"/></bsdf>
    <bsdf type="twosided" id="both"><ref id="inner"/></bsdf>
    <shape type="rectangle"><ref id="both"/></shape>
    <shape type="rectangle"><bsdf type="diffuse"><spectrum name="reflectance" value="400:0.2, 450:0.4, 500:0.6, 550:0.6, 600:0.9, 700:0.7"/></bsdf></shape>
    <shape type="rectangle"><bsdf type="dielectric"><string name="int_ior" value="water"/><float name="ext_ior" value="1.0"/></bsdf></shape>
    <shape type="rectangle"><bsdf type="conductor"><rgb name="specular_reflectance" value="0.9, 0.8, 0.7"/></bsdf></shape>
    </scene>"""
    p = os.path.join(str(tmp_path), "bsdfs.xml")
    open(p, "w").write(xml)
    rc, info, err = _info(host, p)
    assert rc == 0, err
    assert info["triangles"] == 8 and info["reflecting_triangles"] == 4 and info["twosided_triangles"] == 2
    assert np.allclose(info["mean_reflectance"], [(0.1 + 0.8) / 2, (0.2 + 0.6) / 2, (0.3 + 0.3) / 2], rtol=1e-6)
    assert "rfilter type=\"gaussian\" (the film's default)" in err         # no <rfilter> in this film: Mitsuba would filter with a gaussian
    assert "wavelength table" in err and info["dielectric_triangles"] == 2 and info["mirror_triangles"] == 2


def test_bsdf_parameter_errors(host, tmp_path):
    """An index of refraction is a number or one of Mitsuba's material names; anything else fails like the reference's lookup."""
    xml = """<scene version="2.0.0">
    <sensor type="perspective"><float name="fov" value="40"/><film type="hdrfilm"/></sensor>
    <medium id="m" type="homogeneous"/>
    <shape type="rectangle"><bsdf type="dielectric"><string name="int_ior" value="{ior}"/></bsdf></shape>
    </scene>"""
    p = os.path.join(str(tmp_path), "ior.xml")
    open(p, "w").write(xml.format(ior="unobtainium"))
    rc, info, err = _info(host, p)
    assert rc != 0 and "unknown index of refraction" in err and "unobtainium" in err
    open(p, "w").write(xml.format(ior="diamond"))
    rc, info, err = _info(host, p)
    assert rc == 0 and info["dielectric_triangles"] == 2, err


def test_property_and_schema_errors(host, tmp_path):
    path = _write_scene(str(tmp_path), extra='<integer name="vrlTargetNum" value="5"/>')
    rc, info, err = _info(host, path)
    assert rc != 0 and "unreferenced property" in err and "vrlTargetNum" in err
    path = _write_scene(str(tmp_path), extra='<string name="target_vrls_x" value="$nope"/>')
    rc, info, err = _info(host, path)
    assert rc != 0 and "undefined parameter" in err
    path = _write_scene(str(tmp_path), extra='<boolean name="use_light_cut" value="true"/>'.replace("use_light_cut", "rr_vrl"))
    assert _info(host, path)[0] == 0
    old = os.path.join(str(tmp_path), "old.xml")
    open(old, "w").write('<scene version="0.6.0"><integrator type="vrl"/></scene>')
    rc, info, err = _info(host, old)
    assert rc != 0 and "no sensor" in err                                # 0.x files are upgraded now, not refused
    open(old, "w").write('<scene><integrator type="vrl"/></scene>')
    rc, info, err = _info(host, old)
    assert rc != 0 and "missing version attribute" in err
    bad = os.path.join(str(tmp_path), "bad.xml")
    open(bad, "w").write('<scene version="2.0.0"><sensor type="perspective"></scene>')
    rc, info, err = _info(host, bad)
    assert rc != 0 and "mismatched closing tag" in err and "line 1" in err


def _legacy(txt):
    """The synthetic scene respelt the way the 0.x scenes of data/FBD15ALVRL_scenes are written: camelCase names, <lookAt>,
    sigmaS / sigmaA, a medium boundary without BSDF, an integrator property of the LightSlice implementation."""
    txt = txt.replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
    txt = txt.replace('<scene version="2.0.0">', '<scene version="0.6.0">', 1)
    for a, b in (("fov_axis", "fovAxis"), ("near_clip", "nearClip"), ("far_clip", "farClip"), ("to_world", "toWorld"),
                 ("sample_count", "sampleCount"), ("<lookat", "<lookAt"), ("target_vrls", "targetVrls"),
                 ("samples_per_query", "samplesPerQuery"), ("use_light_cut", "useLightCut"), ("min_vrl_length", "minVrlLength")):
        txt = txt.replace(a, b)
    txt = txt.replace('<rgb name="sigma_t" value="0.2, 0.22, 0.24"/>', '<spectrum name="sigmaS" value="0.18, 0.198, 0.216"/>'
                      '<spectrum name="sigmaA" value="0.02, 0.022, 0.024"/>').replace('<rgb name="albedo" value="0.9"/>', "")
    txt = txt.replace('<shape type="cube">\n        <bsdf type="null"/>', '<shape type="cube" name="fogBox">')
    assert "sigmaS" in txt and "fogBox" in txt and "lookAt" in txt
    return txt


def test_legacy_schema_upgrade_equals_current_schema(host, tmp_path):
    """xml.cpp:350-392 (camelCase -> underscore_case, lookAt) plus the FBD15 translation (sigma_s/sigma_a, boundary without
    BSDF, foreign integrator properties): the 0.6-spelt scene loads to the same numbers as the 2.0-spelt one."""
    path = _write_scene(str(tmp_path), extra='<integer name="vrlTargetNum" value="300"/>')
    txt = open(path).read()
    new = os.path.join(str(tmp_path), "new.xml")
    open(new, "w").write(txt.replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
                         .replace('<integer name="vrlTargetNum" value="300"/>', ""))
    old = os.path.join(str(tmp_path), "old.xml")
    open(old, "w").write(_legacy(txt))
    rc_n, a, err_n = _info(host, new)
    rc_o, b, err_o = _info(host, old)
    assert rc_n == 0 and rc_o == 0, (err_n, err_o)
    for w in ("in-memory version upgrade (v0.6.0", "vrl_target_num", "sigma_t / albedo = 0.2 / 0.9, 0.22 / 0.9, 0.24 / 0.9", "index-matched boundary"):
        assert w in err_o, w
    assert "upgrade" not in err_n
    assert "target_vrls = 300" in err_o and "samples_per_query = 1" in err_o        # the renamed properties reached the plugin
    for k in ("triangles", "null_triangles", "emitter_triangles", "interior_triangles", "bbox", "emitters", "integrator"):
        assert a[k] == b[k], k
    assert a["sensor"] == b["sensor"]
    ma, mb = a["media"][0], b["media"][0]
    assert np.allclose(ma["sigma_t"], mb["sigma_t"], rtol=1e-7) and np.allclose(ma["albedo"], mb["albedo"], rtol=1e-7)
    assert {k: v for k, v in ma.items() if k not in ("sigma_t", "albedo")} == {k: v for k, v in mb.items() if k not in ("sigma_t", "albedo")}


FBD15 = "/root/reference/data/FBD15ALVRL_scenes"


@pytest.mark.skipif(not os.path.isdir(FBD15), reason="reference fixtures not mounted")
def test_fbd15_benchmark_scenes_load(host):
    """BASELINE config 4 names data/FBD15ALVRL_scenes: box.xml (Mitsuba 0.6 schema) loads after the upgrade + translation."""
    rc, info, err = _info(host, os.path.join(FBD15, "box", "box.xml"))
    assert rc == 0, err
    assert info["triangles"] == 7 * 2 + (2 * 32 - 2) * 64 + 12 and info["emitter_triangles"] == 2 and info["null_triangles"] == 2
    assert np.allclose(info["bbox"], [-10, -10, -10, 10, 10, 10], atol=1e-5) and info["emitters"] == [[20.0, 20.0, 20.0]]
    m = info["media"][0]
    assert m["id"] == "pm" and m["kind"] == abi.MEDIUM_HOMOGENEOUS and m["scale"] == 1.0
    assert np.allclose(m["sigma_t"], [0.11] * 3, rtol=1e-6) and np.allclose(m["albedo"], [0.1 / 0.11] * 3, rtol=1e-6)
    s = info["sensor"]
    assert (s["width"], s["height"], s["spp"], s["medium"]) == (256, 256, 1, -1)
    cam = scenes.make_camera(256, 256, (0, 0, -40), (0, 0, -39), fov=36, fov_axis="x", near_clip=0.1, far_clip=100)
    assert np.allclose(s["sample_to_camera"], list(cam.sample_to_camera), rtol=2e-6, atol=1e-7)
    assert np.allclose(s["to_world"], list(cam.to_world), rtol=2e-6, atol=1e-7)
    assert "max_passes" in err and info["mirror_triangles"] == (2 * 32 - 2) * 64       # the conductor sphere: specular chains
    rc, info, err = _info(host, os.path.join(FBD15, "glass", "glass.xml"))
    assert rc == 0, err
    assert [m["id"] for m in info["media"]] == ["grapefruitJuiceIsotropic", "absorbingGlass"]
    assert np.allclose(info["media"][0]["sigma_t"], [0.86, 1.27, 4.96], rtol=1e-6) and "envmap" in err
    assert info["dielectric_triangles"] == info["triangles"] > 10000                     # glass, liquid surface, interface: all dielectric


@pytest.mark.skipif(not os.path.isdir("/root/reference/data/hetvol"), reason="reference fixtures not mounted")
def test_hetvol_scene_loads(host):
    """BASELINE config "data/hetvol": a Mitsuba-0.4 scene (volpath, constvolume albedo, gridvolume placed by the file's
    bounding box, ldsampler, gaussian filter).  After the upgrade + translation the smoke grid sits exactly in bounds.obj."""
    rc, info, err = _info(host, "/root/reference/data/hetvol/hetvol.xml")
    assert rc == 0, err
    m = info["media"][0]
    assert (m["id"], m["kind"], m["phase"]) == ("smoke", abi.MEDIUM_GRID, abi.PHASE_HG) and m["scale"] == 100.0
    assert m["grid_res"] == [128, 128, 50] and np.allclose(m["albedo"], [0.9] * 3)
    assert np.allclose(m["grid_bbox"], [-0.5, -0.5, -0.1953125, 0.5, 0.5, 0.1953125], atol=1e-6)
    assert info["null_triangles"] == 12 and info["interior_triangles"] == 12 and info["emitters"] == [[400.0, 400.0, 400.0]]
    assert info["sensor"]["width"] == 768 and info["sensor"]["spp"] == 64 and info["integrator"] == "volpath"
    assert np.allclose(info["mean_reflectance"], [0.2, 0.2, 0.3])      # diffuseReflectance -> diffuse_reflectance -> reflectance
    for w in ("v0.4.0", "use_grid_bbox", "ldsampler", "gaussian"):
        assert w in err, w


@pytest.mark.skipif(not os.path.isfile("/root/reference/data/cbox/cbox.xml"), reason="reference fixtures not mounted")
def test_data_cbox_scene_loads(host):
    """BASELINE config "data/cbox": the Mitsuba-0.4 Cornell box (surface only; spectra as wavelength tables)."""
    rc, info, err = _info(host, "/root/reference/data/cbox/cbox.xml")
    assert rc == 0, err
    assert info["triangles"] == 38 and info["emitter_triangles"] == 2 and info["media"] == []
    assert np.allclose(info["emitters"], [[17.0, 8.0, 0.0]])          # 400:0, 500:8, 600:15.6, 700:18.4 by band means
    assert "wavelength table" in err and info["reflecting_triangles"] == 36


@pytest.mark.skipif(not os.path.isdir(REF_SCENES), reason="reference fixtures not mounted")
def test_reference_fixture_scenes_load(host):
    rc, info, err = _info(host, os.path.join(REF_SCENES, "cbox-volume-vrl.xml"), "use_light_cut=false")
    assert rc == 0, err
    assert info["integrator"] == "vrl" and info["null_triangles"] == 12 and info["emitter_triangles"] == 2
    assert np.allclose(info["bbox"], [-545, -275, -545, 545, 275, 545]) and info["sensor"]["width"] == 512
    m = info["media"][0]
    assert m["id"] == "medium1" and m["kind"] == abi.MEDIUM_HOMOGENEOUS and abs(m["scale"] - 0.011) < 1e-8
    assert "target_vrls = 10000" in err                                  # the fixture's integrator properties reached the plugin
    # the fixture asks for the light cut: it reaches the plugin with the fixture's thresholds (the tree is built in preprocess)
    rc, _, err = _info(host, os.path.join(REF_SCENES, "cbox-volume-vrl.xml"))
    assert rc == 0 and "use_light_cut = 1" in err, err
    rc, info, err = _info(host, os.path.join(REF_SCENES, "cbox-nonlinear-homo-vrl.xml"), "use_light_cut=false")
    assert rc == 0, err
    m = info["media"][0]
    assert m["kind"] == abi.MEDIUM_NONLINEAR and m["nl_res"] == [1, 640, 1] and np.allclose(m["nl_bbox"], [-400, -150, -400, 400, 150, 400])
    assert abs(m["top_ior"] - 0.95) < 1e-6 and m["bottom_ior"] == 1.0
    rc, info, err = _info(host, os.path.join(REF_SCENES, "cbox-heterogeneous-vrl.xml"), "use_light_cut=false")
    assert rc == 0, err
    m = info["media"][0]
    assert m["kind"] == abi.MEDIUM_GRID and m["grid_res"] == [16, 16, 16] and m["scale"] == 8.0 and m["grid_max"] > 0


@pytest.mark.gpu
def test_render_xml_end_to_end(host, tmp_path):
    """`vrl_host render-xml`: XML -> VRLIntegrator -> light tracing on the device -> gather -> XYZAW film, against the
    Python front-end rendering the same scene built from the loader's own numbers."""
    from mitsuba_nlvrl_b200 import api
    path = _write_scene(str(tmp_path))
    txt = open(path).read().replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
    open(path, "w").write(txt)
    out = os.path.join(str(tmp_path), "film.f32")
    r = subprocess.run([host, "render-xml", path, out, "-Dspp=1"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "VRL Map Created" in r.stderr
    film = np.fromfile(out, np.float32).reshape(48, 64, 5)
    assert np.isfinite(film).all() and (film[..., :3] >= 0).all() and film[..., 1].mean() > 0 and np.all(film[..., 4] == 1.0)
    # deterministic: light tracing seeds particle i with seed + i, the gather addresses its streams -> identical film
    out2 = os.path.join(str(tmp_path), "film2.f32")
    r2 = subprocess.run([host, "render-xml", path, out2, "-Dspp=1"], capture_output=True, text=True)
    assert r2.returncode == 0 and np.array_equal(np.fromfile(out2, np.float32), film.reshape(-1))
    # the developed film (XYZ / W -> linear sRGB) as a PFM
    pfm = os.path.join(str(tmp_path), "film.pfm")
    r3 = subprocess.run([host, "render-xml", path, pfm, "-Dspp=1", "target_vrls=200"], capture_output=True, text=True)
    assert r3.returncode == 0 and open(pfm, "rb").read(2) == b"PF"


@pytest.mark.gpu
def test_render_xml_legacy_schema_renders_the_same_film(host, tmp_path):
    """The 0.6-spelt scene renders the film of the 2.0-spelt one (the translation is exact up to one float rounding of the
    medium coefficients, so: same VRL count, film within 1e-5 relative)."""
    path = _write_scene(str(tmp_path))
    txt = open(path).read()
    new, old = os.path.join(str(tmp_path), "new.xml"), os.path.join(str(tmp_path), "old.xml")
    open(new, "w").write(txt.replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">'))
    open(old, "w").write(_legacy(txt))
    films = []
    for p in (new, old):
        out = p + ".f32"
        r = subprocess.run([host, "render-xml", p, out, "-Dspp=1", "use_direct_illum=false"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        films.append(np.fromfile(out, np.float32).reshape(48, 64, 5))
    a, b = films
    assert a[..., 1].mean() > 0
    assert np.allclose(a, b, rtol=1e-5, atol=1e-6 * float(a[..., :3].max()))


@pytest.mark.gpu
def test_render_xml_global_photon_map(host, tmp_path):
    """The plugin builds the global surface map from the same light tracing (global_photons, global_lookup_radius_relative,
    global_lookup_size) and the camera paths that end on a diffuse surface collect it; global_photons=0 switches it off."""
    path = _write_scene(str(tmp_path))
    txt = open(path).read().replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
    open(path, "w").write(txt)
    films = {}
    for n in ("20000", "0"):
        out = os.path.join(str(tmp_path), f"film_g{n}.f32")
        r = subprocess.run([host, "render-xml", path, out, "-Dspp=1", "use_direct_illum=false", f"global_photons={n}",
                            "global_lookup_size=40"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert ("Building global PM, size: 20000" in r.stderr) == (n == "20000"), r.stderr
        films[n] = np.fromfile(out, np.float32).reshape(48, 64, 5)
    a, b = films["20000"][..., 1], films["0"][..., 1]
    assert np.isfinite(a).all() and np.all(a >= b - 1e-7) and (a > b + 1e-9).sum() > 20 and a.sum() > b.sum()


@pytest.mark.gpu
def test_render_xml_with_delta_bsdfs(host, tmp_path):
    """A glass sphere and a mirror triangle in the XML: the scene renders through the specular chains (camera paths carry on
    behind the sphere) and the plugin traces the caustic map because a dielectric is present."""
    path = _write_scene(str(tmp_path))
    txt = open(path).read().replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
    open(path, "w").write(txt)
    glass = txt.replace('<bsdf type="diffuse"/>', '<bsdf type="dielectric"><string name="int_ior" value="water"/></bsdf>')
    assert glass != txt
    p2 = os.path.join(str(tmp_path), "glass.xml")
    open(p2, "w").write(glass)
    films = []
    for p in (path, p2):
        out = p + ".f32"
        r = subprocess.run([host, "render-xml", p, out, "-Dspp=1", "global_photons=20000", "caustic_photons=2000"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert ("caustic" in r.stderr) == (p == p2), r.stderr
        films.append(np.fromfile(out, np.float32).reshape(48, 64, 5))
    a, b = films
    assert np.isfinite(b).all() and (b[..., :3] >= 0).all() and b[..., 1].mean() > 0
    assert (np.abs(a[..., 1] - b[..., 1]) > 1e-6).sum() > 20             # the pixels that see the sphere changed


@pytest.mark.gpu
def test_render_xml_direct_term_switch(host, tmp_path):
    """use_direct_illum (default true, vrl.cpp:57): the plugin builds the volume photon map from the same light tracing and
    adds the direct in-scattering term; with use_direct_illum=false the film holds the VRL term only and is darker."""
    path = _write_scene(str(tmp_path))
    txt = open(path).read().replace('<medium id="fog" type="nonlinear">', '<medium id="fog" type="homogeneous">')
    open(path, "w").write(txt)
    films = {}
    for flag in ("true", "false"):
        out = os.path.join(str(tmp_path), f"film_{flag}.f32")
        r = subprocess.run([host, "render-xml", path, out, "-Dspp=1", f"use_direct_illum={flag}", "volume_photons=20000",
                            "volume_lookup_radius_relative=0.05"], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert ("Building volume PM" in r.stderr) == (flag == "true")
        films[flag] = np.fromfile(out, np.float32).reshape(48, 64, 5)
    a, b = films["true"][..., 1], films["false"][..., 1]
    assert np.isfinite(a).all() and np.all(a >= b - 1e-7) and a.sum() > 1.02 * b.sum()

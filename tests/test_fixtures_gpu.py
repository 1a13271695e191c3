"""The reference's OWN fixture scenes on the GPU ("Config 1-3 use the real fixture scenes"): the XML files,
meshes and volumes under tests/golden/scenes/ are copies of data/scenes/cbox/* and data/hetvol/* of the reference.  Each
is loaded by the C++ XML loader (`vrl_host dump-xml`), light-traced on the device into a VRL set of the size BASELINE.json
names, the set is exported, and the CUDA gather is compared with the CPU oracle on the same exported set and PCG32
streams, per camera segment against the oracle's fp64 sum at the north-star tolerance 1e-4.  A segment above the
tolerance has to be EXPLAINED: its evals are re-run one VRL at a time on both sides and the difference has to sit in a
handful of evals that took a different hit / accept decision (one side exactly 0, or a discrete jump), with everything
else agreeing."""
import os
import subprocess

import numpy as np
import pytest

import mitsuba_nlvrl_b200 as pkg
import oracle

scenes, api, abi = pkg.scenes, pkg.api, pkg.abi
pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FIX = os.path.join(ROOT, "tests", "golden", "scenes")

@pytest.fixture(scope="module")
def host():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "host")], stdout=subprocess.DEVNULL)
    return os.path.join(ROOT, "host", "vrl_host")


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def load_fixture(host, tmp_path, rel, *defs):
    out = str(tmp_path / (os.path.basename(rel) + ".dump"))
    r = subprocess.run([host, "dump-xml", os.path.join(FIX, rel), out, *defs], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return scenes.load_scene_dump(out)


from parity_tools import REL_TOL, check_pixels, check_segments, explain, rel_err, seg_key


def generate_set(ctx, n_target, emission_medium, **kw):
    gen = scenes.generate_desc(n_target, seed=7, start_medium=emission_medium, **kw)
    n, particles = ctx.generate_vrls(gen)
    soa, med = ctx.download_vrls()
    assert n >= n_target and not med.any()
    return np.ascontiguousarray(soa[:, :n_target]), ctx.vrls_scale()


def test_config1_cbox_volume_vrl_256_1k(ctx, host, tmp_path):
    """configs[0]: data/scenes/cbox/cbox-volume-vrl.xml (Cornell box + homogeneous medium), 256 x 256, 1k VRLs traced from the
    scene's area light on the device; the WHOLE frame against the oracle."""
    sc, cam, spp, seed, em = load_fixture(host, tmp_path, "cbox/cbox-volume-vrl.xml", "-Dres=256")
    assert len(sc.tri_flags) == 4004 and cam.width == cam.height == 256
    ctx.set_scene(sc)
    soa, scale = generate_set(ctx, 1000, em)
    ctx.upload_vrls(scenes.vrlset_desc(soa, scale=scale))
    opt = scenes.default_options(visibility=1, samples_per_query=1, base_seed=seed)
    n_segs = ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    assert n_segs >= 0.9 * 256 * 256                           # nearly every camera ray crosses the fog cube
    li = ctx.gather_segments(opt).astype(np.float64)
    osc = oracle.OracleScene(sc)
    # the segments themselves: bit for bit the oracle's (camera, null-BSDF crossings, stream positions)
    osegs, _, _ = osc.build_segments(cam, opt, 1000)
    assert len(osegs) == n_segs
    a, b = segs[seg_key(segs)], osegs[seg_key(osegs)]          # (the kernel appends with an atomic cursor: compare as sets)
    for f in ("o", "d", "maxt", "pixel", "rng_state"):
        assert np.array_equal(a[f], b[f]), f
    wd = np.abs(a["weight"] - b["weight"]).max(axis=1) / np.maximum(np.abs(b["weight"]).max(axis=1), 1e-30)
    if wd.max() > 5e-6:
        k = int(np.argmax(wd))
        print("weight mismatch", wd.max(), (wd > 2e-6).sum(), a[k], b[k])
    assert wd.max() <= 5e-6                                      # throughput behind the fog: CUDA expf vs glibc
    err = check_segments(ctx, osc, segs, np.arange(n_segs), li, soa, scale, opt, "cbox-volume-vrl.xml 256^2 x 1k")
    assert np.quantile(err, 0.999) <= REL_TOL
    # and the film through vrlb_render from host buffers
    film = ctx.render(cam, opt, scenes.vrlset_desc(soa, scale=scale))
    assert np.isfinite(film).all() and film[..., 4].min() == 1.0 and film[..., :3].max() > 0


def test_config2_hetvol_smoke_512_10k(ctx, host, tmp_path):
    """configs[1]: data/hetvol/hetvol.xml -- smoke.vol (128 x 128 x 50 grid) in bounds.obj -- at 512 x 512 with 10k VRLs and the
    settings of cbox-heterogeneous-vrl.xml (uniform sampling); bands of rows through the plume against the oracle."""
    sc, cam0, spp, seed, em = load_fixture(host, tmp_path, "hetvol/hetvol.xml")
    m = sc.media[0]
    assert m.kind == abi.MEDIUM_GRID and list(m.grid_res) == [128, 128, 50]
    # the file's sensor at 512 x 512 (BASELINE's size): same pose and field of view
    import ctypes as C
    cam = abi.Camera.from_buffer_copy(bytes(cam0))
    s2c = np.array(cam0.sample_to_camera, np.float64).reshape(4, 4)
    cam.width = cam.height = 512
    # sample_to_camera encodes the aspect ratio: rebuild it for a square film from the x field of view of the file
    x_extent = abs((s2c @ np.array([1, 0.5, 0, 1]))[0] / (s2c @ np.array([1, 0.5, 0, 1]))[3] - (s2c @ np.array([0, 0.5, 0, 1]))[0] / (s2c @ np.array([0, 0.5, 0, 1]))[3])
    fov_x = 2 * np.degrees(np.arctan(0.5 * x_extent / cam0.near_clip))
    sq = scenes.make_camera(512, 512, (0, 0, 0), (0, 0, 1), fov=fov_x, near_clip=cam0.near_clip, far_clip=cam0.far_clip)
    cam.sample_to_camera[:] = list(sq.sample_to_camera)
    ctx.set_scene(sc)
    soa, scale = generate_set(ctx, 10000, em)
    ctx.upload_vrls(scenes.vrlset_desc(soa, scale=scale))
    opt = scenes.default_options(visibility=1, samples_per_query=1, use_uniform_sampling=1, base_seed=seed)
    n_segs = ctx.build_segments(cam, opt)
    assert n_segs > 512 * 64
    segs = ctx.download_segments()
    li = ctx.gather_segments(opt).astype(np.float64)
    y = segs["pixel"] // 512
    rows = np.nonzero((y == 200) | (y == 330))[0][::4]          # every 4th pixel of two rows: ~250 segments x 10k VRLs
    osc = oracle.OracleScene(sc)
    check_segments(ctx, osc, segs, rows, li, soa, scale, opt, "hetvol.xml 512^2 x 10k", max_explain=2)


def test_cbox_heterogeneous_vrl_fixture(ctx, host, tmp_path):
    """data/scenes/cbox/cbox-heterogeneous-vrl.xml as shipped (sigmat.vol 16^3, scale 8, uniform sampling, 100 VRLs) at 128^2."""
    sc, cam, spp, seed, em = load_fixture(host, tmp_path, "cbox/cbox-heterogeneous-vrl.xml", "-Dres=128")
    assert sc.media[0].kind == abi.MEDIUM_GRID and list(sc.media[0].grid_res) == [16, 16, 16]
    ctx.set_scene(sc)
    soa, scale = generate_set(ctx, 100, em)
    ctx.upload_vrls(scenes.vrlset_desc(soa, scale=scale))
    opt = scenes.default_options(visibility=1, samples_per_query=1, use_uniform_sampling=1, base_seed=seed)
    n_segs = ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    li = ctx.gather_segments(opt).astype(np.float64)
    osc = oracle.OracleScene(sc)
    check_segments(ctx, osc, segs, np.arange(0, n_segs, 3), li, soa, scale, opt, "cbox-heterogeneous-vrl.xml 128^2 x 100")


@pytest.mark.parametrize("name,size", [("cbox/cbox-nonlinear-homo-vrl.xml", (1280, 720)), ("cbox/mirage-vrl.xml", (1024, 512))])
def test_config3_graded_ior_fixtures(ctx, host, tmp_path, name, size):
    """configs[2]: the graded-IOR fixtures -- cbox-nonlinear-homo-vrl.xml at 1280 x 720 and mirage-vrl.xml (from_bottom = false)
    at its own 1024 x 512 -- with curved camera paths; 1k VRLs traced through the graded medium on the device."""
    sc, cam0, spp, seed, em = load_fixture(host, tmp_path, name)
    assert sc.media[0].kind == abi.MEDIUM_NONLINEAR
    cam = cam0
    if (cam0.width, cam0.height) != size:                        # 512 x 256 in the file: same aspect ratio 2... 720p is 16:9, rebuild
        s2c = np.array(cam0.sample_to_camera, np.float64).reshape(4, 4)
        p1, p0 = s2c @ np.array([1, 0.5, 0, 1]), s2c @ np.array([0, 0.5, 0, 1])
        x_extent = abs(p1[0] / p1[3] - p0[0] / p0[3])
        fov_x = 2 * np.degrees(np.arctan(0.5 * x_extent / cam0.near_clip))
        cam = scenes.make_camera(size[0], size[1], (0, 0, 0), (0, 0, 1), fov=fov_x, near_clip=cam0.near_clip, far_clip=cam0.far_clip)
        cam.to_world[:] = list(cam0.to_world)
        cam.medium = cam0.medium
    ctx.set_scene(sc)
    if name.endswith("mirage-vrl.xml"):                          # its light is an environment map (not on the path): a laser stands in
        laser, start = ((0.0, 200.0, -400.0), (0.0, -0.35, 1.0), (50.0, 50.0, 50.0)), 0
    else:                                                        # the fixture's own laser (use_laser, laser_origin, laser_direction), from outside the fog
        laser, start = ((-810.0, 154.9, -149.9), (0.9, -0.1, 0.0), (7.0, 30.0, 2.0)), -1
    d = np.array(laser[1]); laser = (laser[0], tuple(d / np.linalg.norm(d)), laser[2])
    gen = scenes.generate_desc(1000, seed=7, start_medium=start, laser=laser, min_vrl_length=50.0, max_steps=1000000)
    n, particles = ctx.generate_vrls(gen)
    soa, med = ctx.download_vrls()
    soa = np.ascontiguousarray(soa[:, :1000]); scale = ctx.vrls_scale()
    ctx.upload_vrls(scenes.vrlset_desc(soa, scale=scale))
    opt = scenes.default_options(visibility=1, samples_per_query=1, use_non_linear_camera=1, base_seed=seed)
    n_segs = ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    n_paths = len(np.unique(segs["pixel"]))
    assert n_segs > 1.5 * n_paths                               # curved paths: several parts per pixel
    li = ctx.gather_segments(opt).astype(np.float64)
    osc = oracle.OracleScene(sc)
    y, x = segs["pixel"] // cam.width, segs["pixel"] % cam.width
    row = cam.height // 2
    in_row = np.nonzero(y == row)[0]
    band = in_row[x[in_row] % 5 == 0]                           # every 5th pixel of the row, all the parts of its path
    # the curved parts of one film row, bit for bit the oracle's (the whole film is minutes of single-threaded marching)
    osegs, _, _ = osc.build_segments(cam, opt, 1000, cap=len(in_row) + 1024, crop=(0, row, cam.width, row + 1))
    a, b = segs[in_row][seg_key(segs[in_row])], osegs[seg_key(osegs)]
    assert len(osegs) == len(in_row) and np.array_equal(a["o"], b["o"]) and np.array_equal(a["d"], b["d"]) and np.array_equal(a["maxt"], b["maxt"])
    assert np.array_equal(a["rng_state"], b["rng_state"])
    check_pixels(ctx, osc, segs, band, li, soa, scale, opt, f"{name} {size[0]}x{size[1]} x 1k")

"""GPU suite: the closest-hit BVH built on the device (bvh_lbvh.cu, vrlb_set_bvh_builder(1)) against the host build.

The closest hit does not depend on the tree (ties go to the lowest original triangle index, scene.cpp:114-131), so everything
that walks it must come out IDENTICAL: the camera segments (origin, direction, length, throughput, stream position) and the
light tracer's VRL set -- on the box-fog scene and on a fog box filled with 20 000 random triangles."""
import numpy as np
import pytest

from mitsuba_nlvrl_b200 import abi, api, scenes
from parity_tools import seg_key

pytestmark = pytest.mark.gpu


def _segments(builder, sc, cam, opt, gen=None):
    ctx = api.Context(0)
    ctx.set_bvh_builder(builder)
    ctx.set_scene(sc)
    info = ctx.bvh_info()
    ctx.upload_vrls(scenes.vrlset_desc(scenes.synthetic_vrls(16, seed=1)))
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    segs = segs[seg_key(segs)]                   # seg_key is the sorting permutation (segments are appended through an atomic cursor)
    soa = None
    if gen is not None:
        ctx.generate_vrls(gen)
        soa, _ = ctx.download_vrls()
    ctx.close()
    return segs, soa, info


def _same(a, b):
    assert len(a) == len(b)
    for f in a.dtype.names:
        assert np.array_equal(a[f], b[f]), f


def test_box_fog_scene_is_walked_identically():
    sc = scenes.box_fog_scene(occluders=True, wall_reflectance=(0.7, 0.6, 0.5))
    cam, opt = scenes.box_fog_camera(320, 180), scenes.default_options(visibility=1)
    gen = scenes.generate_desc(3000, seed=2, laser=((0, 9.5, 0), (0.1, -1, 0.05), (50, 50, 50)), start_medium=0)
    host_segs, host_vrls, hi = _segments(0, sc, cam, opt, gen)
    dev_segs, dev_vrls, di = _segments(1, sc, cam, opt, gen)
    n_tris = len(sc.tri_flags)
    assert hi["builder"] == 0 and di["builder"] == 1 and di["nodes"] == 2 * n_tris - 1 and 8 <= di["depth"] <= 44
    print(f"box-fog, {n_tris} triangles: host SAH depth {hi['depth']}, {hi['nodes']} nodes, {hi['build_us']} us; device LBVH depth {di['depth']}, {di['nodes']} nodes, {di['build_us']} us")
    _same(host_segs, dev_segs)
    assert np.array_equal(host_vrls, dev_vrls)


def test_random_triangle_soup_is_walked_identically():
    rng = np.random.default_rng(11)
    n = 20000
    c = rng.uniform(-9, 9, (n, 1, 3))
    tri = (c + rng.normal(0, 0.25, (n, 3, 3))).astype(np.float32)
    sc = scenes.Scene()
    med = sc.add_medium(scenes.make_medium(sigma_t=(0.11,) * 3, albedo=(0.9,) * 3, scale=1.0))
    sc.add_tris(scenes.box_tris((-10.01,) * 3, (10.01,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(med))
    sc.add_tris(tri, abi.SURF_BSDF_SMOOTH)
    cam, opt = scenes.box_fog_camera(256, 144), scenes.default_options(visibility=1)
    host_segs, _, hi = _segments(0, sc, cam, opt)
    dev_segs, _, di = _segments(1, sc, cam, opt)
    print(f"soup, {n + 12} triangles: host SAH depth {hi['depth']}, {hi['build_us']} us; device LBVH depth {di['depth']}, {di['build_us']} us (builder used: {di['builder']})")
    assert di["builder"] == 1
    _same(host_segs, dev_segs)
    assert len(host_segs) > 30000 and host_segs["maxt"].std() > 1.0          # the paths do end on the soup


@pytest.mark.parametrize("math_mode", [0, 1])
def test_shadow_trees_from_the_device_give_the_same_film(math_mode):
    """With builder 1 the trees of the shadow structure (main geometry, shells, solids) come from the device as well.  A shadow
    query is a yes / no: the exactly rounded kernel must produce a bit-identical film; the fast kernel may meet a grazing
    triangle before or after a certain hit depending on the tree, and then re-evaluates that one eval exactly (1e-5 of it)."""
    from parity_tools import rel_err
    sc, cam = scenes.box_fog_scene(occluders=True), scenes.box_fog_camera(320, 180)
    opt = scenes.default_options(visibility=1, math_mode=math_mode)
    vd = scenes.vrlset_desc(scenes.synthetic_vrls(2048, seed=6))
    films, infos = [], []
    for builder in (0, 1):
        ctx = api.Context(0)
        ctx.set_bvh_builder(builder)
        ctx.set_scene(sc)
        infos.append(ctx.bvh_info())
        ctx.upload_vrls(vd)
        ctx.build_segments(cam, opt)
        ctx.clear_film(); ctx.gather(opt)
        films.append(ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64))
        ctx.close()
    print(f"shadow structure: host depth {infos[0]['shadow_depth']}; device: {infos[1]['shadow_trees_on_device']} trees, depth {infos[1]['shadow_depth']}")
    assert infos[0]["shadow_trees_on_device"] == 0 and infos[1]["shadow_trees_on_device"] >= 3
    differing = (films[0] != films[1]).any(axis=1)
    if math_mode == 0:
        assert not differing.any(), differing.sum()
    else:
        # measured: 1 of 57 600 pixels differs, by 1.03e-4 (one eval of 2048 decided by the exact re-evaluation instead of the
        # fast test, DESIGN.md section 5 kernel 4 "graze"); every other pixel is bit-identical
        err = rel_err(films[1], films[0])
        assert err.max() <= 2e-3 and (err > 1e-4).sum() <= 3 and differing.mean() <= 0.002, (err.max(), differing.sum())

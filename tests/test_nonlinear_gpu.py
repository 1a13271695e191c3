"""The CUDA graded-IOR kernels against the reference's OWN non-linear medium (tests/golden/ref_nonlinear.json, made by
oracle/ref_probe_nl.cpp from src/media/nonlinear.cpp compiled in place): `segments_kernel` (camera paths) and
`generate_kernel` (light tracer) march with the fixed-step eikonal integrator the north star prescribes; where the
reference completes its cell walk their exit points / directions lie within a few cells of it, for from_bottom true and
false.  The classes and bounds are those of tests/test_ref_nonlinear_cpu.py."""
import numpy as np
import pytest

import mitsuba_nlvrl_b200 as pkg
import oracle
from test_ref_nonlinear_cpu import GOLD, classify, grazing, medium_scene

scenes, api, abi = pkg.scenes, pkg.api, pkg.abi
pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gold():
    import json
    with open(GOLD) as f:
        return json.load(f)


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def camera_path(ctx, opt, o, d):
    """The parts segments_kernel makes of the camera ray through o along d: a 1x1 film whose only sample looks along d.  The
    sensor sits OUTSIDE the container, 50 units back along the ray (the reference never queries a medium the sensor is
    inside of, vrl.cpp:523), so the in-medium path starts on the container's face, <= 0.4 units before the golden ray's origin."""
    o, d = np.asarray(o, np.float64), np.asarray(d, np.float64)
    up = (0, 1, 0) if abs(d[1]) < 0.9 else (1, 0, 0)
    cam = scenes.make_camera(1, 1, origin=o - 50.0 * d, target=o, up=up, fov=1.0, near_clip=1e-3, far_clip=1e5, medium=-1)
    n = ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    assert n == len(segs) >= 1
    return segs


@pytest.mark.parametrize("name", ["cbox-nonlinear-homo.xml", "mirage.xml", "cbox-nonlinear-homo.xml res_y 160"])
def test_camera_paths_follow_the_reference_walk(ctx, gold, name):
    m = [x for x in gold["media"] if x["name"] == name][0]
    cell = (m["hi"][1] - m["lo"][1]) / m["res_y"]
    sc, _ = medium_scene(m)
    ctx.set_scene(sc)
    osc = oracle.OracleScene(sc)
    opt = scenes.default_options(use_non_linear_camera=1, visibility=0)
    regular, turning, stalled = classify(m["rays"])
    regular = regular & ~grazing(m["rays"])
    n_checked, worst = 0, 0.0
    for r, reg in zip(m["rays"], regular):
        segs = camera_path(ctx, opt, r["o"], r["d"])
        end = segs["o"][-1].astype(np.float64) + segs["d"][-1].astype(np.float64) * float(segs["maxt"][-1])
        length = float(segs["maxt"].astype(np.float64).sum())
        assert np.linalg.norm(segs["o"][0] + segs["d"][0] * (0.37 / r["d"][2]) - np.array(r["o"])) <= 0.05
        # the oracle's marcher from where the path enters the medium
        ok, p, dd, olen, steps = osc.nl_march(segs["o"][0], r["d"], mint=1e-3)      # (the first part's d is a chord, not the entry direction)
        assert ok and np.linalg.norm(end - p) <= 0.05 * cell + 5e-3
        assert abs(length - olen) <= 1e-4 * olen + 5e-3                    # chords of the marched arc; the oracle sums thousands of fp32 steps
        # the chain is connected and every part stays within 0.25 degrees of the path's direction where it starts
        nxt = segs["o"][:-1].astype(np.float64) + segs["d"][:-1].astype(np.float64) * segs["maxt"][:-1, None]
        assert np.allclose(nxt, segs["o"][1:], atol=2e-3)
        if not reg or np.sign(segs["d"][-1][1]) != np.sign(r["d"][1]):     # (a path that turns in the continuous profile: the
            continue                                                        # turning class of test_ref_nonlinear_cpu.py)
        # the reference's walk (complete, no reflection, not a grazing start): exit point, path length, exit direction
        e = np.linalg.norm(end - np.array(r["exit_p"]))
        assert e <= 6.0 * cell, (name, r["o"], e / cell)
        assert abs(length - 0.37 / r["d"][2] - r["length"]) <= 6.0 * cell
        assert np.linalg.norm(segs["d"][-1] - np.array(r["exit_d"])) <= 4.4e-3 + 4e-3 * cell       # last chord: within the 0.25 degree part tolerance
        n_checked += 1
        worst = max(worst, e / cell)
    assert n_checked >= 5
    print(f"{name}: {n_checked} complete walks, camera-path exit within {worst:.2f} cells of the reference")
    if stalled.any():                                  # mirage.xml: curved where the reference goes straight (it stalls)
        turned = 0
        for r, s in zip(m["rays"], stalled):
            if s:
                segs = camera_path(ctx, opt, r["o"], r["d"])
                turned += int(len(segs) > 1)
        assert turned >= 0.5 * stalled.sum()


@pytest.mark.parametrize("name", ["cbox-nonlinear-homo.xml", "mirage.xml"])
def test_light_tracer_flights_follow_the_reference_walk(ctx, gold, name):
    """generate_kernel: a laser particle through the (practically non-scattering) medium leaves a chain of VRLs; its end
    point and last direction against the reference's walk of the same ray."""
    m = [x for x in gold["media"] if x["name"] == name][0]
    cell = (m["hi"][1] - m["lo"][1]) / m["res_y"]
    sc = scenes.Scene()
    med = scenes.make_nonlinear_medium(m["lo"], m["hi"], m["res_y"], bottom_ior=m["bottom_ior"], top_ior=m["top_ior"],
                                       sigma_t=(1.0,) * 3, albedo=(0.5,) * 3, scale=1e-9)
    med.from_bottom = m["from_bottom"]
    mid = sc.add_medium(med)
    sc.add_tris(scenes.box_tris(m["lo"], m["hi"]), abi.SURF_BSDF_NULL | abi.surf_interior(mid))
    ctx.set_scene(sc)
    regular, turning, stalled = classify(m["rays"])
    regular = regular & ~grazing(m["rays"])
    n_checked = 0
    for r, reg in zip(m["rays"], regular):
        if not reg or np.sign(r["exit_d"][1]) != np.sign(r["d"][1]):
            continue
        gen = scenes.generate_desc(1500, seed=1, min_vrl_length=4.0, laser=(r["o"], r["d"], (1, 1, 1)), start_medium=0, max_steps=1000000)
        n, particles = ctx.generate_vrls(gen)
        soa, _ = ctx.download_vrls()
        o, dd, ln = soa[0:3].T.astype(np.float64), soa[3:6].T.astype(np.float64), soa[6].astype(np.float64)
        k = 1
        while k < n and np.allclose(o[k], o[k - 1] + dd[k - 1] * ln[k - 1], atol=2e-3):
            k += 1
        end = o[k - 1] + dd[k - 1] * ln[k - 1]
        assert np.linalg.norm(end - np.array(r["exit_p"])) <= 6.0 * cell, (name, r["o"], np.linalg.norm(end - np.array(r["exit_p"])) / cell)
        assert np.linalg.norm(dd[k - 1] - np.array(r["exit_d"])) <= 4e-3 * cell + 1e-3
        assert abs(ln[:k].sum() - r["length"]) <= 6.0 * cell
        n_checked += 1
    assert n_checked >= 5

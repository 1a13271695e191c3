"""Parity of the CUDA gather (through the C ABI) against the CPU oracle on the same exported VRL set and
the same PCG32 streams.  Tolerance from BASELINE.json north_star: per-pixel relative error <= 1e-4
(pixels below 1e-6 * max use that as an absolute floor)."""
import numpy as np
import pytest

import oracle
from mitsuba_nlvrl_b200 import abi, api, scenes

pytestmark = pytest.mark.gpu

REL_TOL = 1e-4


def film_err_fp64(ctx, osc, cam, opt, vd, n_vrls):
    """Per-pixel relative error of the device's RGB radiance film (after ctx.render) against the fp64-summed oracle:
    sum over the pixel's camera segments of weight * Li (fp64 sum over the VRLs) + the emitter radiance seen directly."""
    rgb = ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64)
    segs = ctx.download_segments()
    o32, o64, ev = osc.gather(vd, segs, opt)
    ref = np.zeros_like(rgb)
    np.add.at(ref, segs["pixel"], o64 * segs["weight"])
    _, direct, _ = osc.build_segments(cam, opt, n_vrls, cap=len(segs) + 1024)
    ref += direct.reshape(-1, 3)
    return rel_err(rgb, ref)


def rel_err(got, ref):
    floor = 1e-6 * np.abs(ref).max()
    return np.abs(got - ref) / np.maximum(np.abs(ref), floor)


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def _setup(occluders, phase=abi.PHASE_ISOTROPIC, n_vrls=2000, w=48, h=32, **optkw):
    sc = scenes.box_fog_scene(occluders=occluders, phase=phase)
    cam = scenes.box_fog_camera(w, h)
    opt = scenes.default_options(**optkw)
    soa = scenes.synthetic_vrls(n_vrls, seed=3)
    return sc, cam, opt, soa


@pytest.mark.parametrize("occluders,vis", [(False, 0), (True, 1)])
@pytest.mark.parametrize("uniform", [0, 1])
@pytest.mark.parametrize("math_mode", [0, 1])
def test_segments_and_gather_match_oracle(ctx, occluders, vis, uniform, math_mode):
    sc, cam, opt, soa = _setup(occluders, visibility=vis, use_uniform_sampling=uniform, math_mode=math_mode)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ref_segs, ref_direct, ref_aw = osc.build_segments(cam, opt, soa.shape[1])
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    n = ctx.build_segments(cam, opt)
    assert n == len(ref_segs)
    segs = ctx.download_segments()
    key = lambda s: np.lexsort((s["part"], s["sample"], s["pixel"]))
    a, b = segs[key(segs)], ref_segs[key(ref_segs)]
    for f in ("pixel", "sample", "part", "medium", "rng_state"):
        assert np.array_equal(a[f], b[f]), f
    for f in ("o", "d", "maxt", "weight"):
        assert np.array_equal(a[f], b[f]), f            # same IEEE op order: bit-exact
    # gather on the GPU's own segments, oracle on the same records
    li = ctx.gather_segments(opt)
    o32, o64, ev = osc.gather(vd, segs, opt)
    assert ev == len(segs) * soa.shape[1]
    err = rel_err(li.astype(np.float64), o64)
    assert err.max() <= REL_TOL, f"max per-segment rel err {err.max():.3e}"
    # film: weight * Li accumulated per pixel + alpha/weight channels
    rgb = ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64)
    ref = np.zeros_like(rgb)
    np.add.at(ref, segs["pixel"], o64 * segs["weight"])
    ref += ref_direct.reshape(-1, 3)
    assert rel_err(rgb, ref).max() <= REL_TOL
    xyzaw = ctx.film_xyzaw(cam.width, cam.height)
    assert np.array_equal(xyzaw[..., 3:], ref_aw)


@pytest.mark.parametrize("math_mode", [0, 1])
def test_hg_rr_q2_match_oracle(ctx, math_mode):
    sc, cam, opt, soa = _setup(True, phase=abi.PHASE_HG, n_vrls=1500, visibility=1, rr_vrl=1, samples_per_query=2, math_mode=math_mode)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    li = ctx.gather_segments(opt)
    o32, o64, ev = osc.gather(vd, segs, opt)
    err = rel_err(li.astype(np.float64), o64)
    assert err.max() <= REL_TOL, f"{err.max():.3e}"


@pytest.mark.parametrize("rr,q", [(0, 1), (1, 2)])
def test_nl_atomic_query_matches_oracle(ctx, rr, q):
    """use_nl_atomic_query (vrl.cpp:613-617): ONE uniform-sampled query over the whole curved camera path.  The parts
    of a sample are consecutive records that share the query's stream position (sample = j | n_parts << 16); every
    eval lands on exactly one part (NLRay::at), the distance roulette is measured from the first part."""
    sc = scenes.nl_fog_scene(occluders=True, res_y=160, top_ior=0.8)
    cam = scenes.box_fog_camera(32, 18)
    opt = scenes.default_options(visibility=1, use_nl_atomic_query=1, rr_vrl=rr, samples_per_query=q)
    soa = scenes.synthetic_vrls(600, seed=12)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ref_segs, _, _ = osc.build_segments(cam, opt, soa.shape[1], cap=32 * 18 * 200)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    n = ctx.build_segments(cam, opt)
    assert n == len(ref_segs) and n > 3 * 32 * 18
    segs = ctx.download_segments()
    key = lambda s: np.lexsort((s["part"], s["sample"], s["pixel"]))
    a, b = segs[key(segs)], ref_segs[key(ref_segs)]
    for f in ("pixel", "sample", "part", "medium", "rng_state", "o", "d", "maxt", "weight"):
        assert np.array_equal(a[f], b[f]), f
    assert (segs["sample"] >> 16).min() >= 1
    first = segs["part"] == 0
    groups = np.cumsum(first) - 1                                # records of a sample are consecutive, part 0 first
    assert np.array_equal(segs["rng_state"], segs["rng_state"][first][groups])
    assert np.array_equal(np.bincount(groups), (segs["sample"] >> 16)[first])
    li = ctx.gather_segments(opt).astype(np.float64)
    o32, o64, ev = osc.gather(vd, segs, opt)
    assert o64.max() > 0
    got, ref = np.zeros((32 * 18, 3)), np.zeros((32 * 18, 3))
    np.add.at(got, segs["pixel"], li * segs["weight"])
    np.add.at(ref, segs["pixel"], o64 * segs["weight"])
    assert rel_err(got, ref).max() <= REL_TOL
    assert np.quantile(rel_err(li, o64), 0.99) <= REL_TOL       # an eval on a part boundary may land on the neighbour
    out = ctx.render(cam, opt, vd)
    ref_xyzaw, _, _ = osc.render(cam, opt, vd)
    assert np.array_equal(out[..., 3:], ref_xyzaw[..., 3:])
    assert film_err_fp64(ctx, osc, cam, opt, vd, soa.shape[1]).max() <= REL_TOL      # the film against the fp64-summed oracle


@pytest.mark.parametrize("occluders,vis,uniform,rr,math_mode", [(False, 0, 0, 0, 0), (True, 1, 0, 0, 1), (True, 1, 1, 1, 0)])
def test_grid_medium_matches_oracle(ctx, occluders, vis, uniform, rr, math_mode):
    """Grid (heterogeneous) medium, SURVEY 8a a8-a10: trilinear sigma_s at both points and the reference's stochastic
    majorant transmittance on the per-eval sub-stream.  Every math mode runs the exactly-rounded kernel here."""
    sc = scenes.grid_fog_scene(occluders=occluders, res=(32, 25, 32))
    cam = scenes.box_fog_camera(32, 18)
    opt = scenes.default_options(visibility=vis, use_uniform_sampling=uniform, rr_vrl=rr, math_mode=math_mode)
    soa = scenes.synthetic_vrls(600, seed=5)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    ref_segs, _, _ = osc.build_segments(cam, opt, soa.shape[1])
    key = lambda s: np.lexsort((s["part"], s["sample"], s["pixel"]))
    for f in ("o", "d", "maxt", "weight", "rng_state"):
        assert np.array_equal(segs[key(segs)][f], ref_segs[key(ref_segs)][f]), f
    li = ctx.gather_segments(opt).astype(np.float64)
    o32, o64, ev = osc.gather(vd, segs, opt)
    assert o64.max() > 0
    # a free-flight accept / reject decided differently by one ulp of logf changes that eval by a discrete factor: every
    # segment above the tolerance has to be explained by such evals, everything else agreeing (parity_tools.explain)
    import parity_tools
    parity_tools.check_segments(ctx, osc, segs, np.arange(len(segs)), li, soa, 1.0 / soa.shape[1], opt, "grid medium 32x18", max_explain=4, max_bad=len(segs) // 25)
    ctx.upload_vrls(vd); ctx.upload_segments(segs, cam.width, cam.height)


def test_vrl_shards_sum_to_whole(ctx):
    """VRL sharding: two shards with global stream addressing add up to the single-set result."""
    sc, cam, opt, soa = _setup(True, n_vrls=2048, visibility=1)
    ctx.set_scene(sc)
    n = soa.shape[1]
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    whole = ctx.gather_segments(opt).astype(np.float64)
    parts = np.zeros_like(whole)
    for lo, hi in ((0, 777), (777, n)):
        shard = np.ascontiguousarray(soa[:, lo:hi])
        ctx.upload_vrls(scenes.vrlset_desc(shard, scale=1.0 / n, first_index=lo, total=n))
        parts += ctx.gather_segments(opt)
    assert rel_err(parts, whole).max() <= 1e-5


def test_render_end_to_end_matches_oracle(ctx):
    sc, cam, opt, soa = _setup(True, n_vrls=1000, w=64, h=36, visibility=1)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ref_xyzaw, ref_rgb, ev = osc.render(cam, opt, vd)
    ctx.set_scene(sc)
    out = ctx.render(cam, opt, vd)
    assert np.array_equal(out[..., 3:], ref_xyzaw[..., 3:])
    t = ctx.timing()
    assert film_err_fp64(ctx, osc, cam, opt, vd, soa.shape[1]).max() <= REL_TOL      # the film against the fp64-summed oracle
    assert t["n_evals"] == ev and t["gather_launches"] == 1


@pytest.mark.parametrize("math_mode", [0, 1])
def test_nonlinear_camera_parts_match_oracle(ctx, math_mode):
    """Graded-IOR medium (config 3): curved camera paths are marched with the eikonal integrator and split into
    straight parts, one query per part with the preceding parts' transmittance in the weight (vrl.cpp:550-629)."""
    sc = scenes.nl_fog_scene(occluders=True, res_y=160, top_ior=0.8)
    cam = scenes.box_fog_camera(32, 18)
    opt = scenes.default_options(visibility=1, math_mode=math_mode)
    soa = scenes.synthetic_vrls(600, seed=12)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ref_segs, ref_direct, ref_aw = osc.build_segments(cam, opt, soa.shape[1], cap=32 * 18 * 200)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    n = ctx.build_segments(cam, opt)
    assert n == len(ref_segs) and n > 3 * 32 * 18, "camera paths must split into several parts"
    segs = ctx.download_segments()
    key = lambda s: np.lexsort((s["part"], s["sample"], s["pixel"]))
    a, b = segs[key(segs)], ref_segs[key(ref_segs)]
    for f in ("pixel", "sample", "part", "medium", "rng_state", "o", "d", "maxt"):
        assert np.array_equal(a[f], b[f]), f                 # same IEEE operation order: bit-exact parts
    assert np.allclose(a["weight"], b["weight"], rtol=2e-6, atol=0)   # products of exp(): CUDA expf vs glibc differ by ulps
    li = ctx.gather_segments(opt).astype(np.float64)
    o32, o64, ev = osc.gather(vd, segs, opt)
    # the north-star tolerance is per PIXEL: sum over the parts of weight * Li.  (A part can be arbitrarily short --
    # the piece between the last cut and the surface -- and then theta_B - theta_A of the equi-angular pdf cancels, so
    # single parts are only held to a loose bound; they carry a negligible share of their pixel.)
    got, ref = np.zeros((32 * 18, 3)), np.zeros((32 * 18, 3))
    np.add.at(got, segs["pixel"], li * segs["weight"])
    np.add.at(ref, segs["pixel"], o64 * segs["weight"])
    assert rel_err(got, ref).max() <= REL_TOL
    assert rel_err(li, o64).max() <= 2e-2 and np.quantile(rel_err(li, o64), 0.99) <= REL_TOL
    # straight rays when the option is off (vrl.cpp:550 m_useNonLinearCameraRays)
    opt.use_non_linear_camera = 0
    assert ctx.build_segments(cam, opt) == 32 * 18
    # whole frame
    opt.use_non_linear_camera = 1
    out = ctx.render(cam, opt, vd)
    ref_xyzaw, _, _ = osc.render(cam, opt, vd)
    assert np.array_equal(out[..., 3:], ref_xyzaw[..., 3:])
    assert film_err_fp64(ctx, osc, cam, opt, vd, soa.shape[1]).max() <= REL_TOL      # the film against the fp64-summed oracle


@pytest.mark.parametrize("math_mode", [0, 1])
def test_gather_matches_reference_compiled_golden(ctx, math_mode):
    """CUDA gather against outputs of the REFERENCE'S OWN getContrib (tests/golden/ref_vrl_struct.json, made by
    oracle/_ref/refprobe = src/integrators/vrl/vrl_struct.h compiled in place): 96 VRLs, q = 1/2, importance and
    uniform sampling, with and without occluders.  Only the golden queries whose sequential stream consumed the full
    2 draws per eval are used (no early-out), where addressed == sequential stream positions."""
    import ctypes as C
    import json
    import os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vrl_struct.json")) as f:
        gold = json.load(f)
    soa = np.ascontiguousarray(np.array(gold["query_vrls"], np.float32).T)
    vd = scenes.vrlset_desc(soa, scale=gold["query_scale"])
    m = gold["medium"]
    used = 0
    for c in gold["query"]:
        if c["draws"] != 2 * c["samples"] * soa.shape[1]:
            continue
        used += 1
        sc = scenes.Scene()
        sc.add_medium(scenes.make_medium(sigma_t=m["sigma_t"], albedo=m["albedo"], scale=m["scale"]))
        if c["vis"]:
            sc.add_tris(np.array(gold["triangles"], np.float32).reshape(-1, 3, 3), abi.SURF_BSDF_SMOOTH)
        opt = scenes.default_options(samples_per_query=c["samples"], use_uniform_sampling=c["uniform"],
                                     visibility=c["vis"], math_mode=math_mode)
        st, inc = C.c_uint64(), C.c_uint64()
        oracle.lib().vrlo_pcg32_seed(c["seed"], 0xda3e39cb94b95bdb, C.byref(st), C.byref(inc))
        seg = np.zeros(1, abi.SEGMENT_DTYPE)
        r = np.array(c["ray"], np.float32)
        seg["o"][0], seg["d"][0], seg["maxt"][0] = r[0:3], r[3:6], r[6]
        seg["weight"][0] = 1.0
        seg["rng_state"][0] = st.value
        ctx.set_scene(sc)
        ctx.upload_vrls(vd)
        ctx.upload_segments(seg, 1, 1)
        li = ctx.gather_segments(opt)
        want = np.array(c["out"], np.float64)
        err = np.abs(li[0] - want) / np.abs(want)
        assert err.max() <= REL_TOL, (c["seed"], math_mode, li[0], want)
    assert used >= 6


@pytest.mark.parametrize("phase", [abi.PHASE_ISOTROPIC, abi.PHASE_HG])
def test_photon_direct_term_matches_oracle(ctx, phase):
    """Direct in-scattering term from the volume photon map (vrl.cpp:523-611, photonmap.h:216-271): the grid-binned CUDA
    gather against the oracle's loop over all photons, per segment; then the film: with a resident photon set
    vrlb_gather adds the term on top of the VRL gather, without one it does not."""
    sc = scenes.box_fog_scene(occluders=True, phase=phase, g=0.6)
    cam = scenes.box_fog_camera(48, 32)
    opt = scenes.default_options(visibility=1)
    soa = scenes.synthetic_vrls(300, seed=3)
    vd = scenes.vrlset_desc(soa)
    pos, pdir, power = scenes.synthetic_photons(20000, seed=9)
    pd = scenes.photonset_desc(pos, pdir, power, scale=1.0 / 19999, lookup_radius=0.7)
    osc = oracle.OracleScene(sc)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    ctx.upload_photons(pd)
    assert ctx.photons_count() == 20000
    got = ctx.photon_gather_segments(opt).astype(np.float64)
    ref = osc.photon_gather(pd, segs, opt).astype(np.float64)
    assert (ref > 0).mean() > 0.5                                       # most camera segments see photons
    assert rel_err(got, ref).max() <= 2e-5                              # same fp32 operations, different photon visiting order
    # film = VRL term + direct term (the direct term already carries the path throughput)
    ctx.clear_film(); ctx.gather(opt)
    with_direct = ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64)
    ctx.clear_photons()
    assert ctx.photons_count() == 0
    ctx.clear_film(); ctx.gather(opt)
    vrl_only = ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64)
    want = vrl_only.copy()
    np.add.at(want, segs["pixel"], ref)
    assert rel_err(with_direct, want).max() <= REL_TOL
    assert with_direct.sum() > 1.05 * vrl_only.sum()


@pytest.mark.parametrize("atomic", [0, 1])
def test_photon_direct_term_on_curved_paths_matches_oracle(ctx, atomic):
    """Direct in-scattering term along curved camera paths (vrl.cpp:553-608): the parts of a path are consecutive records
    carrying the part count, the thread of the first record walks them with the step phase carried across parts; against
    the oracle's walk over the same records, per path, then on the film."""
    sc = scenes.nl_fog_scene(occluders=True, res_y=160, top_ior=0.8)
    cam = scenes.box_fog_camera(32, 18)
    opt = scenes.default_options(visibility=1, use_nl_atomic_query=atomic)
    soa = scenes.synthetic_vrls(300, seed=12)
    vd = scenes.vrlset_desc(soa)
    pos, pdir, power = scenes.synthetic_photons(20000, seed=9)
    pd = scenes.photonset_desc(pos, pdir, power, scale=1.0 / 19999, lookup_radius=0.7)
    osc = oracle.OracleScene(sc)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    idx, n_parts = segs["part"] & 0xffff, np.maximum(segs["part"] >> 16, segs["sample"] >> 16)
    first = np.flatnonzero(idx == 0)
    assert n_parts.min() >= 1 and n_parts[first].max() >= 3 and n_parts[first].sum() == len(segs)
    for i in first[:: max(1, len(first) // 64)]:                      # the group of a first record follows it, in order
        k = int(n_parts[i])
        assert np.array_equal(idx[i:i + k], np.arange(k)) and np.all(segs["pixel"][i:i + k] == segs["pixel"][i])
    ctx.upload_photons(pd)
    got = ctx.photon_gather_segments(opt).astype(np.float64)
    ref = osc.photon_gather(pd, segs, opt).astype(np.float64)
    assert np.all(got[idx != 0] == 0) and (ref[first] > 0).mean() > 0.5
    assert rel_err(got, ref).max() <= 2e-5
    ctx.clear_film(); ctx.gather(opt)
    with_direct = ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64)
    ctx.clear_photons()
    ctx.clear_film(); ctx.gather(opt)
    vrl_only = ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64)
    want = vrl_only.copy()
    np.add.at(want, segs["pixel"], ref)
    assert rel_err(with_direct, want).max() <= REL_TOL
    assert with_direct.sum() > 1.05 * vrl_only.sum()


@pytest.mark.parametrize("radius,k", [(0.5, 50), (0.2, 120), (0.6, 1), (0.35, 1024)])
def test_surface_photon_estimates_match_oracle(ctx, radius, k):
    """Global + caustic surface estimates (vrl.cpp:671-680, photonmap.h:131-211) through vrlb_surface_gather_queries: grid
    binned k-nearest with the shrinking radius against the oracle's sort over all photons; ~157 / 25 / 226 / 77 photons in
    range for the four cases (so: shrunk, not shrunk, k = 1, k never reached); then the film."""
    w, h = 48, 32
    pos, pdir, power = scenes.synthetic_surface_photons(20000, seed=21)
    cpos, cdir, cpower = scenes.synthetic_surface_photons(5000, seed=23)
    g = scenes.photonset_desc(pos, pdir, power, scale=1.0 / 812, lookup_radius=radius)
    c = scenes.photonset_desc(cpos, cdir, cpower, scale=1.0 / 377, lookup_radius=0.5 * radius)
    q = scenes.synthetic_surface_queries(3000, w, h, seed=22, twosided_every=5)
    # the film below needs a resident scene / VRL set / segments
    sc = scenes.box_fog_scene(occluders=False)
    cam = scenes.box_fog_camera(w, h)
    opt = scenes.default_options(visibility=0)
    ctx.set_scene(sc)
    ctx.upload_vrls(scenes.vrlset_desc(scenes.synthetic_vrls(100, seed=3)))
    ctx.build_segments(cam, opt)
    ctx.upload_surface_queries(q, w, h)
    ctx.upload_surface_photons(abi.PHOTONMAP_GLOBAL, g, k)
    ctx.clear_surface_photons(abi.PHOTONMAP_CAUSTIC)
    assert ctx.surface_photons_count(abi.PHOTONMAP_GLOBAL) == 20000 and ctx.surface_photons_count(abi.PHOTONMAP_CAUSTIC) == 0
    got = ctx.surface_gather_queries().astype(np.float64)
    ref = oracle.surface_gather(q, global_desc=g, k_global=k).astype(np.float64)
    front = (q["wi"][:, 2] > 0) | (q["flags"] != 0)
    assert np.all(ref[~front] == 0) and np.all(got[~front] == 0)          # the back of a one-sided surface
    assert (ref[front] > 0).mean() > 0.6 and (ref[q["wi"][:, 2] < 0] > 0).sum() > 20   # twosided: seen from below too
    assert rel_err(got, ref).max() <= 2e-5
    kc = max(2, k // 2)                                                 # (k = 1 puts the cone filter at its zero)
    ctx.upload_surface_photons(abi.PHOTONMAP_CAUSTIC, c, kc)
    got2 = ctx.surface_gather_queries().astype(np.float64)
    ref2 = oracle.surface_gather(q, global_desc=g, k_global=k, caustic_desc=c, k_caustic=kc).astype(np.float64)
    assert (ref2.sum() > 1.05 * ref.sum()) and rel_err(got2, ref2).max() <= 2e-5
    # film = VRL term + weight * estimate at the query's pixel; without maps it is the VRL term alone
    ctx.clear_film(); ctx.gather(opt)
    with_surface = ctx.film_rgb(w, h).reshape(-1, 3).astype(np.float64)
    ctx.clear_surface_photons(abi.PHOTONMAP_GLOBAL); ctx.clear_surface_photons(abi.PHOTONMAP_CAUSTIC)
    ctx.clear_film(); ctx.gather(opt)
    vrl_only = ctx.film_rgb(w, h).reshape(-1, 3).astype(np.float64)
    want = vrl_only.copy()
    np.add.at(want, q["pixel"], ref2 * q["weight"])
    assert rel_err(with_surface, want).max() <= REL_TOL
    ctx.upload_surface_queries(q[:0], w, h)                              # leave no queries behind for the other tests


def _same_queries(got, ref):
    """GPU surface queries against the oracle's: one per pixel at spp 1, same IEEE operations -> position, normal,
    direction bit-exact; the throughput is a product of exp() (CUDA expf vs glibc: ulps)."""
    assert len(got) == len(ref)
    a, b = got[np.argsort(got["pixel"])], ref[np.argsort(ref["pixel"])]
    for f in ("pixel", "flags", "p", "n", "wi", "reflectance"):
        assert np.array_equal(a[f], b[f]), f
    assert np.allclose(a["weight"], b["weight"], rtol=2e-6, atol=0)


def test_camera_paths_hand_their_first_smooth_surface_to_the_photon_maps(ctx):
    """vrl.cpp:671-680: where a camera path ends on a smooth surface, radiance += (caustic + global estimate) * throughput.
    vrlb_segments_build writes one surface query per such path (position, normal, -ray.d, reflectance, throughput) when the
    scene carries reflectances; vrlb_gather adds weight * estimate to the film."""
    w, h = 64, 36
    refl = (0.7, 0.6, 0.5)
    sc = scenes.box_fog_scene(occluders=True, wall_reflectance=refl)     # sphere / cuboid carry no reflectance: no query there
    cam = scenes.box_fog_camera(w, h)
    opt = scenes.default_options(visibility=1)
    ctx.set_scene(sc)
    ctx.upload_vrls(scenes.vrlset_desc(scenes.synthetic_vrls(200, seed=3)))
    ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    q = ctx.download_surface_queries()
    assert len(q) > 0.5 * w * h and len(np.unique(q["pixel"])) == len(q)  # one path per pixel at spp 1
    # the paths that end on a wall are those whose in-medium segment ends there
    end = segs["o"] + segs["d"] * segs["maxt"][:, None]
    on_wall = np.abs(end).max(axis=1) > scenes.BOX - 1e-3
    assert set(segs["pixel"][on_wall]) == set(q["pixel"])
    _same_queries(q, oracle.OracleScene(sc).build_surface_queries(cam, opt, 200))
    p = q["p"].astype(np.float64)
    assert np.allclose(np.abs(p).max(axis=1), scenes.BOX, atol=1e-3)
    assert np.allclose(q["reflectance"], np.array(refl, np.float32)[None, :]) and np.all(q["flags"] == 0)
    assert ((q["wi"] * q["n"]).sum(1) > 0).all()                           # inward walls seen from inside
    # throughput = transmittance over the stretch inside the fog box [-10.01, 10.01]^3 (camera at z = -40 outside it)
    o = np.array([0.0, 0.0, -40.0])
    d = p - o
    dist = np.linalg.norm(d, axis=1)
    t_in = (-(scenes.BOX + 0.01) - o[2]) / (d[:, 2] / dist)
    assert np.allclose(q["weight"][:, 0], np.exp(-0.11 * (dist - t_in)), rtol=2e-4)
    assert np.allclose(q["wi"], -d / dist[:, None], atol=2e-5)
    # film: a global map on the back wall (photons travelling +z arrive on its front)
    pos, pdir, power = scenes.synthetic_surface_photons(20000, seed=31, half=10.0, z=10.0)
    g = scenes.photonset_desc(pos, pdir, power, scale=1.0 / 5000, lookup_radius=0.8)
    ref = oracle.surface_gather(q, global_desc=g, k_global=60).astype(np.float64)
    assert (ref.sum(1) > 0).sum() > 0.1 * len(q)
    ctx.upload_surface_photons(abi.PHOTONMAP_GLOBAL, g, 60)
    assert rel_err(ctx.surface_gather_queries().astype(np.float64), ref).max() <= 2e-5
    ctx.clear_film(); ctx.gather(opt)
    with_surface = ctx.film_rgb(w, h).reshape(-1, 3).astype(np.float64)
    ctx.clear_surface_photons(abi.PHOTONMAP_GLOBAL)
    ctx.clear_film(); ctx.gather(opt)
    vrl_only = ctx.film_rgb(w, h).reshape(-1, 3).astype(np.float64)
    want = vrl_only.copy()
    np.add.at(want, q["pixel"], ref * q["weight"])
    assert rel_err(with_surface, want).max() <= REL_TOL and with_surface.sum() > 1.05 * vrl_only.sum()
    # a scene without reflectances collects no queries
    ctx.set_scene(scenes.box_fog_scene(occluders=True))
    ctx.build_segments(cam, opt)
    assert len(ctx.download_surface_queries()) == 0


def test_specular_chains_on_the_camera_path_match_oracle(ctx):
    """vrl.cpp:682-716: a camera path that meets a delta BSDF (mirror sphere, glass cuboid) is reflected / refracted with
    bsdf->sample, multiplies its throughput and carries on through the fog: more in-medium segments per pixel, each with
    its own query.  Segments bit-exact against the oracle's restatement, then the gather per pixel."""
    w, h = 64, 36
    sc = scenes.specular_fog_scene()
    cam = scenes.box_fog_camera(w, h)
    opt = scenes.default_options(visibility=1)
    soa = scenes.synthetic_vrls(400, seed=3)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ref_segs, ref_direct, ref_aw = osc.build_segments(cam, opt, soa.shape[1], cap=w * h * 40)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    n = ctx.build_segments(cam, opt)
    assert n == len(ref_segs) and n > 1.3 * w * h                         # the chains add segments
    segs = ctx.download_segments()
    end = lambda s: np.round((s["o"] + s["d"] * s["maxt"][:, None]).astype(np.float64).sum(1) * 1e3)
    key = lambda s: np.lexsort((s["rng_state"], s["pixel"]))              # a pixel's segments in path order: stream position
    a, b = segs[key(segs)], ref_segs[key(ref_segs)]
    for f in ("pixel", "sample", "part", "medium", "rng_state", "o", "d", "maxt"):
        assert np.array_equal(a[f], b[f]), f
    assert np.allclose(a["weight"], b["weight"], rtol=2e-6, atol=0)
    # tinted mirror: some segments carry an RGB-distinct throughput; the glass scales by eta^-2 / eta^2
    assert (np.abs(a["weight"][:, 0] - a["weight"][:, 2]) > 1e-3).sum() > 100
    per_pixel = np.bincount(segs["pixel"], minlength=w * h)
    assert per_pixel.max() >= 3
    li = ctx.gather_segments(opt).astype(np.float64)
    o32, o64, ev = osc.gather(vd, segs, opt)
    got, ref = np.zeros((w * h, 3)), np.zeros((w * h, 3))
    np.add.at(got, segs["pixel"], li * segs["weight"])
    np.add.at(ref, segs["pixel"], o64 * segs["weight"])
    assert rel_err(got, ref).max() <= REL_TOL
    # the paths still end on diffuse walls: surface queries carry the chain's throughput; bit-exact against the oracle's
    q = ctx.download_surface_queries()
    assert len(q) > 0.8 * w * h and (np.abs(q["weight"][:, 0] - q["weight"][:, 2]) > 1e-3).sum() > 50
    _same_queries(q, osc.build_surface_queries(cam, opt, soa.shape[1]))


def test_surface_estimates_match_reference_compiled_photonmap(ctx):
    """The CUDA surface estimates through the C ABI against outputs of the REFERENCE'S OWN PhotonMap (photonmap.h + kdtree.h
    compiled in place, tests/golden/ref_photonmap.json): global and caustic estimates of 120 queries, three (radius, k)."""
    import json
    import os
    gold = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_photonmap.json")))
    ph = np.array(gold["surface_photons"], np.float32)
    cs = ph[::gold["caustic_every"]]
    cases = {}
    for c in gold["surface"]:
        cases.setdefault((c["radius"], c["k"]), []).append(c)
    assert len(cases) == 3
    for (radius, k), cs_ in cases.items():
        q = np.zeros(len(cs_), abi.SURFACE_QUERY_DTYPE)
        for i, c in enumerate(cs_):
            q["p"][i], q["n"][i], q["wi"][i], q["reflectance"][i] = c["p"], c["n"], c["wi"], c["reflectance"]
        q["weight"] = 1.0
        ctx.upload_surface_queries(q, 4, 4)
        want_g = np.array([c["global"] for c in cs_], np.float64)
        want_c = np.array([c["caustic"] for c in cs_], np.float64)
        g = scenes.photonset_desc(ph[:, 0], ph[:, 1], ph[:, 2], scale=gold["global_scale"], lookup_radius=radius)
        cd = scenes.photonset_desc(cs[:, 0], cs[:, 1], cs[:, 2], scale=gold["caustic_scale"], lookup_radius=float(np.float32(0.8) * np.float32(radius)))
        ctx.upload_surface_photons(abi.PHOTONMAP_GLOBAL, g, k)
        ctx.clear_surface_photons(abi.PHOTONMAP_CAUSTIC)
        assert rel_err(ctx.surface_gather_queries().astype(np.float64), want_g).max() <= 2e-5
        ctx.clear_surface_photons(abi.PHOTONMAP_GLOBAL)
        ctx.upload_surface_photons(abi.PHOTONMAP_CAUSTIC, cd, k)
        assert rel_err(ctx.surface_gather_queries().astype(np.float64), want_c).max() <= 2e-5
    ctx.clear_surface_photons(abi.PHOTONMAP_CAUSTIC)
    ctx.upload_surface_queries(q[:0], 4, 4)


@pytest.mark.parametrize("math_mode", [0, 1])
def test_spp4_matches_oracle(ctx, math_mode):
    """sampler sample_count 4: every sample of a pixel draws its anti-aliasing jitter first (2 draws, only when spp > 1,
    integrator.cpp:249), then wavelength and channel; the four samples of a pixel continue one PCG32 stream, and the film
    divides by W = 4 (imageblock.cpp:162-169)."""
    sc, cam, opt, soa = _setup(True, n_vrls=600, w=40, h=24, visibility=1, math_mode=math_mode, spp=4)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    ref_segs, ref_direct, ref_aw = osc.build_segments(cam, opt, soa.shape[1])
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    n = ctx.build_segments(cam, opt)
    assert n == len(ref_segs) == 40 * 24 * 4
    segs = ctx.download_segments()
    key = lambda s: np.lexsort((s["part"], s["sample"], s["pixel"]))
    a, b = segs[key(segs)], ref_segs[key(ref_segs)]
    for f in ("pixel", "sample", "part", "medium", "rng_state", "o", "d", "maxt", "weight"):
        assert np.array_equal(a[f], b[f]), f                    # jittered camera rays and stream positions: bit-exact
    assert set(np.unique(segs["sample"])) == {0, 1, 2, 3}
    # the four samples of a pixel look through different points of it
    first = segs[segs["pixel"] == segs["pixel"][0]]
    assert len(first) == 4 and len({tuple(d) for d in first["d"]}) == 4
    li = ctx.gather_segments(opt).astype(np.float64)
    o32, o64, ev = osc.gather(vd, segs, opt)
    import parity_tools
    parity_tools.check_segments(ctx, osc, segs, np.arange(n), li, soa, 1.0 / soa.shape[1], opt, f"spp 4, math_mode {math_mode}", max_explain=3)
    ctx.upload_vrls(vd)
    out = ctx.render(cam, opt, vd)
    ref_xyzaw, _, _ = osc.render(cam, opt, vd)
    assert np.array_equal(out[..., 3:], ref_xyzaw[..., 3:]) and out[..., 4].min() == 4.0
    assert film_err_fp64(ctx, osc, cam, opt, vd, soa.shape[1]).max() <= 5e-3          # (the per-segment check above carries the tolerance)


@pytest.mark.parametrize("scale", [0.05, 50.0])
def test_fast_path_parity_is_scale_invariant(ctx, scale):
    """The conditioning guard of math_mode 1 is 1 % of the scene's extent, not an absolute length: the same frame with every
    length multiplied by 0.05 / 50 (extinction divided accordingly) keeps the per-segment parity, and the share of evals the
    guard sends to the exactly-rounded path stays what it is on the +-10 box (timing: no 5x slow-down at the small scale)."""
    sc0, cam0, opt, soa0 = _setup(True, n_vrls=1500, w=48, h=32, visibility=1, math_mode=1)
    sc = scenes.Scene()
    m = scenes.make_medium(sigma_t=(0.11 / scale,) * 3, albedo=(0.1 / 0.11,) * 3)
    sc.add_medium(m)
    flags = sc0.tri_flags
    start = 0
    tv = sc0.tri_verts * np.float32(scale)
    while start < len(flags):
        end = start + 1
        while end < len(flags) and flags[end] == flags[start]:
            end += 1
        sc.add_tris(tv[start:end], int(flags[start]))
        start = end
    cam = scenes.make_camera(48, 32, origin=(0, 0, -40 * scale), target=(0, 0, -39 * scale), fov=36.0, fov_axis="x",
                             near_clip=0.1 * scale, far_clip=1000.0 * scale)
    soa = soa0.copy()
    soa[0:3] *= np.float32(scale); soa[6] *= np.float32(scale)
    soa = np.ascontiguousarray(soa)
    vd = scenes.vrlset_desc(soa)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    n = ctx.build_segments(cam, opt)
    assert n == 48 * 32
    segs = ctx.download_segments()
    li = ctx.gather_segments(opt).astype(np.float64)
    osc = oracle.OracleScene(sc)
    import parity_tools
    err = parity_tools.check_segments(ctx, osc, segs, np.arange(n), li, soa, 1.0 / soa.shape[1], opt, f"box-fog x {scale}", max_explain=3)
    assert np.quantile(err, 0.99) <= 1e-5

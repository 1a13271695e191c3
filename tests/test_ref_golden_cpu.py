"""CPU suite (-m "not gpu"): pins the oracle's VRL maths against outputs of the REFERENCE'S OWN code.

tests/golden/ref_vrl_struct.json was produced by oracle/_ref/refprobe = the reference's
src/integrators/vrl/vrl_struct.h compiled where it lies (against the mock Mitsuba of oracle/ref_mock.h and the
reference's vendored Enoki; generator oracle/ref_probe.cpp + oracle/make_golden.py).  It holds, with their inputs:
  closest        VRL::findClosetPoint      vrl_struct.h:163-228
  sampling       VRL::samplingVRL          vrl_struct.h:239-360   (+ how many sampler draws each call consumed)
  transmittance  VRL::evalTransmittance    vrl_struct.h:509-572
  contrib        VRL::getContrib           vrl_struct.h:584-662
  query          getContrib summed over a VRL list on one sequential stream (vrl_map.h:151-169)

tests/golden/ref_photonmap.json comes from oracle/_ref/refprobe_pm = the reference's photonmapper/photonmap.h + kdtree.h
compiled the same way (oracle/ref_probe_pm.cpp): estimateRadiance / estimateCausticRadiance / estimateRadianceVolume.

Tolerances: findClosetPoint is pure +,*,/ in the same order -> bit-exact.  Everything that goes through
log/sinh/atan/tan/exp is compared at 2e-5 relative: the reference evaluates Enoki's polynomial approximations,
the oracle glibc's correctly-rounded-ish libm, and the inverse CDFs amplify an ulp of difference (the north-star
tolerance on the image is 1e-4)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import oracle
from mitsuba_nlvrl_b200 import abi, scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vrl_struct.json")
RTOL = 2e-5


def _num(x):
    return float(x) if not isinstance(x, str) else float(x.replace("nan", "nan"))


def _arr(xs):
    return np.array([_num(x) for x in xs], np.float32)


@pytest.fixture(scope="module")
def gold():
    with open(GOLD) as f:
        return json.load(f)


def _medium(gold, hg=False):
    m = gold["medium"]
    return scenes.make_medium(sigma_t=m["sigma_t"], albedo=m["albedo"], scale=m["scale"],
                              phase=abi.PHASE_HG if hg else abi.PHASE_ISOTROPIC, g=m["g"])


def _scene(gold, occluders, hg=False):
    sc = scenes.Scene()
    sc.add_medium(_medium(gold, hg))
    if occluders:
        sc.add_tris(np.array(gold["triangles"], np.float32).reshape(-1, 3, 3), abi.SURF_BSDF_SMOOTH)
    return sc


def _close(got, want, rtol=RTOL, atol=0.0):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    return np.all(np.abs(got - want) <= rtol * np.abs(want) + atol)


def test_find_closest_point_bit_exact(gold):
    L = oracle.lib()
    for c in gold["closest"]:
        v, r = oracle.f32(c["vrl"]), oracle.f32(c["ray"])
        tc, tv = C.c_float(), C.c_float()
        L.vrlo_find_closest_point(oracle.fp(v[0:3]), oracle.fp(v[3:6]), v[6], oracle.fp(r[0:3]), oracle.fp(r[3:6]), r[6],
                                  C.byref(tc), C.byref(tv))
        want = np.array(c["out"], np.float32)
        assert np.float32(tc.value) == want[0] and np.float32(tv.value) == want[1], c


def test_sampling_vrl_matches_reference(gold):
    L = oracle.lib()
    n_invalid = n_nan = 0
    for c in gold["sampling"]:
        v, r = oracle.f32(c["vrl"]), oracle.f32(c["ray"])
        out = np.zeros(9, np.float32)
        valid = L.vrlo_sampling_vrl(oracle.fp(v[0:3]), oracle.fp(v[3:6]), v[6], oracle.fp(r[0:3]), oracle.fp(r[3:6]), r[6],
                                    c["uniform"], c["u"][0], c["u"][1], oracle.fp(out))
        want = _arr(c["out"])
        # the early-outs (h == 0, parallel) return before any draw: vrl_struct.h:264-265, 275-278
        assert (c["draws"] == 0) == (valid == 0), c
        if not valid:
            n_invalid += 1
            assert np.all(want == 0)
            continue
        if np.isnan(want[8]):
            # asinh by log(z + sqrt(z*z+1)) cancelling -> NaN pdf; getContrib rejects it (vrl_struct.h:594)
            n_nan += 1
            assert np.isnan(out[8]), c
            continue
        scale = max(1.0, float(np.abs(want[:8]).max()))
        assert _close(out[:8], want[:8], atol=RTOL * scale), (c, out)
        assert _close(out[8], want[8], rtol=5e-5), (c, out)
    assert n_invalid >= 10      # the fixture does exercise the early-outs


def test_eval_transmittance_matches_reference(gold):
    sc = oracle.OracleScene(_scene(gold, True))
    blocked = 0
    for c in gold["transmittance"]:
        got = sc.transmittance(c["p1"], c["p2"], 0)
        want = _arr(c["out"])
        blocked += int(want[0] == 0)
        assert (got[0] == 0) == (want[0] == 0), c
        assert _close(got, want), (c, got)
    assert blocked >= 10


@pytest.mark.parametrize("hg", [0, 1])
def test_get_contrib_matches_reference(gold, hg):
    L = oracle.lib()
    med = _medium(gold, bool(hg))
    sc_occ = _scene(gold, True, bool(hg))
    desc = sc_occ.desc()
    nz = 0
    for c in gold["contrib"]:
        if c["hg"] != hg:
            continue
        v, r = oracle.f32(c["vrl"]), oracle.f32(c["ray"])
        out = np.zeros(3, np.float32)
        L.vrlo_get_contrib(C.byref(desc) if c["vis"] else None, C.byref(med), oracle.fp(v), oracle.fp(r[0:3]),
                           oracle.fp(r[3:6]), r[6], c["uniform"], c["u"][0], c["u"][1], oracle.fp(out))
        want = _arr(c["out"])
        if np.any(np.isnan(want)):
            assert np.any(np.isnan(out)) or np.all(out == 0), (c, out)
            continue
        nz += int(np.any(want != 0))
        assert np.all((out == 0) == (want == 0)), (c, out)
        assert _close(out, want, rtol=5e-6), (c, out)   # measured max 3.5e-7
    assert nz >= 50


def test_query_sum_matches_reference_sequential_stream(gold):
    """VRLMap::query over 96 VRLs, q = 1 or 2, on the reference's one sequential PCG32 stream per pixel: the
    oracle in VRLO_RNG_SEQUENTIAL mode must consume the same draws (early-outs consume none) and sum the same Li."""
    L = oracle.lib()
    soa = np.ascontiguousarray(np.array(gold["query_vrls"], np.float32).T)
    vd = scenes.vrlset_desc(soa, scale=gold["query_scale"])
    for c in gold["query"]:
        sc = oracle.OracleScene(_scene(gold, bool(c["vis"])))
        opt = scenes.default_options(samples_per_query=c["samples"], use_uniform_sampling=c["uniform"],
                                     visibility=c["vis"], math_mode=0)
        st, inc = C.c_uint64(), C.c_uint64()
        L.vrlo_pcg32_seed(c["seed"], 0xda3e39cb94b95bdb, C.byref(st), C.byref(inc))
        seg = np.zeros(1, abi.SEGMENT_DTYPE)
        r = oracle.f32(c["ray"])
        seg["o"][0], seg["d"][0], seg["maxt"][0] = r[0:3], r[3:6], r[6]
        seg["weight"][0] = 1.0
        seg["rng_state"][0] = st.value
        o32, o64, ev = sc.gather(vd, seg, opt, rng_mode=oracle.RNG_SEQUENTIAL, n_threads=1)
        want = _arr(c["out"])
        assert _close(o32[0], want, rtol=5e-6), (c, o32[0])


def test_distance_roulette_query_matches_reference_sequential_stream(gold):
    """VRLMap::query with EDistanceRoulette (vrl_map.h:171-190) on the reference's sequential stream: the roulette draw
    comes before the eval's draws, rrWeight = min(1 / (distance * scale_rr), 1), survivors are divided by it."""
    soa = np.ascontiguousarray(np.array(gold["query_vrls"], np.float32).T)
    vd = scenes.vrlset_desc(soa, scale=gold["query_scale"])
    L = oracle.lib()
    n_pruned = 0
    for c in gold["query_rr"]:
        sc = oracle.OracleScene(_scene(gold, bool(c["vis"])))
        opt = scenes.default_options(samples_per_query=c["samples"], use_uniform_sampling=c["uniform"], visibility=c["vis"],
                                     math_mode=0, rr_vrl=1, scale_rr=c["scale_rr"])
        st, inc = C.c_uint64(), C.c_uint64()
        L.vrlo_pcg32_seed(c["seed"], 0xda3e39cb94b95bdb, C.byref(st), C.byref(inc))
        seg = np.zeros(1, abi.SEGMENT_DTYPE)
        r = oracle.f32(c["ray"])
        seg["o"][0], seg["d"][0], seg["maxt"][0] = r[0:3], r[3:6], r[6]
        seg["weight"][0] = 1.0
        seg["rng_state"][0] = st.value
        o32, o64, ev = sc.gather(vd, seg, opt, rng_mode=oracle.RNG_SEQUENTIAL, n_threads=1)
        assert _close(o32[0], _arr(c["out"]), rtol=5e-6), (c, o32[0])
        n_pruned += c["pruned"]
    assert n_pruned > 500                                  # the roulette does prune


def test_dice_matches_reference(gold):
    """VRL::dice compiled from the reference (pieces of mean length / 3, flux attenuated by the medium transmittance):
    piece count, origins and lengths bit-exact (fp64 position arithmetic rounded once), flux to 1e-6 (Enoki exp vs libm)."""
    src = np.ascontiguousarray(np.array(gold["dice_in"], np.float32).T)
    want = np.array(gold["dice_out"], np.float64)
    sc = oracle.OracleScene(_scene(gold, False))
    got, med = sc.vrls_dice(src, gold["dice_n"])
    assert got.shape[1] == len(want)
    w = want[:, 1:].T.astype(np.float32)
    assert np.array_equal(got[0:7], w[0:7])
    assert _close(got[7:10], w[7:10], rtol=1e-6)
    # the reference emits a zero-length tail piece whenever the length is an exact multiple of the dice length, and one
    # zero-length piece for a zero-length VRL (while (current_distance <= length))
    counts = np.bincount(want[:, 0].astype(int), minlength=src.shape[1])
    assert counts[3] == 1 and counts.min() >= 1


def test_to_long_matches_reference(gold):
    """VRL::convertToLong compiled from the reference against the brute-force mock scene: si.t of the closest surface."""
    src = np.ascontiguousarray(np.array(gold["dice_in"], np.float32).T)
    sc = scenes.Scene()
    sc.add_medium(_medium(gold))
    sc.add_tris(np.array(gold["room_triangles"], np.float32).reshape(-1, 3, 3), abi.SURF_BSDF_SMOOTH)
    got, missed = oracle.OracleScene(sc).vrls_to_long(src)
    assert missed == 0
    assert _close(got, np.array(gold["long"], np.float32), rtol=2e-6)


# ---- the reference's photon map (photonmapper/photonmap.h + kdtree.h compiled in place, oracle/ref_probe_pm.cpp) ----
PM_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_photonmap.json")


@pytest.fixture(scope="module")
def pm_gold():
    with open(PM_GOLD) as f:
        return json.load(f)


def _rel(got, want):
    want = np.asarray(want, np.float64)
    scale = np.abs(want).max()
    return np.abs(np.asarray(got, np.float64) - want).max() / scale if scale > 0 else np.abs(got).max()


def test_surface_estimates_match_reference_compiled_photonmap(pm_gold):
    """PhotonMap::estimateRadiance / estimateCausticRadiance over the reference's own PointKDTree::build + nnSearch, on
    1500 photons and 120 queries (k nearest with the radius shrunk to the k-th, not shrunk, k = 2; arbitrary surface
    normals; photons and camera on either side).  Same fp32 arithmetic, only the summation order differs: 2e-6."""
    ph = np.array(pm_gold["surface_photons"], np.float32)
    every = pm_gold["caustic_every"]
    n_lit = n_shrunk = 0
    for c in pm_gold["surface"]:
        q = np.zeros(1, abi.SURFACE_QUERY_DTYPE)
        q["p"], q["n"], q["wi"], q["reflectance"], q["weight"] = c["p"], c["n"], c["wi"], c["reflectance"], 1.0
        g = scenes.photonset_desc(ph[:, 0], ph[:, 1], ph[:, 2], scale=pm_gold["global_scale"], lookup_radius=c["radius"])
        got = oracle.surface_gather(q, global_desc=g, k_global=c["k"])[0]
        assert _rel(got, c["global"]) <= 2e-6, c
        cs = ph[::every]
        cd = scenes.photonset_desc(cs[:, 0], cs[:, 1], cs[:, 2], scale=pm_gold["caustic_scale"],
                                   lookup_radius=float(np.float32(0.8) * np.float32(c["radius"])))
        got_c = oracle.surface_gather(q, caustic_desc=cd, k_caustic=c["k"])[0]
        assert _rel(got_c, c["caustic"]) <= 2e-6, c
        n_lit += max(c["global"]) > 0
        d2 = ((ph[:, 0] - np.array(c["p"], np.float32)) ** 2).sum(1)
        n_shrunk += (d2 < np.float32(c["radius"]) ** 2).sum() > c["k"]
    assert n_lit >= 30 and 40 <= n_shrunk <= 100          # lit and unlit sides, shrunk and unshrunk radii all occur


def test_volume_estimate_matches_reference_compiled_photonmap(pm_gold):
    """PhotonMap::estimateRadianceVolume over the reference's PointKDTree::executeQuery, isotropic and HG (g = 0.6)."""
    vp = np.array(pm_gold["volume_photons"], np.float32)
    seen = 0
    for c in pm_gold["volume"]:
        med = scenes.make_medium(sigma_t=(1, 1, 1), albedo=(0.5,) * 3, phase=abi.PHASE_HG if c["hg"] else abi.PHASE_ISOTROPIC, g=pm_gold["hg_g"])
        pd = scenes.photonset_desc(vp[:, 0], vp[:, 1], vp[:, 2], scale=1.0, lookup_radius=c["radius"])
        got = oracle.estimate_radiance_volume(med, pd, c["p"], c["wo"], c["radius"])
        assert _rel(got, c["out"]) <= 2e-6, c
        seen += c["M"] > 0
    assert seen >= 40


def test_delta_bsdf_samples_match_reference_compiled_fresnel():
    """include/mitsuba/render/fresnel.h compiled in place (oracle/ref_probe_fresnel.cpp -> tests/golden/ref_fresnel.json):
    fresnel(), reflect(), refract() as SmoothDielectric::sample calls them, 300 directions x 5 indices (both sides, normal
    and grazing incidence, total internal reflection, index-matched).  The oracle's sample_delta reproduces them BIT-EXACT:
    the reflect / refract decision flips between r and the next float above it, the directions and the eta_ti^2 radiance
    scaling are identical.  The default conductor's Fresnel term (eta = 0, k = 1) is exactly 1 at every angle, which is
    what the surface word's MIRROR class assumes."""
    with open(os.path.join(os.path.dirname(GOLD), "ref_fresnel.json")) as f:
        gold = json.load(f)
    n_r = n_t = n_tir = 0
    for c in gold["dielectric"]:
        wi, r = np.array(c["wi"], np.float32), np.float32(c["r"])
        prm = (c["eta"], 1.0, 1.0)
        if r > 0:
            wo, wt, transmitted = oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, prm, (0, 0, 1), -wi, float(r))
            assert not transmitted and np.array_equal(wo, np.array(c["reflect"], np.float32)) and np.all(wt == 1.0), c
            n_r += 1
        if r < 1:
            above = float(np.nextafter(r, np.float32(2)))
            wo, wt, transmitted = oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, prm, (0, 0, 1), -wi, above)
            assert transmitted and np.array_equal(wo, np.array(c["refract"], np.float32)), c
            assert np.all(wt == np.float32(c["eta_ti"]) * np.float32(c["eta_ti"])), c
            n_t += 1
        n_tir += r == 1 and wi[2] != 0
    assert n_r >= 200 and n_t >= 200 and n_tir >= 10
    assert all(f == 1.0 for _, f in gold["conductor_default"])


def test_grid_transmittance_matches_reference_compiled_medium():
    """Medium::evalTransmittance of the reference (include/mitsuba/render/medium.h:109-155 over src/librender/medium.cpp's
    sample_interaction / eval_tr_and_pdf, compiled in place by oracle/ref_probe_medium.cpp) for a grid medium, on the same
    PCG32 sub-streams: the oracle's grid_transmittance gives the same value -- the same number of accepted free-flight steps
    (the product of exp(-majorant * step) over them) -- for segments inside the grid box, starting outside it, leaving it,
    axis-parallel and very short; three majorants."""
    import json, os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_medium.json")) as f:
        g = json.load(f)
    worst, n_multi = 0.0, 0
    by_maj = {}
    for c in g["cases"]:
        by_maj.setdefault(round(c["majorant"], 6), []).append(c)
    assert len(by_maj) == 3
    for maj, cases in by_maj.items():
        sc = scenes.Scene()
        grid = np.full((4, 4, 4), 0.8, np.float32)
        m = scenes.make_medium(kind=abi.MEDIUM_GRID, sigma_t=(1, 1, 1), albedo=(0.5,) * 3, scale=maj / 0.8)
        m.grid_bbox_min[:] = g["grid_lo"]
        m.grid_bbox_max[:] = g["grid_hi"]
        sc.add_medium(m, grid=grid)
        osc = oracle.OracleScene(sc)
        for c in cases:
            tr = osc.grid_transmittance(c["o"], c["d"], c["maxt"], c["seed"])
            err = abs(tr - c["tr"]) / max(c["tr"], 1e-30)
            worst = max(worst, err)
            assert err <= 3e-6, (c, tr)                      # Enoki's exp / log polynomials against glibc's over <= ~20 factors
            n_multi += int(c["draws"] >= 4)
    assert n_multi >= 100                                    # most cases chain several accept steps
    # homoEvalTransmittance, medium.h:103-107
    for length, tr in g["homo"]:
        assert abs(np.exp(np.float32(-g["homo_density"]) * np.float32(length)) - tr) <= 2e-7 * max(tr, 1e-3) + 1e-9
    print(f"grid transmittance vs reference: max rel err {worst:.2e}")

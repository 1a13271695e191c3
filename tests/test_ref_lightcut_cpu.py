"""CPU suite (-m "not gpu"): pins the oracle's VRL light cut against the REFERENCE'S OWN code.

tests/golden/ref_lightcut.json was produced by oracle/_ref/refprobe_lc = the reference's src/integrators/vrl/vrl_lightcut.h
(with vrl_struct.h, include/mitsuba/core/bbox.h and core/vector.h) compiled where it lies against the mock Mitsuba of
oracle/ref_mock.h (generator oracle/ref_probe_lc.cpp + oracle/make_golden.py).  It holds, with their inputs:
  merge      Node::dist and Node(n1, n2, v, ...)            vrl_lightcut.h:101-151
  tree       the light tree of 192 VRLs in preorder          vrl_lightcut.h:196-212, 515-627
  box_dist   BoundingBox::getMinDistanceSqr(ray)             bbox.h:335-816
  seg_dist   VRLLightCut::sqrDistanceVRL                     vrl_lightcut.h:772-838
  bound      VRLLightCut::getClusterUpperBound               vrl_lightcut.h:429-512
  query      VRLLightCut::query: Li, nb_evaluation, draws    vrl_lightcut.h:243-393

Tolerances: dist / merge / tree are +,*,/,max in the reference's order -> bit-exact.  The distances are compared at 1e-5
relative (+1e-6 absolute, the line-box query cancels), bounds at 1e-5, query sums at 2e-5 (as getContrib in
test_ref_golden_cpu.py)."""
import json
import os

import numpy as np
import pytest

import oracle
from mitsuba_nlvrl_b200 import abi, scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_lightcut.json")


def _num(x):
    return float(x) if not isinstance(x, str) else float(x)


def _arr(xs):
    return np.array([_num(x) for x in xs], np.float32)


@pytest.fixture(scope="module")
def gold():
    with open(GOLD) as f:
        return json.load(f)


def _scene(gold, occluders, hg=False):
    m = gold["medium"]
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=m["sigma_t"], albedo=m["albedo"], scale=m["scale"],
                                     phase=abi.PHASE_HG if hg else abi.PHASE_ISOTROPIC, g=m["g"]))
    if occluders:
        sc.add_tris(np.array(gold["triangles"], np.float32).reshape(-1, 3, 3), abi.SURF_BSDF_SMOOTH)
    return sc


def _vrl_desc(gold):
    soa = np.ascontiguousarray(np.array(gold["vrls"], np.float32).T)
    return soa, scenes.vrlset_desc(soa, scale=1.0)


def test_merge_metric_and_representative_are_bit_exact(gold):
    for c in gold["merge"]:
        a, b = _arr(c["a"]), _arr(c["b"])
        assert oracle.lc_node_dist(a, b) == np.float32(c["dist"][0])
        assert oracle.lc_node_dist(b, a) == np.float32(c["dist"][1])
        m = oracle.lc_merge(a, b, c["v"])
        assert m["pick"] == c["pick"]
        for k in ("flux", "add_flux", "lo", "hi"):
            assert np.array_equal(m[k], _arr(c[k])), (k, m[k], c[k])
        assert m["max_len"] == np.float32(c["max_len"])


def test_tree_is_the_reference_tree(gold):
    """Same clusters, same order of merges, same representatives (one sampler draw per merge), bit-exact fluxes and boxes."""
    soa, vd = _vrl_desc(gold)
    lc = oracle.OracleLightcut(vd, seed_offset=gold["build_seed"], n_threads=2)
    assert lc.build_draws == gold["build_draws"] == len(gold["vrls"]) - 1
    f, i = lc.nodes()
    assert len(f) == len(gold["tree"]) == 2 * len(gold["vrls"]) - 1
    for k, n in enumerate(gold["tree"]):
        assert i[k, 0] == n["leaf"] and i[k, 1] == n["rep"] and i[k, 2] == n["same"] and i[k, 3] == n["n_leaves"], (k, i[k], n)
        assert np.array_equal(f[k, 0:3], _arr(n["flux"])) and np.array_equal(f[k, 3:6], _arr(n["add_flux"])), (k, f[k], n)
        assert np.array_equal(f[k, 6:9], _arr(n["lo"])) and np.array_equal(f[k, 9:12], _arr(n["hi"])), (k, f[k], n)
        assert f[k, 12] == np.float32(n["max_len"])


def test_box_distance_matches_reference(gold):
    """Includes the fork's quirks: distance 0 whenever the supporting LINE meets the box, and the line query run on a
    box of the full extents (oracle/vrl_oracle.cpp, box_ray_sqr)."""
    zeros = 0
    for c in gold["box_dist"]:
        got = oracle.lc_box_dist(c["lo"], c["hi"], c["ray"])
        want = _num(c["out"])
        zeros += want == 0.0
        assert abs(got - want) <= 1e-5 * abs(want) + 1e-6, (c, got)
    assert 10 <= zeros <= len(gold["box_dist"]) - 100


def test_segment_distance_matches_reference(gold):
    for c in gold["seg_dist"]:
        got = oracle.lc_seg_dist(_arr(c["vrl"]), c["ray"])
        want = _num(c["out"])
        assert abs(got - want) <= 1e-5 * abs(want) + 1e-7, (c, got)


def _parts(gold, k):
    return np.array(gold["rays"][k], np.float32).reshape(-1, 7)


def test_cluster_bound_matches_reference(gold):
    soa, vd = _vrl_desc(gold)
    lc = oracle.OracleLightcut(vd, seed_offset=gold["build_seed"], n_threads=2)
    osc = [oracle.OracleScene(_scene(gold, False, hg=False)), oracle.OracleScene(_scene(gold, False, hg=True))]
    n_max, n_box, n_exact = 0, 0, 0
    f, i = lc.nodes()
    for hg, node, k, want in gold["bound"]:
        want = _num(want)
        got = lc.bound(osc[hg], node, _parts(gold, k))
        if want >= 3e38:
            n_max += 1
            assert got >= 3e38, (hg, node, k, got)
            continue
        n_box += i[node, 3] > 8
        n_exact += i[node, 3] <= 8
        assert abs(got - want) <= 1e-5 * abs(want), (hg, node, k, got, want)
    assert n_max >= 20 and n_box >= 50 and n_exact >= 300, (n_max, n_box, n_exact)


def test_query_matches_reference_sequential_stream(gold):
    """VRLLightCut::query on the reference's one sequential PCG32 stream: same Li, same number of fresh evaluations, same
    number of draws, for importance / uniform sampling, with occluders, HG phase, thresholds 0.02 / 0.1 / 0.5, single parts
    and 3-part paths.  A query whose cut hinges on a decision closer than 1e-4 to its threshold may differ (expf / atan
    rounding): at most a few, and then only within the cut's own error bound."""
    soa, vd = _vrl_desc(gold)
    lc = oracle.OracleLightcut(vd, seed_offset=gold["build_seed"], n_threads=2)
    scenes_by = {}
    fragile = 0
    for c in gold["query"]:
        key = (c["vis"], c["hg"])
        if key not in scenes_by:
            scenes_by[key] = oracle.OracleScene(_scene(gold, bool(c["vis"]), hg=bool(c["hg"])))
        opt = scenes.default_options(use_uniform_sampling=c["uniform"], visibility=c["vis"], math_mode=0, use_light_cut=1,
                                     threshold_error=c["err"], threshold_better_dist=c["better"])
        got, st = lc.query(scenes_by[key], _parts(gold, c["ray"]), opt, c["seed"], rng_mode=oracle.RNG_SEQUENTIAL)
        want = _arr(c["out"])
        same = st["evals"] == c["nb"] and st["draws"] == c["draws"] and np.all(np.abs(got - want) <= 2e-5 * np.abs(want) + 1e-12)
        if not same:
            assert st["min_margin"] < 1e-4, (c, got, st)
            assert np.all(np.abs(got - want) <= 2 * c["err"] * np.abs(want).max()), (c, got, st)
            fragile += 1
    assert fragile <= 2, fragile


def test_addressed_draws_give_the_same_cut_statistics(gold):
    """The GPU's addressing ((sample * N + representative VRL) * 2) changes which uniforms an evaluation reads, nothing else:
    over the 40 rays the mean Li agrees with the sequential stream's within Monte-Carlo noise and the evaluation counts
    stay in the same range."""
    soa, vd = _vrl_desc(gold)
    lc = oracle.OracleLightcut(vd, seed_offset=gold["build_seed"], n_threads=2)
    osc = oracle.OracleScene(_scene(gold, False))
    opt = scenes.default_options(visibility=0, math_mode=0, use_light_cut=1, threshold_error=0.02, threshold_better_dist=8)
    a, b, ea, eb = [], [], 0, 0
    for k in range(len(gold["rays"])):
        for rep in range(8):
            ga, sa = lc.query(osc, _parts(gold, k), opt, 900 + 40 * rep + k, rng_mode=oracle.RNG_ADDRESSED)
            gb, sb = lc.query(osc, _parts(gold, k), opt, 900 + 40 * rep + k, rng_mode=oracle.RNG_SEQUENTIAL)
            a.append(ga); b.append(gb); ea += sa["evals"]; eb += sb["evals"]
    a, b = np.array(a, np.float64), np.array(b, np.float64)
    assert abs(ea - eb) <= 0.05 * eb
    finite = np.all(a < 1e3, axis=1) & np.all(b < 1e3, axis=1)           # rays through a VRL have unbounded variance
    assert abs(a[finite].mean() - b[finite].mean()) <= 0.1 * b[finite].mean()

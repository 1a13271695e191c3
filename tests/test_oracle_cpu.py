"""CPU suite (-m "not gpu"): pins the oracle against the golden vectors generated from the reference's own
Enoki headers (tests/golden/enoki_pcg32.json, oracle/enoki_probe.cpp), the known answers the reference's
tests hold for this path (src/samplers/tests/test_independent.py:28-33, src/phase/tests/test_isotropic.py:14-26)
and closed-form properties of the sampling routines ("closed-form checks")."""
import ctypes as C
import json
import math
import os

import numpy as np
import pytest

import oracle
from mitsuba_nlvrl_b200 import abi, api, scenes

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DEFAULT_STATE, DEFAULT_STREAM = 0x853c49e6748fea9b, 0xda3e39cb94b95bdb


@pytest.fixture(scope="module")
def gold():
    with open(os.path.join(GOLD, "enoki_pcg32.json")) as f:
        return json.load(f)


def _pcg(initstate, initseq=DEFAULT_STREAM):
    st, inc = C.c_uint64(), C.c_uint64()
    oracle.lib().vrlo_pcg32_seed(initstate, initseq, C.byref(st), C.byref(inc))
    return st, inc


def test_pcg32_default_stream_matches_enoki(gold):
    st, inc = _pcg(DEFAULT_STATE)
    got = [oracle.lib().vrlo_pcg32_next_u32(C.byref(st), inc.value) for _ in range(30)]
    assert got == gold["default_u32"]
    f = np.array([((u >> 9) | 0x3f800000) for u in got], np.uint32).view(np.float32) - np.float32(1.0)
    assert np.array_equal(f, np.array(gold["default_f32"], np.float32))


def test_sampler_seeding_matches_enoki(gold):
    for seed, want in gold["seeded"].items():
        out_f, out_u = np.zeros(16, np.float32), np.zeros(16, np.uint32)
        oracle.lib().vrlo_sampler_stream(0, int(seed), 16, oracle.fp(out_f), out_u.ctypes.data_as(C.POINTER(C.c_uint32)))
        assert out_u.tolist() == want, seed
        # base_seed + offset is what is seeded (sampler.cpp:87-94)
        oracle.lib().vrlo_sampler_stream(int(seed), 0, 16, oracle.fp(out_f), out_u.ctypes.data_as(C.POINTER(C.c_uint32)))
        assert out_u.tolist() == want, seed


def test_pcg32_advance_matches_enoki(gold):
    for delta, state, nxt in gold["advance"]:
        st, inc = _pcg(42)
        oracle.lib().vrlo_pcg32_advance(C.byref(st), inc.value, delta)
        # the probe's printf evaluates its arguments right to left: `state` is the state AFTER the draw
        assert oracle.lib().vrlo_pcg32_next_u32(C.byref(st), inc.value) == nxt, delta
        assert st.value == int(state), delta


def test_morton_matches_enoki(gold):
    for i, x, y in gold["morton"]:
        gx, gy = C.c_uint32(), C.c_uint32()
        oracle.lib().vrlo_morton_decode(i, C.byref(gx), C.byref(gy))
        assert (gx.value, gy.value) == (x, y)


def test_spiral_ids_are_a_permutation_starting_at_centre():
    w, h, bs = 1920, 1080, 32
    nbx, nby = (w + bs - 1) // bs, (h + bs - 1) // bs
    ids = np.zeros(nbx * nby, np.uint32)
    n = oracle.lib().vrlo_spiral_block_ids(w, h, bs, ids.ctypes.data_as(C.POINTER(C.c_uint32)), len(ids))
    assert n == nbx * nby and sorted(ids.tolist()) == list(range(n))
    assert ids.reshape(nby, nbx)[nby // 2, nbx // 2] == 0          # spiral.cpp:27-33 starts at the centre block


def test_isotropic_known_answer():
    """src/phase/tests/test_isotropic.py:14-26: eval == 1/(4 pi) for any directions."""
    m = scenes.make_medium(phase=abi.PHASE_ISOTROPIC)
    rng = np.random.default_rng(0)
    for _ in range(20):
        wi, wo = (oracle.f32(v / np.linalg.norm(v)) for v in rng.normal(size=(2, 3)))
        assert oracle.lib().vrlo_phase_eval(C.byref(m), oracle.fp(wi), oracle.fp(wo)) == pytest.approx(1 / (4 * math.pi), rel=1e-7)


def test_hg_normalised_and_matches_formula():
    """hg.cpp:51-54: 1/4pi (1-g^2) / (1+g^2+2g cos)^1.5, integrates to one over the sphere (chi2 pin of test_hg.py)."""
    g = 0.6
    m = scenes.make_medium(phase=abi.PHASE_HG, g=g)
    wi = oracle.f32([0, 0, 1])
    mu = np.linspace(-1, 1, 4001)
    vals = []
    for c in mu:
        wo = oracle.f32([math.sqrt(max(0, 1 - c * c)), 0, c])
        vals.append(oracle.lib().vrlo_phase_eval(C.byref(m), oracle.fp(wi), oracle.fp(wo)))
    vals = np.array(vals, np.float64)
    want = (1 - g * g) / (4 * math.pi * (1 + g * g + 2 * g * mu) ** 1.5)
    assert np.allclose(vals, want, rtol=2e-6)
    assert 2 * math.pi * np.trapezoid(vals, mu) == pytest.approx(1.0, rel=1e-4)


def _closest(vo, vd, vl, ro, rd, rl):
    tc, tv = C.c_float(), C.c_float()
    a = [oracle.f32(x) for x in (vo, vd, ro, rd)]
    oracle.lib().vrlo_find_closest_point(oracle.fp(a[0]), oracle.fp(a[1]), vl, oracle.fp(a[2]), oracle.fp(a[3]), rl,
                                         C.cast(C.byref(tc), C.POINTER(C.c_float)), C.cast(C.byref(tv), C.POINTER(C.c_float)))
    return tc.value, tv.value


def test_find_closest_point_vs_brute_force():
    rng = np.random.default_rng(1)
    for _ in range(50):
        vo, ro = rng.uniform(-5, 5, (2, 3))
        vd, rd = (v / np.linalg.norm(v) for v in rng.normal(size=(2, 3)))
        vl, rl = rng.uniform(0.5, 8, 2)
        tc, tv = _closest(vo, vd, vl, ro, rd, rl)
        s = np.linspace(0, vl, 801)[:, None, None]
        t = np.linspace(0, rl, 801)[None, :, None]
        d2 = (((vo + s * vd) - (ro + t * rd)) ** 2).sum(-1)
        got = np.linalg.norm((vo + tv * vd) - (ro + tc * rd))
        assert got <= math.sqrt(d2.min()) * (1 + 1e-4) + 1e-4
        assert -1e-6 <= tc <= rl * (1 + 1e-6) and -1e-6 <= tv <= vl * (1 + 1e-6)


def test_find_closest_point_perpendicular_skew():
    tc, tv = _closest([0, 0, 0], [1, 0, 0], 4.0, [2, 3, -2], [0, 0, 1], 4.0)
    assert (tc, tv) == pytest.approx((2.0, 2.0))


def _sample(vrl, ray, uniform, u1, u2):
    out = np.zeros(9, np.float32)
    a = [oracle.f32(x) for x in (vrl[0], vrl[1], ray[0], ray[1])]
    ok = oracle.lib().vrlo_sampling_vrl(oracle.fp(a[0]), oracle.fp(a[1]), vrl[2], oracle.fp(a[2]), oracle.fp(a[3]), ray[2],
                                        uniform, u1, u2, oracle.fp(out))
    return ok, out


def test_importance_and_uniform_sampling_estimate_the_same_integral():
    """E[f * invPDF] must equal the double integral of f = 1/d^2 over both segments for both strategies
    (vrl_struct.h:239-243 vs 251-360); checked against midpoint quadrature."""
    vrl = (np.array([0.0, 0, 0]), np.array([1.0, 0, 0]), 3.0)
    ray = (np.array([1.0, 2.0, -2.0]), np.array([0.0, 0.0, 1.0]), 5.0)
    n = 400
    s = (np.arange(n) + 0.5) / n * vrl[2]
    t = (np.arange(n) + 0.5) / n * ray[2]
    P = vrl[0] + s[:, None, None] * vrl[1]
    Q = ray[0] + t[None, :, None] * ray[1]
    quad = (1.0 / ((P - Q) ** 2).sum(-1)).mean() * vrl[2] * ray[2]
    m = 96
    us = (np.arange(m) + 0.5) / m
    for uniform in (0, 1):
        acc = 0.0
        for u1 in us:
            for u2 in us:
                ok, o = _sample(vrl, ray, uniform, float(u1), float(u2))
                assert ok
                d2 = ((o[0:3].astype(np.float64) - o[4:7]) ** 2).sum()
                acc += o[8] / d2
        assert acc / (m * m) == pytest.approx(quad, rel=2e-3 if not uniform else 2e-2)


def test_sampling_early_outs():
    # parallel segments: sin(theta) == 0 -> invalid (vrl_struct.h:275-278)
    ok, _ = _sample((np.zeros(3), np.array([0, 0, 1.0]), 2.0), (np.array([1.0, 0, 0]), np.array([0, 0, 1.0]), 2.0), 0, 0.3, 0.6)
    assert ok == 0
    # intersecting segments: h == 0 -> invalid (vrl_struct.h:264-265)
    ok, _ = _sample((np.array([-1.0, 0, 0]), np.array([1.0, 0, 0]), 2.0), (np.array([0, -1.0, 0]), np.array([0, 1.0, 0]), 2.0), 0, 0.3, 0.6)
    assert ok == 0


def test_homogeneous_transmittance_is_grey_mean_sigma_t():
    """medium.h:103-117 + constant3d.cpp:37: exp(-scale * mean(sigma_t) * t), the same for the three channels."""
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=(0.3, 0.6, 0.9), albedo=(0.5,) * 3, scale=2.0))
    osc = oracle.OracleScene(sc)
    tr = osc.transmittance([0, 0, 0], [0, 0, 1.5], medium=0)
    assert np.allclose(tr, math.exp(-2.0 * 0.6 * 1.5), rtol=1e-6) and tr[0] == tr[1] == tr[2]


def test_grid_volume_trilinear_known_answers():
    """GridVolume::eval (grid3d.cpp:236-340), as src/textures/tests/test_texture3d.py checks it: voxel centres return
    the stored value, a grid that is linear in x is reproduced between centres, outside points see the clamped border."""
    res = (8, 5, 6)
    sc = scenes.grid_fog_scene(occluders=False, res=res)
    grid = scenes.smoke_grid(res)
    osc = oracle.OracleScene(sc)
    e = scenes.BOX + 0.01
    ext = 2 * e
    for (ix, iy, iz) in [(0, 0, 0), (3, 2, 4), (7, 4, 5)]:
        p = [-e + (ix + .5) / res[0] * ext, -e + (iy + .5) / res[1] * ext, -e + (iz + .5) / res[2] * ext]
        assert abs(osc.grid_eval(p) - grid[iz, iy, ix]) <= 2e-6
    assert abs(osc.grid_eval([-e - 3.0, -e - 1.0, -e - 2.0]) - grid[0, 0, 0]) <= 1e-7
    lin = scenes.Scene()
    g = np.broadcast_to(np.arange(8, dtype=np.float32)[None, None, :], (6, 5, 8))
    m = scenes.make_medium(kind=abi.MEDIUM_GRID, scale=1.0)
    m.grid_bbox_min[:] = [0, 0, 0]
    m.grid_bbox_max[:] = [8, 5, 6]
    lin.add_medium(m, grid=g)
    ol = oracle.OracleScene(lin)
    for x in (0.5, 1.25, 3.9, 7.5):
        assert abs(ol.grid_eval([x, 2.2, 3.3]) - (x - 0.5)) <= 1e-5


def test_grid_transmittance_is_the_references_biased_majorant_estimator():
    """medium.h:119-154: the product of exp(-majorant * step) over the ACCEPTED free-flight steps only, i.e.
    exp(-majorant * T_last) with T_last the last Poisson arrival before L.  Closed form of its mean:
    E = exp(-mu L) (1 + mu L)  (the true transmittance would be exp(-mu L)); every sample lies in [exp(-mu L), 1]."""
    sc = scenes.grid_fog_scene(occluders=False, res=(8, 5, 6), majorant=0.4)
    osc = oracle.OracleScene(sc)
    mu, L = 0.4, 6.0
    o, d = [-3.0, 0.5, -1.0], [1.0, 0.0, 0.0]
    v = np.array([osc.grid_transmittance(o, d, L, seed) for seed in range(20000)])
    assert v.min() >= np.exp(-mu * L) * (1 - 1e-5) and v.max() <= 1.0
    want = np.exp(-mu * L) * (1 + mu * L)
    assert abs(v.mean() - want) <= 4 * v.std() / np.sqrt(len(v))
    assert osc.grid_transmittance(o, d, L, 5) == osc.grid_transmittance(o, d, L, 5)


def test_shadow_query_occluder_and_null_boundary():
    """vrl_struct.h:525-569: a non-null surface zeroes the transmittance; Null::eval returns 0 (null.cpp:59-62),
    so does an index-matched boundary."""
    sc = scenes.box_fog_scene(occluders=True)
    osc = oracle.OracleScene(sc)
    assert np.all(osc.transmittance([4.5, -2, -9], [4.5, -2, 9]) == 0)           # through the sphere
    free = osc.transmittance([-8, 8, -8], [-8, 8, 8])
    assert np.allclose(free, math.exp(-0.11 * 16), rtol=1e-5)
    assert np.all(osc.transmittance([0, 0, -9], [0, 0, -10.5]) == 0)             # across the null fog boundary


def test_oracle_ray_intersect_vs_naive():
    """kd-tree vs naive (src/librender/tests/test_kdtrees.py): the oracle's BVH must return the brute-force closest hit."""
    sc = scenes.box_fog_scene(occluders=True)
    osc = oracle.OracleScene(sc)
    V = sc.tri_verts.astype(np.float64)
    rng = np.random.default_rng(5)
    for _ in range(60):
        o = rng.uniform(-9, 9, 3)
        d = rng.normal(size=3)
        d /= np.linalg.norm(d)
        e1, e2 = V[:, 1] - V[:, 0], V[:, 2] - V[:, 0]
        p = np.cross(d, e2)
        det = (e1 * p).sum(1)
        with np.errstate(divide="ignore", invalid="ignore"):
            inv = 1 / det
            tv = o - V[:, 0]
            u = (tv * p).sum(1) * inv
            q = np.cross(tv, e1)
            v = (q * d).sum(1) * inv
            t = (e2 * q).sum(1) * inv
        okm = (u >= 0) & (v >= 0) & (u + v <= 1) & (t > 1e-4) & np.isfinite(t)
        prim, tt, _, _ = osc.intersect(o, d, 1e-4)
        if okm.any():
            assert prim >= 0 and tt == pytest.approx(t[okm].min(), rel=1e-4, abs=1e-4)
        else:
            assert prim < 0


def test_sequential_and_addressed_streams_agree_without_early_outs():
    """the addressed PCG32 convention equals the reference's sequential consumption whenever no
    eval takes an early-out before its draws (the generic case)."""
    sc = scenes.box_fog_scene(occluders=True)
    cam = scenes.box_fog_camera(16, 9)
    opt = scenes.default_options()
    soa = scenes.synthetic_vrls(300, seed=9)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    a, _, ev_a = osc.render(cam, opt, vd, rng_mode=oracle.RNG_ADDRESSED)
    b, _, ev_b = osc.render(cam, opt, vd, rng_mode=oracle.RNG_SEQUENTIAL)
    assert ev_a == ev_b == 16 * 9 * 300
    assert np.array_equal(a, b)


def test_film_contract_xyzaw():
    """integrator.cpp:282-294 / imageblock.cpp:162-169: XYZ = srgb_to_xyz(rgb), A = valid ray, W = 1 per sample."""
    sc = scenes.box_fog_scene(occluders=False)
    cam = scenes.box_fog_camera(16, 9)
    opt = scenes.default_options(visibility=0, spp=2)
    soa = scenes.synthetic_vrls(64, seed=2)
    xyzaw, rgb, _ = oracle.OracleScene(sc).render(cam, opt, scenes.vrlset_desc(soa))
    assert np.all(xyzaw[..., 4] <= 2.0) and xyzaw[..., 4].sum() == pytest.approx(16 * 9 * 2, abs=8)   # jitter may splat next door
    M = np.array([[0.412453, 0.357580, 0.180423], [0.212671, 0.715160, 0.072169], [0.019334, 0.119193, 0.950227]])
    one = scenes.default_options(visibility=0, spp=1)
    xyzaw1, rgb1, _ = oracle.OracleScene(sc).render(cam, one, scenes.vrlset_desc(soa))
    assert np.allclose(xyzaw1[..., :3], rgb1 @ M.T, rtol=1e-5, atol=1e-12)
    assert np.all(xyzaw1[..., 4] == 1.0)


# ------------------------------------------------------------------------------------------ C ABI (no GPU)
def test_library_exports_every_declared_symbol():
    import re
    hdr = open(os.path.join(os.path.dirname(GOLD), "..", "include", "vrl_b200.h")).read()
    declared = set(re.findall(r"\b(vrlb_[a-z0-9_]+)\s*\(", hdr))
    lib = api.load_library()
    missing = [s for s in sorted(declared) if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(api.EXPORTS), declared ^ set(api.EXPORTS)
    assert lib.vrlb_abi_version() == 3


def test_pod_layouts_match_header(tmp_path):
    """sizeof of every POD as the C compiler lays out include/vrl_b200.h == the ctypes mirror."""
    import subprocess
    names = {"vrlb_segment": abi.Segment, "vrlb_options": abi.Options, "vrlb_camera": abi.Camera, "vrlb_vrlset_desc": abi.VrlSetDesc,
             "vrlb_timing": abi.Timing, "vrlb_medium": abi.Medium, "vrlb_scene_desc": abi.SceneDesc, "vrlb_emitter": abi.Emitter,
             "vrlb_lightcut_stats": abi.LightcutStats, "vrlb_generate_desc": abi.GenerateDesc, "vrlb_photonset_desc": abi.PhotonSetDesc}
    src = tmp_path / "sz.c"
    body = "".join(f'printf("{n} %zu\\n", sizeof({n}));' for n in names)
    src.write_text('#include <stdio.h>\n#include "vrl_b200.h"\nint main(void){' + body + "return 0;}")
    exe = tmp_path / "sz"
    inc = os.path.join(os.path.dirname(GOLD), "..", "include")
    subprocess.check_call(["gcc", "-std=c99", "-I", inc, "-o", str(exe), str(src)])      # the header is plain C
    for line in subprocess.check_output([str(exe)]).decode().splitlines():
        n, sz = line.split()
        assert C.sizeof(names[n]) == int(sz), n
    assert C.sizeof(abi.Segment) == 64 and abi.SEGMENT_DTYPE.itemsize == 64


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(api.VrlbError) as e:
        api.Context(0)
    assert e.value.code == abi.ERR_CUDA


def test_photon_direct_term_single_photon_closed_form():
    """Direct in-scattering term (vrl.cpp:523-611): one photon sitting on the first gather point of a segment.  The radius
    jitter is the draw just before the segment's query position; the first gather point is at t = r; the estimate is
    power * pf * (1 - 0)^2 / r^2 * 3 / pi, weighted by the throughput, exp(-sigma r) and the map's scale."""
    L = oracle.lib()
    sc = scenes.Scene()
    med = scenes.make_medium(sigma_t=(0.3, 0.3, 0.3), albedo=(0.8,) * 3, scale=2.0)
    sc.add_medium(med)
    osc = oracle.OracleScene(sc)
    st, inc = _pcg(1234)
    bits = L.vrlo_pcg32_next_u32(C.byref(st), inc.value)
    u = np.array([(bits >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - np.float32(1.0)
    R = np.float32(0.5)
    r = R * (np.float32(1.25) * u + (np.float32(0.75) - np.float32(0.75) * u))
    seg = np.zeros(1, abi.SEGMENT_DTYPE)
    seg["o"][0], seg["d"][0], seg["maxt"][0] = (1, 2, 3), (0, 0, 1), 1.5 * r        # room for exactly one gather point
    seg["weight"][0] = (0.5, 0.25, 1.0)
    seg["rng_state"][0] = st.value                                                  # the query starts after the jitter draw
    power = np.array([[2.0, 3.0, 4.0]], np.float32)
    pd = scenes.photonset_desc([[1, 2, 3 + r]], [[0, 1, 0]], power, scale=0.125, lookup_radius=float(R))
    opt = scenes.default_options()
    got = osc.photon_gather(pd, seg, opt)[0]
    want = np.array([0.5, 0.25, 1.0]) * math.exp(-0.6 * r) * power[0] / (4 * math.pi) * 3 / (math.pi * r * r) * 0.125
    assert np.allclose(got, want, rtol=2e-6)
    # a photon just outside the radius contributes nothing; a segment shorter than r has no gather point
    pd2 = scenes.photonset_desc([[1 + 1.001 * r, 2, 3 + r]], [[0, 1, 0]], power, scale=0.125, lookup_radius=float(R))
    assert np.all(osc.photon_gather(pd2, seg, opt) == 0)
    seg["maxt"][0] = 0.99 * r
    assert np.all(osc.photon_gather(pd, seg, opt) == 0)
    # gather points at r, 3r, 5r ...: a photon at 3r is seen through exp(-sigma 3r) (two transmittance factors multiplied)
    seg["maxt"][0] = 3.5 * r
    pd3 = scenes.photonset_desc([[1, 2, 3 + 3 * r]], [[0, 1, 0]], power, scale=1.0, lookup_radius=float(R))
    got3 = osc.photon_gather(pd3, seg, opt)[0]
    want3 = np.array([0.5, 0.25, 1.0]) * math.exp(-0.6 * 3 * r) * power[0] / (4 * math.pi) * 3 / (math.pi * r * r)
    assert np.allclose(got3, want3, rtol=1e-5)


def test_photon_direct_term_carries_step_phase_across_parts():
    """Curved camera paths (vrl.cpp:553-608): a path is n consecutive part records, n in the high half of `part`.  Cutting a
    straight segment into collinear parts must not move the gather points (the step phase `leftOver` is carried), and on
    a bent path the transmittance runs over the straight distance between consecutive gather points."""
    L = oracle.lib()
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=(0.3, 0.3, 0.3), albedo=(0.8,) * 3, scale=2.0))
    osc = oracle.OracleScene(sc)
    st, inc = _pcg(99)
    bits = L.vrlo_pcg32_next_u32(C.byref(st), inc.value)
    u = np.array([(bits >> 9) | 0x3f800000], np.uint32).view(np.float32)[0] - np.float32(1.0)
    R = 0.4
    r = float(np.float32(R) * (np.float32(1.25) * u + (np.float32(0.75) - np.float32(0.75) * u)))
    pos, pdir, power = scenes.synthetic_photons(4000, seed=4, box=3.0)
    pd = scenes.photonset_desc(pos, pdir, power, scale=1.0 / 3999, lookup_radius=R)
    opt = scenes.default_options()
    whole = np.zeros(1, abi.SEGMENT_DTYPE)
    whole["o"][0], whole["d"][0], whole["maxt"][0] = (-2.5, 0.2, -0.3), (1, 0, 0), 5.0
    whole["weight"][0] = (1.0, 0.5, 0.25)
    whole["rng_state"][0] = st.value
    want = osc.photon_gather(pd, whole, opt)[0]
    assert want.min() > 0
    cuts = [0.0, 0.3 * r, 1.7 * r, 1.7 * r + 2.0, 5.0]              # a part shorter than r, one without any gather point
    parts = np.zeros(len(cuts) - 1, abi.SEGMENT_DTYPE)
    for k in range(len(parts)):
        parts["o"][k] = (-2.5 + cuts[k], 0.2, -0.3)
        parts["d"][k], parts["maxt"][k] = (1, 0, 0), cuts[k + 1] - cuts[k]
        parts["part"][k] = k | (len(parts) << 16)
    parts["weight"][:] = whole["weight"][0]
    parts["weight"][1:] *= 0.123                                      # only the first part's throughput is read
    parts["rng_state"][:] = st.value
    got = osc.photon_gather(pd, parts, opt)
    assert np.all(got[1:] == 0)                                       # the path's term sits on its first record
    assert np.allclose(got[0], want, rtol=2e-5)
    # without the group size every part would restart its own walk: the carried phase matters
    solo = parts.copy(); solo["part"] = 0
    assert not np.allclose(osc.photon_gather(pd, solo, opt).sum(0), want, rtol=1e-3)
    # a right-angle bend at t = 2 r with one photon on the second gather point (1 r past the corner): it is seen through
    # exp(-sigma (r + sqrt(2) r)), chord not arc length, and its phase function takes the FIRST part's direction
    hg = scenes.Scene()
    hg.add_medium(scenes.make_medium(sigma_t=(0.3, 0.3, 0.3), albedo=(0.8,) * 3, scale=2.0, phase=abi.PHASE_HG, g=0.5))
    bend = np.zeros(2, abi.SEGMENT_DTYPE)
    bend["o"][0], bend["d"][0], bend["maxt"][0] = (0, 0, 0), (0, 0, 1), 2 * r
    bend["o"][1], bend["d"][1], bend["maxt"][1] = (0, 0, 2 * r), (1, 0, 0), 1.5 * r
    bend["part"] = [0 | 2 << 16, 1 | 2 << 16]
    bend["weight"][:] = 1.0
    bend["rng_state"][:] = st.value
    p1 = scenes.photonset_desc([[r, 0, 2 * r]], [[0, 0, -1]], [[1.0, 1.0, 1.0]], scale=1.0, lookup_radius=R)
    got = oracle.OracleScene(hg).photon_gather(p1, bend, opt)[0]
    g = 0.5
    cos_theta = 1.0                                                   # wo = d0 = +z, wi = -photon.direction = +z
    pf = (1 - g * g) / (4 * math.pi) / (1 + g * g + 2 * g * cos_theta) ** 1.5
    want = math.exp(-0.6 * (r + math.sqrt(2.0) * r)) * pf * 3 / (math.pi * r * r)
    assert np.allclose(got, want, rtol=1e-5)


def _sq(p=(0, 0, 0), n=(0, 0, 1), wi=(0, 0, 1), refl=(0.5, 0.6, 0.7), flags=0):
    q = np.zeros(1, abi.SURFACE_QUERY_DTYPE)
    q["p"], q["n"], q["wi"], q["reflectance"], q["weight"], q["flags"] = p, n, wi, refl, 1.0, flags
    return q


def test_surface_photon_estimates_closed_form():
    """photonmap.h:131-211 over kdtree.h:319-400: k nearest inside R; the radius of the density estimate shrinks to the
    k-th nearest when more than k lie inside; SmoothDiffuse::eval = reflectance / pi * cos_theta_o on the front side."""
    refl = np.array([0.5, 0.6, 0.7])
    R = 1.0
    # one photon, straight down, 0.3 away: N = 1 <= k, the radius stays R
    one = scenes.photonset_desc([[0.3, 0, 0]], [[0, 0, -1]], [[2.0, 3.0, 4.0]], scale=0.5, lookup_radius=R)
    got = oracle.surface_gather(_sq(), global_desc=one, k_global=10)[0]
    want = np.array([2.0, 3.0, 4.0]) * refl / math.pi * 1.0 * 0.5 / (math.pi * R * R)
    assert np.allclose(got, want, rtol=2e-6)
    # arriving at 60 degrees: cos_theta_o = 0.5
    tilt = scenes.photonset_desc([[0.3, 0, 0]], [[math.sin(math.pi / 3), 0, -0.5]], [[2.0, 3.0, 4.0]], scale=0.5, lookup_radius=R)
    assert np.allclose(oracle.surface_gather(_sq(), global_desc=tilt, k_global=10)[0], 0.5 * want, rtol=2e-6)
    # from behind the surface, or seen from behind a one-sided surface: nothing; twosided flips the frame
    back = scenes.photonset_desc([[0.3, 0, 0]], [[0, 0, 1]], [[2.0, 3.0, 4.0]], scale=0.5, lookup_radius=R)
    assert np.all(oracle.surface_gather(_sq(), global_desc=back, k_global=10) == 0)
    assert np.all(oracle.surface_gather(_sq(wi=(0, 0, -1)), global_desc=back, k_global=10) == 0)
    assert np.allclose(oracle.surface_gather(_sq(wi=(0, 0, -1), flags=abi.SURF_TWOSIDED), global_desc=back, k_global=10)[0], want, rtol=2e-6)
    # outside R: nothing
    far = scenes.photonset_desc([[1.0001, 0, 0]], [[0, 0, -1]], [[2.0, 3.0, 4.0]], scale=0.5, lookup_radius=R)
    assert np.all(oracle.surface_gather(_sq(), global_desc=far, k_global=10) == 0)
    # five photons at 0.1 .. 0.5, k = 3: the three nearest count and the estimate divides by pi * 0.3^2
    d = np.array([0.4, 0.1, 0.5, 0.3, 0.2], np.float32)
    pos = np.stack([d, np.zeros(5), np.zeros(5)], 1)
    pw = np.stack([np.arange(1, 6)] * 3, 1).astype(np.float32)            # power 1..5 in file order
    five = scenes.photonset_desc(pos, [[0, 0, -1]] * 5, pw, scale=1.0, lookup_radius=R)
    got = oracle.surface_gather(_sq(), global_desc=five, k_global=3)[0]
    r2 = float(np.float32(0.3) * np.float32(0.3))
    want3 = (2 + 4 + 5) * refl / math.pi / (math.pi * r2)                 # photons at 0.1, 0.3, 0.2
    assert np.allclose(got, want3, rtol=1e-5)
    # caustic estimate: cone filter (1 - d / r) with the shrunk r, factor 3
    got_c = oracle.surface_gather(_sq(), caustic_desc=five, k_caustic=3)[0]
    wsum = 2 * (1 - 0.1 / 0.3) + 4 * (1 - 0.3 / 0.3) + 5 * (1 - 0.2 / 0.3)
    assert np.allclose(got_c, 3 * wsum * refl / math.pi / (math.pi * r2), rtol=2e-5)
    # both maps: the sum
    both = oracle.surface_gather(_sq(), global_desc=five, k_global=3, caustic_desc=five, k_caustic=3)[0]
    assert np.allclose(both, got + got_c, rtol=1e-6)


def test_delta_bsdf_samples_closed_form():
    """SmoothConductor / SmoothDielectric::sample of the specular chains (vrl.cpp:320, 682): mirror direction
    d - 2 (d.n) n whatever the frame, Snell's law, Fresnel at normal incidence, the radiance scaling eta_ti^2, total
    internal reflection, one-sided conductor."""
    rng = np.random.default_rng(3)
    for _ in range(50):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        d = rng.normal(size=3); d /= np.linalg.norm(d)
        if d @ n > 0:
            d = -d                                                       # arrive on the front side
        wo, wt, tr = oracle.sample_delta(abi.SURF_BSDF_MIRROR, (0.9, 0.8, 0.7), n, d, 0.5)
        assert np.allclose(wo, d - 2 * (d @ n) * n, atol=2e-6) and np.allclose(wt, (0.9, 0.8, 0.7)) and not tr
        # the back of a one-sided conductor reflects nothing; the twosided adapter does
        assert oracle.sample_delta(abi.SURF_BSDF_MIRROR, (0.9, 0.8, 0.7), n, -d, 0.5) is None
        wo2, _, _ = oracle.sample_delta(abi.SURF_BSDF_MIRROR | abi.SURF_TWOSIDED, (0.9, 0.8, 0.7), n, -d, 0.5)
        assert np.allclose(wo2, -d + 2 * (d @ n) * n, atol=2e-6)
        # dielectric, entering (eta = 1.5): refraction obeys Snell, weight = T / eta^2; reflection weight = R
        eta = 1.5
        cos_i = -(d @ n)
        wo_t, wt_t, tr_t = oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, (eta, 0.8, 0.6), n, d, 1.0)
        assert tr_t and np.isclose(np.linalg.norm(wo_t), 1.0, atol=1e-5)
        sin_i, sin_t = math.sqrt(max(0.0, 1 - cos_i ** 2)), np.linalg.norm(np.cross(wo_t, n))
        assert np.isclose(sin_t, sin_i / eta, atol=1e-5) and wo_t @ n < 0
        assert np.allclose(wt_t, 0.6 / eta ** 2, rtol=1e-6)
        tang_in, tang_out = d - (d @ n) * n, wo_t - (wo_t @ n) * n       # same plane of incidence, same side
        assert np.allclose(tang_out, tang_in / eta, atol=1e-5)
        wo_r, wt_r, tr_r = oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, (eta, 0.8, 0.6), n, d, 0.0)
        assert not tr_r and np.allclose(wo_r, d - 2 * (d @ n) * n, atol=2e-6) and np.allclose(wt_r, 0.8)
        # leaving: weight = T * eta^2, or total internal reflection beyond the critical angle
        wo_l, wt_l, tr_l = oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, (eta, 0.8, 0.6), n, -d, 0.999999)
        if sin_i * eta >= 1.0:
            assert not tr_l and np.allclose(wt_l, 0.8)
        else:
            assert tr_l and np.allclose(wt_l, 0.6 * eta ** 2, rtol=1e-6) and wo_l @ n > 0
    # Fresnel reflectance at normal incidence: ((eta - 1) / (eta + 1))^2 = 0.04 for 1.5 -- the switch point of sample1
    n, d = np.array([0, 0, 1.0]), np.array([0, 0, -1.0])
    assert not oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, (1.5, 1, 1), n, d, 0.0399)[2]
    assert oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, (1.5, 1, 1), n, d, 0.0401)[2]
    # index-matched: r_i = 0, straight through with weight T (sample1 == 0 itself still satisfies `sample1 <= r_i`)
    wo, wt, tr = oracle.sample_delta(abi.SURF_BSDF_DIELECTRIC, (1.0, 1, 0.5), n, np.array([0.6, 0, -0.8]), 1e-6)
    assert tr and np.allclose(wo, (0.6, 0, -0.8), atol=1e-6) and np.allclose(wt, 0.5)

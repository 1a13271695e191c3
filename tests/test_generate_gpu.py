"""VRL generation on the device (vrlb_vrls_generate, generate.cu).  The north star replaces the reference's
cell-refraction light tracer by an eikonal marcher, so generation is checked statistically:
free-flight statistics and energy in a homogeneous medium, Snell's invariant along marched paths in a graded-IOR
medium, cosine emission from area emitters, determinism."""
import math

import numpy as np
import pytest

from mitsuba_nlvrl_b200 import abi, api, scenes

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = api.Context(0)
    yield c
    c.close()


def _fog(size, medium):
    sc = scenes.Scene()
    mid = sc.add_medium(medium)
    sc.add_tris(scenes.box_tris((-size,) * 3, (size,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(mid))
    return sc


def test_homogeneous_free_flight_statistics(ctx):
    sigma_t, albedo = 1.0, 0.8
    sc = _fog(200.0, scenes.make_medium(sigma_t=(sigma_t,) * 3, albedo=(albedo,) * 3))
    ctx.set_scene(sc)
    gen = scenes.generate_desc(40000, seed=11, max_depth=100000, rr_depth=5, laser=((0, 0, 0), (0, 0, 1), (2.0, 1.0, 0.5)), start_medium=0)
    n, particles = ctx.generate_vrls(gen)
    # the map takes target_vrls VRLs and no more (VRLMap::push_back, vrl_map.h:43-45); every particle scatters in the unbounded
    # fog, and the one that fills the map is not counted (vrl.cpp:370-371): m_scale = 1 / (particles - 1)
    assert n == 40000 and ctx.vrls_scale() == pytest.approx(1.0 / (particles - 1))
    soa, med = ctx.download_vrls()
    depth = ctx.download_vrl_depths()
    length = soa[6].astype(np.float64)
    # every flight in an unbounded homogeneous medium is Exp(sigma_t): mean 1/sigma_t, variance 1/sigma_t^2
    assert length.mean() == pytest.approx(1 / sigma_t, abs=5 / np.sqrt(n))
    assert length.var() == pytest.approx(1 / sigma_t ** 2, rel=0.08)
    # energy.  A VRL keeps the flux it was created with and is stored at the END of its flight, after that iteration's
    # Russian roulette (vrl.cpp:181-185 before :283): VRL_k is stored with probability S_k = prod_{j<=k, j>=rr_depth} q_j and
    # flux power * t_k, t = throughput at creation.  The reference's set loses (1 - q) of every VRL in flight at a roulette;
    # reproduced, not repaired: E[sum flux] / power = sum_k S_k t_k  (4.52 here, an unbiased tracer would give 5).
    expect, t, surv = 0.0, 1.0, 1.0
    for k in range(400):
        created = t
        if k + 1 > 5:
            q = min(t, 0.95); surv *= q; t /= q
        expect += surv * created
        t *= albedo
    assert 4.4 < expect < 4.6
    assert soa[7].astype(np.float64).sum() / particles == pytest.approx(2.0 * expect, rel=0.04)
    early = depth <= 5
    assert np.allclose(soa[7, early], 2.0 * albedo ** depth[early], rtol=1e-4) and np.allclose(soa[9, early], 0.5 * albedo ** depth[early], rtol=1e-4)
    assert np.allclose(np.linalg.norm(soa[3:6], axis=0), 1.0, atol=1e-5)
    first = depth == 0
    assert first.sum() == particles and np.allclose(soa[3:6, first], np.array([[0], [0], [1]]), atol=1e-6)
    # isotropic phase: scattered directions have zero mean
    assert np.abs(soa[3:6, ~first].mean(axis=1)).max() < 0.02


def test_generation_is_deterministic_and_seed_dependent(ctx):
    sc = _fog(50.0, scenes.make_medium(sigma_t=(0.5,) * 3, albedo=(0.7,) * 3, phase=abi.PHASE_HG, g=0.6))
    ctx.set_scene(sc)
    gen = scenes.generate_desc(5000, seed=3, laser=((0, 0, -40), (0, 0.1, 1), (1, 1, 1)), start_medium=0)
    ctx.generate_vrls(gen)
    a, _ = ctx.download_vrls()
    a = a.copy()
    ctx.generate_vrls(gen)
    b, _ = ctx.download_vrls()
    assert np.array_equal(a, b)
    gen.seed = 4
    ctx.generate_vrls(gen)
    c, _ = ctx.download_vrls()
    assert c.shape != a.shape or not np.array_equal(a, c)
    # HG g = 0.6: mean cosine between consecutive flights of a particle is g; checked through the depth-1 VRLs
    soa, _ = ctx.download_vrls()
    depth = ctx.download_vrl_depths()
    d0 = np.array([0, 0.1, 1.0]) / np.linalg.norm([0, 0.1, 1.0])
    cos1 = (soa[3:6, depth == 1] * d0[:, None]).sum(0)
    assert cos1.mean() == pytest.approx(0.6, abs=0.05)


def test_eikonal_path_keeps_snell_invariant(ctx):
    """In n(y) the horizontal optical momentum n * d_x, n * d_z is conserved and the ray bends towards larger n."""
    lo, hi = (-10.0, -10.0, -10.0), (10.0, 10.0, 10.0)
    med = scenes.make_nonlinear_medium(lo, hi, res_y=2000, bottom_ior=1.0, top_ior=0.8, sigma_t=(1e-5,) * 3, albedo=(0.9,) * 3)
    sc = _fog(10.0, med)
    ctx.set_scene(sc)
    d = np.array([1.0, 0.45, 0.2]); d /= np.linalg.norm(d)
    gen = scenes.generate_desc(40, seed=1, min_vrl_length=0.25, laser=((-9.5, -6.0, 0.0), d, (1, 1, 1)), start_medium=0)
    n, particles = ctx.generate_vrls(gen)
    soa, _ = ctx.download_vrls()
    # the first particle's polyline: consecutive segments until the chain breaks (next particle restarts at the laser)
    o, dd, ln = soa[0:3].T.astype(np.float64), soa[3:6].T.astype(np.float64), soa[6].astype(np.float64)
    k = 1
    while k < n and np.allclose(o[k], o[k - 1] + dd[k - 1] * ln[k - 1], atol=1e-4):
        k += 1
    assert k >= 10, "the curved flight must be cut into several VRLs"
    mid_y = o[:k, 1] + 0.5 * dd[:k, 1] * ln[:k]
    n_mid = 1.0 + (0.8 - 1.0) * (mid_y + 10.0) / 20.0
    inv_x, inv_z = n_mid * dd[:k, 0], n_mid * dd[:k, 2]
    n0 = 1.0 + (0.8 - 1.0) * (-6.0 + 10.0) / 20.0
    assert np.allclose(inv_x, n0 * d[0], rtol=3e-3) and np.allclose(inv_z, n0 * d[2], rtol=3e-3)
    assert np.all(np.diff(dd[:k, 1]) < 0)                  # n decreases upwards: the ray curves back down
    # with use_non_linear = 0 the same particle flies straight: one VRL from the laser to the far boundary
    gen.use_non_linear = 0
    ctx.generate_vrls(gen)
    soa2, _ = ctx.download_vrls()
    assert np.allclose(soa2[3:6, 0], d, atol=1e-6) and soa2[6, 0] > 15


def test_area_emitter_cosine_emission(ctx):
    sc = scenes.Scene()
    mid = sc.add_medium(scenes.make_medium(sigma_t=(0.05,) * 3, albedo=(0.5,) * 3))
    sc.add_tris(scenes.box_tris((-100,) * 3, (100,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(mid))
    e = sc.add_emitter((3.0, 2.0, 1.0))
    quad = np.array([[[-1, -1, 0], [1, -1, 0], [1, 1, 0]], [[-1, -1, 0], [1, 1, 0], [-1, 1, 0]]], np.float32)   # normal +z, area 4
    sc.add_tris(quad, abi.SURF_BSDF_SMOOTH | abi.SURF_EMITTER | abi.surf_emitter_id(e))
    ctx.set_scene(sc)
    gen = scenes.generate_desc(20000, seed=5, start_medium=0)
    n, particles = ctx.generate_vrls(gen)
    soa, _ = ctx.download_vrls()
    depth = ctx.download_vrl_depths()
    f = depth == 0
    assert soa[5, f].mean() == pytest.approx(2.0 / 3.0, abs=0.01)          # E[cos] of a cosine lobe
    assert np.all(np.abs(soa[0, f]) <= 1.001) and np.all(np.abs(soa[1, f]) <= 1.001)
    assert np.allclose(soa[7, f], 3.0 * np.pi * 4.0 * np.pi, rtol=1e-5)     # Le * pi * area * pi (vrl.cpp:151-152)


def test_generated_set_feeds_the_gather(ctx):
    sc = scenes.box_fog_scene(occluders=True)
    ctx.set_scene(sc)
    gen = scenes.generate_desc(3000, seed=2, laser=((0, 9.5, 0), (0.1, -1, 0.05), (50, 50, 50)), start_medium=0)
    n, particles = ctx.generate_vrls(gen)
    m_scale = ctx.vrls_scale()
    cam = scenes.box_fog_camera(32, 18)
    opt = scenes.default_options()
    ctx.build_segments(cam, opt)
    ctx.gather(opt)
    rgb = ctx.film_rgb(32, 18)
    assert np.isfinite(rgb).all() and rgb.max() > 0
    # the same exported set through upload gives the same image (export -> import round trip of the parity protocol)
    soa, med = ctx.download_vrls()
    soa = np.ascontiguousarray(soa)
    ctx.upload_vrls(scenes.vrlset_desc(soa, scale=m_scale))
    ctx.build_segments(cam, opt)
    ctx.gather(opt)
    assert np.allclose(ctx.film_rgb(32, 18), rgb, rtol=1e-5, atol=1e-12)


# ---- the two map transforms of VRLIntegrator::preprocess (vrl.cpp:408-415): long beams and dicing -------------------
def _gold():
    import json, os
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_vrl_struct.json")) as f:
        return json.load(f)


def test_dice_matches_oracle_and_reference_golden():
    """vrlb_vrls_dice against (a) the pieces VRL::dice of the reference produced (tests/golden/ref_vrl_struct.json, made by
    oracle/_ref/refprobe) and (b) the oracle on a 5000-VRL set: count and order identical, origin / direction / length
    bit-exact, flux within 2e-6 (CUDA expf vs Enoki / glibc exp); m_scale untouched."""
    import oracle
    gold = _gold()
    m = gold["medium"]
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=m["sigma_t"], albedo=m["albedo"], scale=m["scale"]))
    ctx = api.Context(0)
    ctx.set_scene(sc)
    src = np.ascontiguousarray(np.array(gold["dice_in"], np.float32).T)
    ctx.upload_vrls(scenes.vrlset_desc(src, scale=0.25))
    n = ctx.dice_vrls(gold["dice_n"])
    want = np.array(gold["dice_out"], np.float64)[:, 1:].T.astype(np.float32)
    got, _ = ctx.download_vrls()
    assert n == want.shape[1] == got.shape[1]
    assert np.array_equal(got[0:7], want[0:7])
    assert np.allclose(got[7:10], want[7:10], rtol=2e-6, atol=0)
    assert ctx.vrls_scale() == 0.25
    # larger set, against the oracle
    sc2 = scenes.box_fog_scene(occluders=True)
    ctx.set_scene(sc2)
    soa = scenes.synthetic_vrls(5000, seed=21)
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    n2 = ctx.dice_vrls(4)
    ref, _ = oracle.OracleScene(sc2).vrls_dice(soa, 4)
    got2, _ = ctx.download_vrls()
    assert n2 == ref.shape[1] and n2 > 4 * 5000
    assert np.array_equal(got2[0:7], ref[0:7])
    assert np.allclose(got2[7:10], ref[7:10], rtol=2e-6, atol=0)
    # energy: the pieces of one VRL carry flux * Tr(start of piece) -- the first piece keeps the flux
    first = np.nonzero(np.all(got2[0:3] == soa[0:3, :1], axis=0))[0][0]
    assert np.array_equal(got2[7:10, first], soa[7:10, 0])
    ctx.close()


def test_to_long_matches_oracle_and_reference_golden():
    """vrlb_vrls_to_long against VRL::convertToLong of the reference (golden) and the oracle; a VRL that sees no surface is
    an error like the reference's Log(Error)."""
    import oracle
    gold = _gold()
    m = gold["medium"]
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=m["sigma_t"], albedo=m["albedo"], scale=m["scale"]))
    sc.add_tris(np.array(gold["room_triangles"], np.float32).reshape(-1, 3, 3), abi.SURF_BSDF_SMOOTH)
    ctx = api.Context(0)
    ctx.set_scene(sc)
    src = np.ascontiguousarray(np.array(gold["dice_in"], np.float32).T)
    ctx.upload_vrls(scenes.vrlset_desc(src))
    ctx.vrls_to_long()
    got, _ = ctx.download_vrls()
    assert np.allclose(got[6], np.array(gold["long"], np.float32), rtol=2e-6, atol=0)
    assert np.array_equal(got[0:6], src[0:6]) and np.array_equal(got[7:10], src[7:10])
    # box-fog: every VRL inside the fog container ends on a surface (container, walls, sphere or cuboid)
    sc2 = scenes.box_fog_scene(occluders=True)
    ctx.set_scene(sc2)
    soa = scenes.synthetic_vrls(4000, seed=22)
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    ctx.vrls_to_long()
    ref, missed = oracle.OracleScene(sc2).vrls_to_long(soa)
    got2, _ = ctx.download_vrls()
    assert missed == 0 and np.array_equal(got2[6], ref)          # same Moeller-Trumbore arithmetic: bit-exact
    assert np.all(got2[6] > 0) and np.all(got2[6] <= 20 * np.sqrt(3) + 1e-3)   # ends on the container, a wall or an occluder
    inside = np.linalg.norm(soa[0:3].T + soa[3:6].T * got2[6][:, None], axis=1, ord=np.inf) <= 10.01 + 1e-3   # the front face is open: the fog container at 10.01 closes it
    assert inside.all()
    # no surface at all -> error
    empty = scenes.Scene()
    empty.add_medium(scenes.make_medium())
    ctx.set_scene(empty)
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    with pytest.raises(api.VrlbError, match="Impossible to convert to long vrl"):
        ctx.vrls_to_long()
    ctx.close()


def test_volume_photons_are_the_first_scatterings_of_the_same_particles(ctx):
    """vrlb_photons_generate traces the same particles as vrlb_vrls_generate: photon k sits at the end of the direct
    (depth-0) VRL of a particle, travels in that VRL's direction and carries the flux of the VRL that starts there
    (power after the albedo, vrl.cpp:270-281).  Capacity and the reference's path count (capacity - 1) are honoured."""
    sigma_t, albedo = 1.0, 0.8
    sc = _fog(200.0, scenes.make_medium(sigma_t=(sigma_t,) * 3, albedo=(albedo,) * 3))
    ctx.set_scene(sc)
    gen = scenes.generate_desc(20000, seed=11, max_depth=100000, rr_depth=5, laser=((0, 0, 0), (0, 0, 1), (2.0, 1.0, 0.5)), start_medium=0)
    ctx.generate_vrls(gen)
    soa, _ = ctx.download_vrls()
    depth = ctx.download_vrl_depths()
    n, particles = ctx.generate_photons(gen, 2000, 0.5)
    assert n == 2000 and particles == 2000                       # every particle scatters in the unbounded fog
    pos, pdir, power, scale = ctx.download_photons()
    assert scale == pytest.approx(1.0 / 1999)
    direct = np.nonzero(depth == 0)[0][:2000]                   # the direct VRLs of particles 0 .. 1999, in particle order
    ends = (soa[0:3, direct] + soa[3:6, direct] * soa[6, direct]).T
    # the resident set is in grid-cell order: match by position
    key = lambda a: np.lexsort((a[:, 2], a[:, 1], a[:, 0]))
    assert np.allclose(pos[key(pos)], ends[key(ends)], atol=2e-5)
    assert np.allclose(pdir, [[0, 0, 1]], atol=1e-6)            # laser direction
    assert np.allclose(power, np.array([[2.0, 1.0, 0.5]]) * albedo, rtol=1e-5)
    # exponential free flight: mean distance of the photons from the laser origin
    assert pos[:, 2].mean() == pytest.approx(1 / sigma_t, abs=5 / np.sqrt(2000))
    # capacity 0 switches the term off (vrl.cpp:79)
    assert ctx.generate_photons(gen, 0, 0.5) == (0, 0) and ctx.photons_count() == 0


def _bounce_box(size, sigma_t, refl, flags=abi.SURF_BSDF_SMOOTH, inward=True, with_table=True):
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=(sigma_t,) * 3, albedo=(0.9,) * 3))
    sc.add_tris(scenes.box_tris((-size,) * 3, (size,) * 3, inward=inward), flags, reflectance=refl if with_table else None)
    return sc


def test_particles_reflect_diffusely_off_smooth_surfaces(ctx):
    """vrl.cpp:307-358 with SmoothDiffuse::sample (diffuse.cpp): a particle that meets a smooth surface closes its VRL there,
    is reflected into a cosine lobe about the normal with weight = reflectance and opens the next VRL at the surface."""
    size, refl, power = 5.0, (0.6, 0.5, 0.4), (2.0, 1.0, 0.5)
    sc = _bounce_box(size, 0.02, refl)                                  # thin fog: most particles reach a wall unscattered
    ctx.set_scene(sc)
    gen = scenes.generate_desc(60000, seed=5, laser=((0, 0, 0), (0.3, 0.2, 1.0), power), start_medium=0)
    n, particles = ctx.generate_vrls(gen)
    soa, _ = ctx.download_vrls()
    depth = ctx.download_vrl_depths()
    o, d, flux = soa[0:3].astype(np.float64), soa[3:6].astype(np.float64), soa[7:10].astype(np.float64)
    end = o + d * soa[6]
    assert np.abs(o).max() <= size + 1e-3 and np.abs(end).max() <= size + 1e-3         # nothing leaves the closed box
    # a reflected VRL starts one ray epsilon (~1e-3 here) off the wall; a few VRLs that start at a scattering happen to lie
    # as close, hence the "almost all" below
    on_wall = np.abs(o).max(axis=0) > size - 2.5e-3
    assert on_wall.sum() > 0.3 * n                                                     # VRLs that start at a reflection
    # the wall a VRL starts on and its inward normal
    ax = np.abs(o).argmax(axis=0)
    normal = np.zeros_like(o)
    normal[ax, np.arange(o.shape[1])] = -np.sign(o[ax, np.arange(o.shape[1])])
    cos_o = (d * normal).sum(0)[on_wall]
    assert (cos_o > 0).mean() > 0.995                                                  # reflected to the side it came from
    assert cos_o.mean() == pytest.approx(2.0 / 3.0, abs=4 / np.sqrt(on_wall.sum()))   # cosine lobe: E[cos] = 2/3
    assert (cos_o ** 2).mean() == pytest.approx(0.5, abs=4 / np.sqrt(on_wall.sum()))  # E[cos^2] = 1/2
    # laser -> wall -> first reflected VRL (depth 1, starts on a wall): flux = power * reflectance exactly
    first_bounce = on_wall & (depth == 1)
    assert first_bounce.sum() > 1000
    assert np.isclose(flux[:, first_bounce], (np.array(power) * np.array(refl))[:, None], rtol=1e-5).all(axis=0).mean() > 0.995
    # k events before the roulette depth, r of them reflections: flux = power * albedo^(k - r) * reflectance^r; check the
    # red channel's admissible set for depth 2
    d2 = flux[0, depth == 2] / power[0]
    allowed = np.array([0.9 ** 2, 0.9 * 0.6, 0.6 ** 2])
    assert np.abs(d2[:, None] - allowed[None, :]).min(axis=1).max() < 1e-5
    # without the reflectance table the particles end at the walls (ABI 2 behaviour): no VRL starts on a wall
    ctx.set_scene(_bounce_box(size, 0.02, refl, with_table=False))
    ctx.generate_vrls(gen)
    soa0, _ = ctx.download_vrls()
    assert (np.abs(soa0[0:3]).max(axis=0) > size - 2.5e-3).sum() < 0.005 * soa0.shape[1]


def test_one_sided_diffuse_surfaces_and_the_twosided_adapter(ctx):
    """diffuse.cpp: nothing is reflected from the back of a one-sided surface (cos_theta_i <= 0 -> weight 0, the particle
    ends); `twosided` (VRLB_SURF_TWOSIDED) reflects on both sides."""
    size, refl = 5.0, (0.6, 0.6, 0.6)
    gen = scenes.generate_desc(20000, seed=6, laser=((0, 0, 0), (0.3, 0.2, 1.0), (1, 1, 1)), start_medium=0)
    ctx.set_scene(_bounce_box(size, 0.02, refl, inward=False))          # outward normals: the particles see the backs
    ctx.generate_vrls(gen)
    back, _ = ctx.download_vrls()
    assert (np.abs(back[0:3]).max(axis=0) > size - 2.5e-3).sum() < 0.005 * back.shape[1]
    ctx.set_scene(_bounce_box(size, 0.02, refl, flags=abi.SURF_BSDF_SMOOTH | abi.SURF_TWOSIDED, inward=False))
    ctx.generate_vrls(gen)
    two, _ = ctx.download_vrls()
    on_wall = np.abs(two[0:3]).max(axis=0) > size - 2.5e-3
    assert on_wall.sum() > 0.3 * two.shape[1]
    ax = np.abs(two[0:3]).argmax(axis=0)
    idx = np.arange(two.shape[1])
    assert ((two[3:6][ax, idx][on_wall] * -np.sign(two[0:3][ax, idx][on_wall])) > 0).mean() > 0.995     # back into the box


def test_particles_end_on_emitter_surfaces_without_storing_the_flight(ctx):
    """vrl.cpp:309-310: `if (si.shape->is_emitter()) break;` comes before the VRL in flight is stored."""
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=(1e-4,) * 3, albedo=(0.9,) * 3))       # practically no scattering
    e = sc.add_emitter((1.0, 1.0, 1.0))
    sc.add_tris(scenes.box_tris((-5,) * 3, (5,) * 3, inward=True), abi.SURF_BSDF_SMOOTH | abi.SURF_EMITTER | (e << 24), reflectance=0.5)
    ctx.set_scene(sc)
    gen = scenes.generate_desc(60, seed=1, laser=((0, 0, 0), (0, 0, 1), (1, 1, 1)), start_medium=0)
    n, particles = ctx.generate_vrls(gen)
    assert n >= 60 and particles > 1000 * n                                         # only the rare scattered particles store anything
    soa, _ = ctx.download_vrls()
    assert np.abs((soa[0:3] + soa[3:6] * soa[6])).max() < 5 - 1e-3                  # every stored VRL ends inside, at a scattering


def test_global_photons_are_the_surface_hits_of_the_same_particles(ctx):
    """handleSurfaceInteraction (vrl.cpp:315, 772-782): every smooth surface a particle meets stores a photon (position,
    travel direction, flux * throughput before the reflection) until the map is full; scale = 1 / surfacePathCount where
    every particle traced while the map was not full counts, except the one that fills it (vrl.cpp:124, 368-369, 379)."""
    size, refl, power = 5.0, (0.6, 0.5, 0.4), (2.0, 1.0, 0.5)
    ctx.set_scene(_bounce_box(size, 0.02, refl))
    gen = scenes.generate_desc(60000, seed=5, laser=((0, 0, 0), (0.3, 0.2, 1.0), power), start_medium=0)
    n, particles = ctx.generate_surface_photons(gen, 5000, 0.4, 50)
    assert n == 5000 and 500 < particles < 5000                           # several bounces per particle
    assert ctx.surface_photons_count(abi.PHOTONMAP_GLOBAL) == 5000
    pos, pdir, pw, scale = ctx.download_surface_photons(abi.PHOTONMAP_GLOBAL)
    assert scale == pytest.approx(1.0 / (particles - 1))
    assert np.allclose(np.abs(pos).max(axis=1), size, atol=1e-3)          # on the walls
    ax = np.abs(pos).argmax(axis=1)
    idx = np.arange(len(pos))
    assert (pdir[idx, ax] * np.sign(pos[idx, ax]) > 0).all()              # travelling into the wall
    # the unscattered laser flight: full power, all at the spot the laser hits on the z = +size wall
    direct = np.isclose(pw, np.array(power, np.float32)[None, :], rtol=1e-6).all(axis=1)
    assert direct.sum() > 0.5 * particles
    assert np.allclose(pos[direct], [[1.5, 1.0, 5.0]], atol=1e-4)
    # below the roulette depth: power = laser power * albedo^a * reflectance^r (red channel, up to two events)
    ratio = pw[:, 0].astype(np.float64) / power[0]
    allowed = np.array([1.0, 0.9, 0.6, 0.81, 0.54, 0.36])
    assert (np.abs(ratio[:, None] - allowed[None, :]).min(axis=1) < 1e-5).mean() > 0.5
    # no reflectance table or capacity 0: no map
    assert ctx.generate_surface_photons(gen, 0, 0.4, 50) == (0, 0) and ctx.surface_photons_count(abi.PHOTONMAP_GLOBAL) == 0
    ctx.set_scene(_bounce_box(size, 0.02, refl, with_table=False))
    assert ctx.generate_surface_photons(gen, 5000, 0.4, 50)[0] == 0
    # the volume map counts its paths the same way: here most particles end on a wall without scattering, and count
    n_v, p_v = ctx.generate_photons(gen, 500, 0.5)
    assert n_v == 500 and p_v > 3 * n_v
    assert ctx.download_photons()[3] == pytest.approx(1.0 / (p_v - 1))


def test_light_side_specular_chain_and_caustic_photons(ctx):
    """vrl.cpp:318-358 with SmoothDielectric::sample: a particle is refracted through a glass slab (Snell twice: the spot on
    the far wall shifts by the slab's lateral offset), keeps its power (eta^-2 in, eta^2 out), and the photon it stores on
    the next diffuse surface goes to the caustic map (wasTransmitted, vrl.cpp:345, 776-780); the 4 % reflected at the slab
    end up in the global map."""
    size, refl, power, eta = 5.0, (0.6, 0.5, 0.4), (2.0, 1.0, 0.5), 1.5
    sc = scenes.Scene()
    sc.add_medium(scenes.make_medium(sigma_t=(0.002,) * 3, albedo=(0.9,) * 3))
    sc.add_tris(scenes.box_tris((-size,) * 3, (size,) * 3, inward=True), abi.SURF_BSDF_SMOOTH, reflectance=refl)
    sc.add_tris(scenes.box_tris((-4.9, -4.9, 2.0), (4.9, 4.9, 3.0)), abi.SURF_BSDF_DIELECTRIC, reflectance=(eta, 1.0, 1.0))
    ctx.set_scene(sc)
    gen = scenes.generate_desc(20000, seed=8, laser=((0, 0, 0), (0.3, 0.0, 1.0), power), start_medium=0)
    n_c, p_c = ctx.generate_surface_photons(gen, 4000, 0.4, 50, which=abi.PHOTONMAP_CAUSTIC)
    assert n_c == 4000
    pos, pdir, pw, scale = ctx.download_surface_photons(abi.PHOTONMAP_CAUSTIC)
    assert scale == pytest.approx(1.0 / (p_c - 1))
    direct = np.isclose(pw, np.array(power, np.float32)[None, :], rtol=2e-5).all(axis=1)
    assert direct.sum() > 0.4 * 4000                                     # straight through the slab onto the z = +size wall (the
                                                                         # rest crossed the slab again after a diffuse bounce)
    sin_i = 0.3 / math.sqrt(1.09)
    tan_t = (sin_i / eta) / math.sqrt(1 - (sin_i / eta) ** 2)
    x_spot = 0.3 * 2.0 + tan_t * 1.0 + 0.3 * 2.0
    # (the 4 % reflected inside the slab leave through its lower face with their full power: caustic photons on z = -size)
    assert (np.abs(pos[direct] - np.array([[x_spot, 0.0, size]])).max(axis=1) < 2e-3).mean() > 0.9
    assert (np.abs(pdir[direct] - np.array([[0.3, 0.0, 1.0]]) / math.sqrt(1.09)).max(axis=1) < 1e-5).mean() > 0.9   # the exit restores the direction
    # the global map: photons whose last surface event was not a transmission -- the slab's reflection lands on z = -size
    n_g, p_g = ctx.generate_surface_photons(gen, 4000, 0.4, 50, which=abi.PHOTONMAP_GLOBAL)
    gpos, gdir, gpw, _ = ctx.download_surface_photons(abi.PHOTONMAP_GLOBAL)
    mirrored = np.isclose(gpw, np.array(power, np.float32)[None, :], rtol=2e-5).all(axis=1)
    assert mirrored.sum() > 10 and np.allclose(gpos[mirrored][:, 2], -size, atol=1e-3)
    assert mirrored.sum() < 0.2 * max(p_g, 1)                            # Fresnel: ~4 % at this angle (twice: both faces)
    # no caustic photon is also a global one: full-power photons on the far wall are all in the caustic map
    assert not (mirrored & (gpos[:, 2] > size - 1e-3)).any()
    # VRLs: the flight continues behind the slab with the entry direction
    ctx.generate_vrls(gen)
    soa, _ = ctx.download_vrls()
    behind = (soa[2] > 3.0) & np.isclose(soa[7], power[0], rtol=2e-5)
    assert behind.sum() > 1000
    assert (np.abs(soa[3:6, behind].T - np.array([[0.3, 0.0, 1.0]]) / math.sqrt(1.09)).max(axis=1) < 1e-5).mean() > 0.9


def test_grid_medium_light_tracer_takes_null_collisions(ctx):
    """Delta tracking in the light tracer (vrl.cpp:246-264): free flights are sampled with the majorant and a tentative
    collision is real with probability sigma_t(x) / majorant.  A laser along +z through a grid whose density steps from
    rho1 (z < 0) to rho2 (z > 0): the first scatterings (the volume photons) follow sigma(z) exp(-int sigma) -- not the
    majorant's exponential -- and a VRL runs through null collisions uncut."""
    L, rho1, rho2, scale = 10.0, 0.25, 1.0, 0.4
    nz = 80
    grid = np.empty((nz, 4, 4), np.float32)
    grid[: nz // 2] = rho1
    grid[nz // 2:] = rho2
    m = scenes.make_medium(kind=abi.MEDIUM_GRID, sigma_t=(1, 1, 1), albedo=(0.8,) * 3, scale=scale)
    m.grid_bbox_min[:] = [-L] * 3
    m.grid_bbox_max[:] = [L] * 3
    sc = scenes.Scene()
    mid = sc.add_medium(m, grid=grid)
    sc.add_tris(scenes.box_tris((-L,) * 3, (L,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(mid))
    ctx.set_scene(sc)
    gen = scenes.generate_desc(1, seed=5, max_depth=100000, rr_depth=50, laser=((0.1, 0.2, -L + 1e-3), (0, 0, 1), (1, 1, 1)), start_medium=0)
    n, particles = ctx.generate_photons(gen, 20000, 0.5)
    pos, pdir, power, pscale = ctx.download_photons()
    assert n == 20000
    s1, s2 = scale * rho1, scale * rho2
    p_first = 1 - np.exp(-s1 * L)                                       # first scattering before the step
    p_second = np.exp(-s1 * L) * (1 - np.exp(-s2 * L))
    frac_first = (pos[:, 2] < 0).mean()
    want = p_first / (p_first + p_second)                               # (particles that cross the box unscattered leave no photon)
    assert abs(frac_first - want) <= 4 * np.sqrt(want * (1 - want) / n) + 0.01, (frac_first, want)
    # inside each half the depths are exponential with that half's extinction: mean depth of a truncated exponential
    mean_trunc = lambda s, ln: 1 / s - ln * np.exp(-s * ln) / (1 - np.exp(-s * ln))
    z1 = pos[pos[:, 2] < -0.3, 2] + L                                   # (0.3 = more than a voxel away from the trilinear ramp at the step)
    z2 = pos[pos[:, 2] > 0.3, 2] - 0.3
    assert z1.mean() == pytest.approx(mean_trunc(s1, L - 0.3), rel=0.04) and z2.mean() == pytest.approx(mean_trunc(s2, L - 0.3), rel=0.04)
    assert np.allclose(power, 0.8, rtol=1e-5)                            # a null collision leaves the throughput alone
    # with the majorant's exponential (the behaviour before delta tracking) the first half would hold
    wrong = (1 - np.exp(-s2 * L))
    assert abs(frac_first - wrong) > 0.2

"""GPU suite: a converged-image test in the manner of the reference's render tests (src/librender/tests/test_renders.py:64-81:
mean image against a reference with a Z-test on the per-pixel variance) -- here against a CLOSED FORM, so that it checks what
no same-set parity test can: VRL generation (laser, free-flight lengths, flux, 1 / path-count scale), the unbiasedness of the
2-D importance sampling, the film accumulation, end to end through vrlb_vrls_generate -> vrlb_segments_build -> vrlb_gather.

Scene: the box-fog without occluders, a laser inside the fog, max_depth = 3 so that the light tracer stores exactly the first
VRL of every particle (laser origin -> first collision or the wall; the second VRL is dropped at vrl.cpp:257-258).  The
expected radiance of a camera segment is then a smooth 2-D integral

    Li = n * scale * power * sigma_s^2 / (4 pi)^2 * Int_0^T dt Int_0^S ds  e^{-mu t} e^{-2 mu s} e^{-mu d(s,t)} / d(s,t)^2

(e^{-mu s} once for the probability that the VRL reaches s and once for the transmittance VRL::getContrib applies along the
VRL, vrl_struct.h:647 -- the reference's short-VRL estimator, reproduced), evaluated here by a midpoint rule."""
import numpy as np
import pytest

from mitsuba_nlvrl_b200 import api, scenes

pytestmark = pytest.mark.gpu

ORIGIN, DIRECTION, POWER = np.array([-9.0, 1.0, -2.0]), np.array([1.0, 0.1, 0.15]), np.array([3.0, 2.0, 1.0])
SIGMA_S, SIGMA_A = 0.1, 0.01


def _expected(seg, beam_len, n_scale, m=1200):
    mu = SIGMA_S + SIGMA_A
    d0 = DIRECTION / np.linalg.norm(DIRECTION)
    t = (np.arange(m) + 0.5) * (float(seg["maxt"]) / m)
    s = (np.arange(m) + 0.5) * (beam_len / m)
    pc = seg["o"].astype(np.float64)[None, :] + t[:, None] * seg["d"].astype(np.float64)[None, :]
    pb = ORIGIN[None, :] + s[:, None] * d0[None, :]
    d2 = ((pc[:, None, :] - pb[None, :, :]) ** 2).sum(axis=2)
    d = np.sqrt(d2)
    f = np.exp(-mu * t)[:, None] * np.exp(-2 * mu * s)[None, :] * np.exp(-mu * d) / d2
    integral = f.sum() * (float(seg["maxt"]) / m) * (beam_len / m)
    return n_scale * SIGMA_S ** 2 / (4 * np.pi) ** 2 * integral * seg["weight"].astype(np.float64), float(d.min())


def test_laser_single_scattering_image_converges_to_the_closed_form():
    ctx = api.Context(0)
    sc, cam = scenes.box_fog_scene(occluders=False, sigma_s=SIGMA_S, sigma_a=SIGMA_A), scenes.box_fog_camera(32, 18)
    ctx.set_scene(sc)
    opt = scenes.default_options(visibility=1, math_mode=1)
    n, K = 2000, 24
    films, survived, lens, beam_len = [], [], [], 0.0
    d0 = DIRECTION / np.linalg.norm(DIRECTION)
    mu = SIGMA_S + SIGMA_A
    for k in range(K):
        gen = scenes.generate_desc(n, seed=1000 + 100003 * k, max_depth=3,      # particle i of a set draws from seed + i: disjoint ranges
                                   rr_depth=100, use_non_linear=0,
                                   laser=(tuple(ORIGIN), tuple(DIRECTION), tuple(POWER)), start_medium=0)
        n_vrls, particles = ctx.generate_vrls(gen)
        assert n_vrls == n, (n_vrls, particles)
        scale = ctx.vrls_scale()
        assert abs(scale * (particles - 1) - 1.0) < 1e-6 or abs(scale * particles - 1.0) < 1e-6       # the reference's path count (vrl.cpp:404, 423)
        # max_depth = 3 also keeps the second VRL of a particle when it ends on a wall; the closed form is about the first
        soa, _ = ctx.download_vrls()
        first = np.all(np.abs(soa[0:3].T - ORIGIN) < 1e-5, axis=1)
        assert particles - 1 <= first.sum() <= particles
        assert np.allclose(soa[3:6, first & (soa[6] > 0.5)].T, d0, atol=1e-4) and np.allclose(soa[7:10, first].T, POWER, rtol=1e-6)
        sub = np.ascontiguousarray(soa[:, first])
        beam_len = max(beam_len, float(sub[6].max()))
        survived.append(float((sub[6] >= sub[6].max() - 1e-2).mean()))
        lens.append(float(sub[6].mean()))
        ctx.upload_vrls(scenes.vrlset_desc(sub, scale=scale))
        opt.base_seed = 77 + k
        ctx.build_segments(cam, opt)
        ctx.clear_film(); ctx.gather(opt)
        films.append(ctx.film_rgb(cam.width, cam.height).reshape(-1, 3).astype(np.float64) / (first.sum() * scale))
    assert abs((ORIGIN + beam_len * d0)[0] - 10.0) < 1e-3                    # the VRLs that are not cut by a collision end on the wall x = +10
    assert abs(np.mean(survived) - np.exp(-mu * beam_len)) < 0.01           # free-flight survival over K * n particles
    n_scale = 1.0
    assert abs(np.mean(lens) * mu / (1 - np.exp(-mu * beam_len)) - 1.0) < 0.02          # mean of the truncated exponential
    segs = ctx.download_segments()
    films = np.array(films)
    mean, se = films.mean(axis=0), films.std(axis=0, ddof=1) / np.sqrt(K)
    rng = np.random.default_rng(5)
    z, ratio, rse = [], [], []
    for i in rng.permutation(len(segs)):
        sg = segs[i]
        want, d_min = _expected(sg, beam_len, n_scale)
        if d_min < 0.75:                     # the camera ray (nearly) crosses the beam: log-singular integrand, heavy-tailed estimator
            continue
        want = want * POWER
        px = int(sg["pixel"])
        z.append((mean[px] - want) / se[px])
        ratio.append(mean[px].sum() / want.sum())
        rse.append((se[px] / mean[px]).mean())
        if len(z) == 40:
            break
    z, ratio = np.array(z), np.array(ratio)
    assert len(z) == 40
    print(f"40 pixels x {K} frames x {n} VRLs: max |Z| {np.abs(z).max():.2f}, mean Z {z.mean():+.2f}, "
          f"mean ratio GPU / closed form {np.mean(ratio):.4f} (relative standard error per pixel {np.median(rse):.3f})")
    assert np.abs(z).max() <= 4.5
    assert abs(z.mean()) <= 0.6
    assert abs(np.mean(ratio) - 1.0) <= 0.015
    ctx.close()


def test_gpu_frame_against_independent_oracle_frames():
    """The recipe of test_renders.py:64-81 with the CPU oracle as the reference renderer: K independent oracle frames (own
    sampler seeds, so own jitter and own sampling draws) give a per-pixel mean and variance; one GPU frame rendered with
    yet another seed must be a draw from that distribution (Z-test per pixel, Bonferroni-sized bound), and its mean over
    the image must agree to a fraction of a percent.  Same exported VRL set on both sides; nothing else is shared -- so
    this catches what a same-stream comparison cannot see (a mistake both sides make at a fixed stream position)."""
    import oracle
    sc, cam = scenes.box_fog_scene(occluders=True), scenes.box_fog_camera(48, 27)
    soa = scenes.synthetic_vrls(256, seed=8)
    vd = scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    K, spp = 16, 4
    frames = []
    for k in range(K):
        opt = scenes.default_options(visibility=1, math_mode=0, spp=spp, base_seed=1000 + 7919 * k)
        xyzaw, rgb, _ = osc.render(cam, opt, vd)
        frames.append((xyzaw[..., :3] / np.maximum(xyzaw[..., 4:5], 1e-30)).reshape(-1, 3).astype(np.float64))     # XYZ / W, what the film develops
    frames = np.array(frames)
    mean, var = frames.mean(axis=0), frames.var(axis=0, ddof=1)                # var of ONE spp-sample frame
    ctx = api.Context(0)
    ctx.set_scene(sc)
    gopt = scenes.default_options(visibility=1, math_mode=1, spp=spp, base_seed=424242)
    out = ctx.render(cam, gopt, vd)
    ctx.close()
    gpu = (out[..., :3] / np.maximum(out[..., 4:5], 1e-30)).reshape(-1, 3).astype(np.float64)
    lit = mean.max(axis=1) > 0.05 * mean.max()
    z = (gpu - mean)[lit] / np.sqrt(var[lit] * (1.0 + 1.0 / K) + 1e-30)       # Student t, K - 1 degrees of freedom
    print(f"{lit.sum()} pixels, K = {K} oracle frames at {spp} spp: max |Z| {np.abs(z).max():.2f}, |Z| > 3 in {100 * (np.abs(z) > 3).mean():.2f} %, "
          f"mean Z {z.mean():+.3f}, image mean GPU / oracle {gpu[lit].sum() / mean[lit].sum():.4f}")
    assert np.abs(z).max() <= 8.0 and (np.abs(z) > 3).mean() <= 0.03 and abs(z.mean()) <= 0.25
    assert abs(gpu[lit].sum() / mean[lit].sum() - 1.0) <= 0.01

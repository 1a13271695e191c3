"""Shared pieces of the GPU parity tests: relative error with the north-star floor, and the EXPLANATION of a camera segment
whose GPU / oracle sums differ by more than the tolerance -- its evals are re-run one VRL at a time on both sides (global
stream addressing keeps every eval on its own draws) and the difference has to sit in a handful of evals that took a
different hit / accept decision (one side exactly 0, or a discrete jump), with everything else agreeing."""
import numpy as np

import mitsuba_nlvrl_b200 as pkg

scenes = pkg.scenes
REL_TOL = 1e-4


def seg_key(s):
    return np.lexsort((s["rng_state"], s["part"], s["sample"], s["pixel"]))


def rel_err(got, ref):
    floor = 1e-6 * np.abs(ref).max()
    return (np.abs(got - ref) / np.maximum(np.abs(ref), floor)).max(axis=-1)


def explain(ctx, osc, seg, soa, scale, opt, max_flips=6):
    """Re-runs one camera segment one VRL at a time on the GPU and on the oracle (global stream addressing keeps every eval
    on its own draws).  Returns (#evals that differ by a decision, relative error of the sum over all OTHER evals)."""
    n = soa.shape[1]
    one = np.ascontiguousarray(seg.reshape(1))
    ctx.upload_segments(one, 2048, 2048)                       # (pixel index < width * height is all the upload checks)
    g = np.zeros((n, 3)); o = np.zeros((n, 3))
    for j in range(n):
        vd = scenes.vrlset_desc(np.ascontiguousarray(soa[:, j:j + 1]), scale=scale, first_index=j, total=n)
        ctx.upload_vrls(vd)
        g[j] = ctx.gather_segments(opt)[0]
        o[j] = osc.gather(vd, one, opt, n_threads=1)[1][0]
    mag = np.abs(o).sum(axis=1) + np.abs(g).sum(axis=1)
    diff = np.abs(g - o).sum(axis=1)
    flip = diff > 1e-3 * np.maximum(mag, 1e-30)              # one side 0 (shadow decision) or a discrete jump (accept / reject)
    rest = rel_err(g[~flip].sum(axis=0)[None], o[~flip].sum(axis=0)[None])[0]
    return int(flip.sum()), float(rest)


def check_segments(ctx, osc, segs, idx, li, soa, scale, opt, label, max_explain=6, max_bad=None):
    vd = scenes.vrlset_desc(soa, scale=scale)
    o32, o64, ev = osc.gather(vd, segs[idx], opt)
    err = rel_err(li[idx], o64)
    bad = np.nonzero(err > REL_TOL)[0]
    print(f"{label}: {len(idx)} segments, {ev:.3g} evals, max rel err {err.max():.2e}, {len(bad)} above 1e-4, median {np.median(err):.1e}")
    assert len(bad) <= (max_bad if max_bad is not None else max(8, len(idx) // 2000)), (label, len(bad))
    for k in bad[:max_explain]:
        flips, rest = explain(ctx, osc, segs[idx[k]], soa, scale, opt)
        print(f"  segment {idx[k]} (pixel {segs['pixel'][idx[k]]}): rel err {err[k]:.2e} = {flips} decision flips; all other evals agree to {rest:.1e}")
        assert 1 <= flips <= 12 and rest <= REL_TOL, (label, idx[k], flips, rest)
    return err




def check_pixels(ctx, osc, segs, idx, li, soa, scale, opt, label, max_explain=4, max_bad=8):
    """Curved camera paths: the north star's tolerance is per PIXEL -- the sum over a path's parts of weight * Li.  (A single
    part can be a stub of 1e-5 of its path; the equi-angular pdf of such a part, theta_B - theta_A of two nearly equal
    arctangents, carries 1e-3 of relative rounding noise in ANY fp32 evaluation, the reference's included, at 1e-8 of the
    pixel.)  `idx` must hold complete paths.  An offending pixel is explained through its worst part."""
    vd = scenes.vrlset_desc(soa, scale=scale)
    o32, o64, ev = osc.gather(vd, segs[idx], opt)
    pix, inv = np.unique(segs["pixel"][idx], return_inverse=True)
    got, ref = np.zeros((len(pix), 3)), np.zeros((len(pix), 3))
    w = segs["weight"][idx].astype(np.float64)
    np.add.at(got, inv, li[idx] * w)
    np.add.at(ref, inv, o64 * w)
    err = rel_err(got, ref)
    bad = np.nonzero(err > REL_TOL)[0]
    print(f"{label}: {len(pix)} pixels / {len(idx)} parts, {ev:.3g} evals, max rel err per pixel {err.max():.2e}, {len(bad)} above 1e-4, median {np.median(err):.1e}")
    assert len(bad) <= max_bad, (label, len(bad))
    for k in bad[:max_explain]:
        members = np.nonzero(inv == k)[0]
        worst = members[np.argmax(np.abs((li[idx][members] - o64[members]) * w[members]).sum(axis=1))]
        flips, rest = explain(ctx, osc, segs[idx[worst]], soa, scale, opt)
        print(f"  pixel {pix[k]}: rel err {err[k]:.2e}; its worst part has {flips} decision flips, all its other evals agree to {rest:.1e}")
        assert 1 <= flips <= 12 and rest <= REL_TOL, (label, pix[k], flips, rest)
    return err

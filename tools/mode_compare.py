"""GPU: per-segment error of math_mode 0/1/2 against the CPU oracle, and gather time per mode (dev tool)."""
import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import oracle
from mitsuba_nlvrl_b200 import abi, api, scenes

def rel_err(got, ref):
    floor = 1e-6 * np.abs(ref).max()
    return np.abs(got - ref) / np.maximum(np.abs(ref), floor)

CFG = [(0, 0, 0), (1, 0, 0), (1, 50, 32), (1, 200, 8), (1, 400, 4), (2, 0, 0)]
ctx = api.Context(0)
for occl in (False, True):
    for phase in (abi.PHASE_ISOTROPIC, abi.PHASE_HG):
        sc = scenes.box_fog_scene(occluders=occl, phase=phase)
        cam = scenes.box_fog_camera(96, 54)
        soa = scenes.synthetic_vrls(int(os.environ.get("NV", 2000)), seed=3)
        vd = scenes.vrlset_desc(soa)
        osc = oracle.OracleScene(sc)
        ctx.set_scene(sc); ctx.upload_vrls(vd)
        ref = None
        for mode, gh, gz in CFG:
            opt = scenes.default_options(visibility=1 if occl else 0, math_mode=mode)
            opt.reserved[0], opt.reserved[1] = gh, gz
            ctx.build_segments(cam, opt)
            segs = ctx.download_segments()
            li = ctx.gather_segments(opt).astype(np.float64)
            _, ref, _ = osc.gather(vd, segs, scenes.default_options(visibility=1 if occl else 0))   # segment order varies per build
            e = rel_err(li, ref).max(axis=1)
            print(f"occl={occl} phase={phase} mode={mode} gh={gh} gz={gz}: max {e.max():.2e} p99.9 {np.quantile(e, 0.999):.2e} p99 {np.quantile(e, 0.99):.2e} "
                  f"median {np.median(e):.2e} n>1e-4 {(e > 1e-4).sum()}/{len(e)}", flush=True)
# timing at 1080p
cam = scenes.box_fog_camera(1920, 1080)
soa = scenes.synthetic_vrls(2000, seed=1)
for occl in (False, True):
    sc = scenes.box_fog_scene(occluders=occl)
    ctx.set_scene(sc); ctx.upload_vrls(scenes.vrlset_desc(soa))
    for mode, gh, gz in CFG:
        opt = scenes.default_options(visibility=1 if occl else 0, math_mode=mode)
        opt.reserved[0], opt.reserved[1] = gh, gz
        n = ctx.build_segments(cam, opt)
        for _ in range(2):
            ctx.gather(opt); ctx.synchronize()
        t = ctx.timing()
        print(f"occl={occl} mode={mode} gh={gh} gz={gz}: gather {t['ms_gather']:.2f} ms  {t['n_evals'] / t['ms_gather'] * 1e-6:.2f} G evals/s", flush=True)

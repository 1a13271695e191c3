"""Diagnostic: whole-set gather vs the sum of two VRL shards on the bench frame (which pixels differ, by how much)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mitsuba_nlvrl_b200 import api, scenes
if os.environ.get('VRLB_LIB'): api.LIB_PATH = os.environ['VRLB_LIB']   # A/B against another build (diagnostic only)

def rel(a, b):
    floor = 1e-6 * np.abs(b).max()
    return np.abs(a - b) / np.maximum(np.abs(b), floor)

ctx = api.Context(0)
sc, cam = scenes.box_fog_scene(occluders=True), scenes.box_fog_camera(1920, 1080)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
soa = scenes.synthetic_vrls(n, seed=1)
def run(mode, n, splits, vis=1):
    soa = scenes.synthetic_vrls(n, seed=1)
    opt = scenes.default_options(visibility=vis, math_mode=mode)
    ctx.set_scene(sc if vis else scenes.box_fog_scene(occluders=False)); ctx.upload_vrls(scenes.vrlset_desc(soa)); ctx.build_segments(cam, opt)
    segs = ctx.download_segments()
    w1 = ctx.gather_segments(opt).astype(np.float64)
    for split in splits:
        parts = np.zeros_like(w1)
        for lo, hi in ((0, split), (split, n)):
            ctx.upload_vrls(scenes.vrlset_desc(np.ascontiguousarray(soa[:, lo:hi]), scale=1.0 / n, first_index=lo, total=n))
            parts += ctx.gather_segments(opt)
        e = rel(parts, w1)
        i = np.unravel_index(np.argmax(e), e.shape)
        px = segs["pixel"][i[0]]
        print(f"mode {mode} vis {vis} n {n} split {split}: max {e.max():.3e} at pixel ({px % 1920},{px // 1920}) whole {w1[i]:.9e} parts {parts[i]:.9e} "
              f"diff/avg-eval {(parts[i]-w1[i])/(w1[i]/n):.3f}  n>1e-5: {(e > 1e-5).sum()}  n>3e-6: {(e > 3e-6).sum()}", flush=True)

run(1, 100000, (37777, 12345, 77777))

"""Where does the visibility kernel's time go?  Same 1080p frame, 5000 VRLs, four variants."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from mitsuba_nlvrl_b200 import api, scenes

n = 5000
soa = scenes.synthetic_vrls(n, seed=1)
cam = scenes.box_fog_camera(1920, 1080)
ctx = api.Context(0)
for name, occ, vis, mode in (("no occluders, vis off (V1 kernel)", False, 0, 1), ("no occluders, vis off, unguarded", False, 0, 2),
                             ("no occluders, vis on (shells only)", False, 1, 1), ("no occluders, vis on, unguarded", False, 1, 2),
                             ("occluders, vis on, guard (bench)", True, 1, 1), ("occluders, vis on, unguarded", True, 1, 2),
                             ("occluders, vis on, exact kernel", True, 1, 0)):
    sc = scenes.box_fog_scene(occluders=occ)
    opt = scenes.default_options(visibility=vis, math_mode=mode)
    ctx.set_scene(sc)
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    nseg = ctx.build_segments(cam, opt)
    ms = []
    for _ in range(2):
        ctx.gather(opt); ctx.synchronize(); ms.append(ctx.timing()["ms_gather"])
    print(f"{name:40s} {ms[-1]:9.1f} ms  {nseg * n / ms[-1] * 1e-6:8.2f} G evals/s", flush=True)

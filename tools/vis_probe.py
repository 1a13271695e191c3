"""Quick look at the visibility kernel: the bench frame (1080p box-fog V2) at a reduced VRL count, one launch each."""
import sys

sys.path.insert(0, ".")
from mitsuba_nlvrl_b200 import api, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
soa = scenes.synthetic_vrls(n, seed=1)
cam = scenes.box_fog_camera(1920, 1080)
ctx = api.Context(0)
for name, occ, vis, mode in (("no occluders, vis off (V1 kernel)", False, 0, 1), ("no occluders, vis on (shells only)", False, 1, 1),
                             ("occluders, vis on, guard (bench)", True, 1, 1), ("occluders, vis on, unguarded", True, 1, 2)):
    sc = scenes.box_fog_scene(occluders=occ)
    opt = scenes.default_options(visibility=vis, math_mode=mode)
    ctx.set_scene(sc)
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    nseg = ctx.build_segments(cam, opt)
    ms = []
    for _ in range(3):
        ctx.gather(opt); ctx.synchronize(); ms.append(ctx.timing()["ms_gather"])
    print(f"{name:40s} {min(ms[1:]):9.1f} ms  {nseg * n / min(ms[1:]) * 1e-6:8.2f} G evals/s", flush=True)

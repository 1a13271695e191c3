"""Where the grid-medium gather (V3: 512 x 512, 10k VRLs, grid 128 x 50 x 128, HG) spends its time: variants, one launch each."""
import sys

sys.path.insert(0, ".")
from mitsuba_nlvrl_b200 import api, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
soa = scenes.synthetic_vrls(n, seed=2)
cam = scenes.box_fog_camera(512, 512)
ctx = api.Context(0)
for name, sc, opt in (
        ("grid medium, occluders, vis on (V3)", scenes.grid_fog_scene(occluders=True, res=(128, 50, 128)), scenes.default_options(visibility=1)),
        ("grid medium, occluders, vis off", scenes.grid_fog_scene(occluders=True, res=(128, 50, 128)), scenes.default_options(visibility=0)),
        ("grid medium, no occluders, vis on (shells only)", scenes.grid_fog_scene(occluders=False, res=(128, 50, 128)), scenes.default_options(visibility=1)),
        ("homogeneous HG, occluders, vis on, exact kernel", scenes.box_fog_scene(occluders=True, phase=1), scenes.default_options(visibility=1, math_mode=0)),
        ("homogeneous HG, occluders, vis off, exact kernel", scenes.box_fog_scene(occluders=True, phase=1), scenes.default_options(visibility=0, math_mode=0)),
        ("homogeneous HG, occluders, vis on, fast kernel", scenes.box_fog_scene(occluders=True, phase=1), scenes.default_options(visibility=1, math_mode=1))):
    ctx.set_scene(sc)
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    nseg = ctx.build_segments(cam, opt)
    ms = []
    for _ in range(2):
        ctx.clear_film(); ctx.gather(opt); ctx.synchronize(); ms.append(ctx.timing()["ms_gather"])
    print(f"{name:55s} {ms[-1]:9.1f} ms  {nseg * n / ms[-1] * 1e-6:8.2f} G evals/s", flush=True)

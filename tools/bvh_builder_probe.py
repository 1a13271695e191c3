"""The bench frame (1080p box-fog with occluders) at a reduced VRL count, shadow trees from the host build (binned SAH) and from the
device build (LBVH): gather time of the visibility kernel, one launch each.  usage: bvh_builder_probe.py [N_vrl]"""
import sys

sys.path.insert(0, ".")
from mitsuba_nlvrl_b200 import api, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
soa = scenes.synthetic_vrls(n, seed=1)
cam = scenes.box_fog_camera(1920, 1080)
sc = scenes.box_fog_scene(occluders=True)
opt = scenes.default_options(visibility=1, math_mode=1)
for builder in (0, 1):
    ctx = api.Context(0)
    ctx.set_bvh_builder(builder)
    ctx.set_scene(sc)
    info = ctx.bvh_info()
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    nseg = ctx.build_segments(cam, opt)
    ms = []
    for _ in range(3):
        ctx.clear_film(); ctx.gather(opt); ctx.synchronize(); ms.append(ctx.timing()["ms_gather"])
    print(f"builder {builder}: closest-hit depth {info['depth']}, shadow depth {info['shadow_depth']} ({info['shadow_trees_on_device']} trees on the device), "
          f"segments {ctx.timing()['ms_segments']:.3f} ms, gather {min(ms[1:]):.1f} ms = {nseg * n / min(ms[1:]) * 1e-6:.2f} G evals/s", flush=True)
    ctx.close()

"""Cost of the photon-map terms next to the VRL gather, at the bench frame's size: box-fog 1920x1080 with reflecting walls,
the plugin's default map sizes (volume_photons 1000, global_photons 250000, lookup 0.005 / 0.05 of the scene radius, k = 120).
Wall clock around the synchronous C-ABI calls (each ends with a stream synchronize); light tracing included.
    python tools/photon_cost_probe.py [width height]"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from mitsuba_nlvrl_b200 import abi, api, scenes  # noqa: E402

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (1920, 1080)
ctx = api.Context(0)
sc = scenes.box_fog_scene(occluders=True, wall_reflectance=(0.7, 0.6, 0.5))
e = sc.add_emitter((20.0, 20.0, 20.0))
sc.add_tris(scenes.box_tris((-3, 9.99, -3), (3, 9.995, 3), faces="y"), abi.SURF_BSDF_SMOOTH | abi.SURF_EMITTER | (e << 24))
cam = scenes.box_fog_camera(w, h)
opt = scenes.default_options(visibility=1)
ctx.set_scene(sc)
radius = float(np.linalg.norm([10.01] * 3))
gen = scenes.generate_desc(1000, seed=0, start_medium=0)


def timed(f, *a):
    t = time.perf_counter(); r = f(*a); return r, (time.perf_counter() - t) * 1e3


out = {"frame": [w, h]}
(nv, pv), out["ms_generate_vrls_1000"] = timed(ctx.generate_vrls, gen)
(n_vol, p_vol), out["ms_generate_volume_photons"] = timed(ctx.generate_photons, gen, 1000, 0.005 * radius)
(n_g, p_g), out["ms_generate_global_photons"] = timed(ctx.generate_surface_photons, gen, 250000, 0.05 * radius, 120)
out.update(n_volume_photons=n_vol, n_global_photons=n_g, global_particles=p_g)
_, out["ms_build_segments"] = timed(ctx.build_segments, cam, opt)
q = ctx.download_surface_queries()
out["n_surface_queries"] = len(q)
ctx.surface_gather_queries()                                   # warm-up
est, out["ms_surface_gather"] = timed(ctx.surface_gather_queries)
ctx.photon_gather_segments(opt)
d, out["ms_volume_photon_gather"] = timed(ctx.photon_gather_segments, opt)
out["mean_surface_estimate"] = float(est.mean()); out["mean_direct_term"] = float(d.mean())
out["note"] = "ms_*_gather include the result download (n x 3 floats); the VRL gather of the same frame with 100k VRLs takes 4930 ms"
print(json.dumps(out))

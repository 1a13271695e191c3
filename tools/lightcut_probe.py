"""The light cut on the bench frame (1080p box-fog with occluders): build time, one gather through the cut, the same frame by
the full gather.  usage: lightcut_probe.py N [threshold_error] [width height]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from mitsuba_nlvrl_b200 import api, scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
err = float(sys.argv[2]) if len(sys.argv) > 2 else 0.02
w, h = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1920, 1080)
soa = scenes.synthetic_vrls(n, seed=1)
cam = scenes.box_fog_camera(w, h)
ctx = api.Context(0)
sc = scenes.box_fog_scene(occluders=True)
ctx.set_scene(sc)
ctx.upload_vrls(scenes.vrlset_desc(soa))
opt = scenes.default_options(visibility=1, math_mode=1, use_light_cut=1, threshold_error=err)
nseg = ctx.build_segments(cam, opt)
t = time.time(); ctx.lightcut_build(seed_offset=1); wall = time.time() - t
st = ctx.lightcut_stats()
print(f"build: {n} VRLs -> {st.n_nodes} nodes, {st.build_searches} partner searches, {st.build_ms:.0f} ms ({1e3 * st.build_ms / max(st.build_searches - n, 1):.1f} us per search)", flush=True)
for rep in range(2):
    ctx.clear_film(); ctx.gather(opt); ctx.synchronize()
    ms = ctx.timing()["ms_gather"]
st = ctx.lightcut_stats()
cut = ctx.film_rgb(w, h).astype(np.float64)
print(f"cut gather: {ms:.1f} ms, {st.evaluations / st.queries:.0f} evaluations per query ({100 * st.evaluations / st.queries / n:.2f} % of the set), "
      f"{st.refinements / st.queries:.0f} refinements, largest heap {st.max_heap}, {st.overflows} overflows; "
      f"{(st.evaluations + st.queries) / ms * 1e-6:.3f} G cut evals/s; equivalent full-gather rate {nseg * n / ms * 1e-6:.1f} G evals/s", flush=True)
full_opt = scenes.default_options(visibility=1, math_mode=1)
if nseg * n <= 3e11:
    ctx.clear_film(); ctx.gather(full_opt); ctx.synchronize()
    ms_full = ctx.timing()["ms_gather"]
    full = ctx.film_rgb(w, h).astype(np.float64)
    rel = np.abs(cut - full).max(axis=1 if cut.ndim == 2 else 2) / np.maximum(np.abs(full).max(axis=1 if cut.ndim == 2 else 2), 1e-30)
    print(f"full gather: {ms_full:.1f} ms ({nseg * n / ms_full * 1e-6:.1f} G evals/s); cut is {ms_full / ms:.2f}x faster; per-pixel |cut - full| / full: "
          f"median {np.median(rel):.4f} q90 {np.quantile(rel, 0.9):.4f}; film mean ratio {cut.sum() / full.sum():.4f}", flush=True)

import os, sys, subprocess
import numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import mitsuba_nlvrl_b200 as pkg
import oracle
from parity_tools import rel_err
scenes, api, abi = pkg.scenes, pkg.api, pkg.abi
subprocess.check_call(["make", "-C", "host"], stdout=subprocess.DEVNULL)
subprocess.check_call(["host/vrl_host", "dump-xml", "tests/golden/scenes/cbox/cbox-nonlinear-homo-vrl.xml", "/tmp/nl.dump"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
sc, cam0, spp, seed, em = scenes.load_scene_dump("/tmp/nl.dump")
s2c = np.array(cam0.sample_to_camera, np.float64).reshape(4, 4)
p1, p0 = s2c @ np.array([1, 0.5, 0, 1]), s2c @ np.array([0, 0.5, 0, 1])
fov_x = 2 * np.degrees(np.arctan(0.5 * abs(p1[0] / p1[3] - p0[0] / p0[3]) / cam0.near_clip))
cam = scenes.make_camera(1280, 720, (0, 0, 0), (0, 0, 1), fov=fov_x, near_clip=cam0.near_clip, far_clip=cam0.far_clip)
cam.to_world[:] = list(cam0.to_world); cam.medium = cam0.medium
ctx = api.Context(0); ctx.set_scene(sc)
laser = ((-810.0, 154.9, -149.9), (0.9, -0.1, 0.0), (7.0, 30.0, 2.0)); d = np.array(laser[1]); laser = (laser[0], tuple(d / np.linalg.norm(d)), laser[2])
gen = scenes.generate_desc(1000, seed=7, start_medium=-1, laser=laser, min_vrl_length=50.0, max_steps=1000000)
n, particles = ctx.generate_vrls(gen)
soa, med = ctx.download_vrls(); soa = np.ascontiguousarray(soa[:, :1000]); scale = ctx.vrls_scale()
opt = scenes.default_options(visibility=1, samples_per_query=1, use_non_linear_camera=1, base_seed=seed)
ctx.upload_vrls(scenes.vrlset_desc(soa, scale=scale))
n_segs = ctx.build_segments(cam, opt); segs = ctx.download_segments()
li = ctx.gather_segments(opt).astype(np.float64)
y = segs["pixel"] // cam.width
band = np.nonzero(y == cam.height // 2)[0][::5]
osc = oracle.OracleScene(sc)
o32, o64, ev = osc.gather(scenes.vrlset_desc(soa, scale=scale), segs[band], opt)
err = rel_err(li[band], o64)
for k in np.nonzero(err > 1e-4)[0]:
    seg = segs[band[k]]
    print("segment", band[k], "pixel", seg["pixel"], "err", err[k], "o", seg["o"], "d", seg["d"], "maxt", seg["maxt"], "part", seg["part"] & 0xffff, "of", seg["part"] >> 16, "Li", o64[k])
    one = np.ascontiguousarray(seg.reshape(1)); ctx.upload_segments(one, 2048, 2048)
    g = np.zeros((1000, 3)); o = np.zeros((1000, 3))
    for j in range(1000):
        vd = scenes.vrlset_desc(np.ascontiguousarray(soa[:, j:j + 1]), scale=scale, first_index=j, total=1000)
        ctx.upload_vrls(vd); g[j] = ctx.gather_segments(opt)[0]; o[j] = osc.gather(vd, one, opt, n_threads=1)[1][0]
    diff = np.abs(g - o).sum(1); mag = np.abs(o).sum(1) + np.abs(g).sum(1)
    rel = diff / np.maximum(mag, 1e-300)
    bad = np.nonzero(rel > 1e-3)[0]
    print("  n bad", len(bad), "share of |Li| in bad evals", np.abs(o[bad]).sum() / np.abs(o).sum())
    # geometry of the bad evals
    so, sd = seg["o"].astype(float), seg["d"].astype(float)
    for j in bad[:8]:
        vo, vdd, vl = soa[0:3, j].astype(float), soa[3:6, j].astype(float), float(soa[6, j])
        st = np.linalg.norm(np.cross(vdd, sd))
        w = vo - so; nrm = np.cross(vdd, sd); hline = abs(np.dot(w, nrm)) / max(np.linalg.norm(nrm), 1e-30)
        print("   vrl", j, "rel", rel[j], "g", g[j][1], "o", o[j][1], "sin", st, "line dist", hline, "vlen", vl, "dist to vrl o", np.linalg.norm(w))
    ctx.upload_vrls(scenes.vrlset_desc(soa, scale=scale))

#!/usr/bin/env python
"""bench.py -- camera-segment x VRL evaluations per second and ms/frame at 1080p (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (config.workload): the synthetic "box-fog" scene of SURVEY.md 8(d), variant V2 (sphere + cuboid
occluders inside a homogeneous isotropic fog box, shadow query through the BVH per eval), 1920x1080, spp 1,
block_size 32, samples_per_query 1, importance sampling, 100k seeded VRLs: BASELINE.json configs[3]/[4]
shape, 2.07e11 evals per frame.  One step = one frame: camera-segment kernel + gather kernel (+ NCCL film
all-reduce when N > 1) + XYZAW film kernel.  N > 1: the VRL set is sharded across ranks (strong scaling: the
frame is fixed, each rank gathers all pixels against N_vrl/N VRLs), films summed with NCCL.

`value`   : whole-job evals/s with the VRL set, scene and camera resident on the device (CUDA events on the
            library's stream, max over ranks).
`e2e`     : same metric through vrlb_render() from HOST buffers: per step the VRL set is copied host->device
            and the XYZAW film device->host inside the timed region (wall clock around synchronised calls).
`roofline`: the gather kernel against the FP32 FMA issue peak measured on this GPU by the library's FFMA
            microkernel (MEASURED_PEAKS.json has no FP32 entry; the path is FP32/SFU-issue bound, not HBM).
`cpu_baseline` / --impl reference: the CPU restatement of the reference's scalar gather (oracle/, kind
            "port": the MSVC-only reference cannot be compiled here) on all host cores, on a crop of the frame.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

FLOPS_PER_EVAL = 210.0        # homogeneous + isotropic + importance sampling, visibility excluded
METRIC = "camera-seg x VRL evals/sec (1080p, 100k VRLs)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--width", type=int, default=1920)
    ap.add_argument("--height", type=int, default=1080)
    ap.add_argument("--vrls", type=int, default=100000)
    ap.add_argument("--q", type=int, default=1)
    ap.add_argument("--no-occluders", action="store_true")
    ap.add_argument("--force-vis", action="store_true", help="with --no-occluders: keep the shadow query (shells only)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--lightcut-1m", action="store_true", help="also time the light cut at 1e6 VRLs (minutes of tree build)")
    return ap.parse_args()


def workload(args):
    import mitsuba_nlvrl_b200 as pkg
    sc = pkg.scenes.box_fog_scene(occluders=not args.no_occluders)
    cam = pkg.scenes.box_fog_camera(args.width, args.height)
    opt = pkg.scenes.default_options(samples_per_query=args.q, visibility=0 if (args.no_occluders and not args.force_vis) else 1)
    soa = pkg.scenes.synthetic_vrls(args.vrls, seed=1)
    name = (f"box-fog {'V1 no occluders' if args.no_occluders else 'V2 occluders+BVH visibility'} "
            f"{args.width}x{args.height} spp1 q{args.q} N_vrl={args.vrls} homogeneous isotropic importance-sampled")
    return sc, cam, opt, soa, name


def bench_config(name, n_segs, frame_evals, world):
    """config of the JSON line; the reference arm prints the same dict key for key."""
    return {"workload": name, "n_segments": int(n_segs), "evals_per_frame": int(frame_evals),
            "parallelism": f"vrl-shard x{world} + NCCL fp32 film all-reduce" if world > 1 else "single GPU",
            "l2": "256 MB flush buffer written between steps; segment buffer 133 MB > L2"}


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 8:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for name, v in zip(self.NAMES, c[4:8]):
                if v == "Active":
                    reasons.add(name)
        os.unlink(self.f.name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # "under load": samples whose power draw is above the midpoint of the observed range
        lo, hi = min(pw), max(pw)
        load = [s for s, w in zip(sm, pw) if w >= lo + 0.5 * (hi - lo)] or sm
        return {"sm_mhz": float(np.median(load)), "sm_max_mhz": float(max(mx)), "power_w_max": float(hi),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_sample(sc, cam, opt, soa, seconds, n_threads=None):
    """Times the CPU oracle (restatement of the reference's scalar gather) on a centred crop of the same frame
    sized for ~`seconds` of work on all host cores.  Returns (evals_per_s, cores, sample description)."""
    import oracle
    import mitsuba_nlvrl_b200 as pkg
    n_threads = n_threads or os.cpu_count()
    vd = pkg.scenes.vrlset_desc(soa)
    osc = oracle.OracleScene(sc)
    cx, cy = cam.width // 2, cam.height // 2

    def run(crop):
        t0 = time.perf_counter()
        _, _, ev = osc.render(cam, opt, vd, n_threads=n_threads, crop=crop)
        return ev, time.perf_counter() - t0
    bs = opt.block_size
    nbx = (cam.width + bs - 1) // bs
    # the oracle parallelises over 32x32 blocks like the reference (integrator.cpp:111-142): calibrate one
    # thread on 4 pixels, then time a band that gives every thread the same number of pixel rows of a block
    bx, by = (cx // bs) * bs, (cy // bs) * bs
    ev, dt = run((bx, by, min(bx + 4, cam.width), by + 1))
    per_thread = ev / max(dt, 1e-9)
    per_px = max(soa.shape[1] * opt.samples_per_query, 1)
    tb = min(n_threads, nbx)                       # blocks side by side
    two_rows = n_threads > nbx and by >= bs        # straddle a block-row boundary to feed more threads
    rows = int(max(1, min(bs, round(per_thread * seconds / per_px / bs))))
    x0 = max(0, min(bx - (tb // 2) * bs, cam.width - tb * bs))
    crop = (x0, by - rows if two_rows else by, min(x0 + tb * bs, cam.width), by + rows)
    ev, dt = run(crop)
    rate = ev / max(dt, 1e-9)
    desc = (f"crop x[{crop[0]},{crop[2]}) y[{crop[1]},{crop[3]}) of the frame = {ev:.3e} evals in {dt:.1f} s, "
            f"{n_threads} std::threads over 32x32 blocks")
    return rate, n_threads, desc, ev, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sc, cam, opt, soa, name = workload(args)
    per_step = max(args.cpu_seconds / max(args.steps, 1), 2.0)
    for _ in range(min(args.warmup, 1)):
        cpu_sample(sc, cam, opt, soa, 1.0)
    tot_ev, tot_t, desc, cores = 0, 0.0, "", os.cpu_count()
    for _ in range(args.steps):
        rate, cores, desc, ev, dt = cpu_sample(sc, cam, opt, soa, per_step)
        tot_ev += ev
        tot_t += dt
    value = tot_ev / tot_t
    frame_evals = cam.width * cam.height * soa.shape[1] * opt.samples_per_query
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * tot_t / args.steps, "ms_per_frame_extrapolated": 1e3 * frame_evals / value,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(name, cam.width * cam.height, frame_evals, args.gpus),     # every pixel of the box-fog frame has one in-medium segment
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": cores, "kind": "port",
                             "sample": desc + "; each step = a bounded crop of the frame on the host cores"},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------ GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    import mitsuba_nlvrl_b200 as pkg
    api, scenes = pkg.api, pkg.scenes

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the VRL gather has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        import datetime
        # a rank that dies must not leave the others in a barrier for the driver's whole time limit
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=240))
    sc, cam, opt, soa, name = workload(args)
    n_total = soa.shape[1]
    lo, hi = pkg.sharding.shard_bounds(n_total, rank, world)
    shard, vd = pkg.sharding.shard_desc(soa, rank, world)

    ctx = api.Context(local)
    stream = torch.cuda.Stream()
    ctx.set_stream(stream.cuda_stream)
    if world > 1:
        ctx.comm_init(pkg.sharding.broadcast_comm_id(dist, api.Context.comm_unique_id, rank), rank, world)
    ctx.set_scene(sc)
    ctx.upload_vrls(vd)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2

    gather_ms = []

    def step():
        with torch.cuda.stream(stream):
            flush.zero_()
        n_segs = ctx.build_segments(cam, opt)
        ctx.gather(opt)
        if world > 1:
            ctx.film_allreduce()
        ctx.film_resolve()
        return n_segs

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def note(msg):
        if rank == 0:
            print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)

    note(f"scene + {hi - lo} VRLs resident; warm-up")
    for _ in range(args.warmup):
        n_segs = step()
        ctx.synchronize()
        note(f"warm-up step done: gather {ctx.timing()['ms_gather']:.1f} ms")
    barrier()
    fp32_peak = ctx.measure_fp32_peak()
    l0 = ctx.timing()["launches_lifetime"]
    clocks = ClockSampler()
    if rank == 0:
        clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        n_segs = step()
        ctx.synchronize()
        gather_ms.append(ctx.timing()["ms_gather"])
    e1.record(stream)
    barrier()
    note("timed steps done")
    clk = clocks.stop() if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    total_ms = float(ms.item())
    launches = ctx.timing()["launches_lifetime"] - l0
    frame_evals = n_segs * n_total * opt.samples_per_query
    value = frame_evals * args.steps / (total_ms * 1e-3)

    # ---- end to end through vrlb_render() from host buffers --------------------------------------------
    e2e = None
    if not args.no_e2e:
        out = np.zeros((cam.height, cam.width, 5), np.float32) if rank == 0 else False
        ctx.render(cam, opt, vd, out)                      # warm-up (allocations, pinned staging)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ctx.render(cam, opt, vd, out)
        barrier()
        dt = torch.tensor([time.perf_counter() - t0], device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": frame_evals * args.steps / float(dt.item()), "unit": "evals/s",
               "h2d_bytes_per_step": int(shard.nbytes) * world, "d2h_bytes_per_step": int(cam.width * cam.height * 5 * 4),
               "ms_per_frame": 1e3 * float(dt.item()) / args.steps}

    # ---- secondary figure: the same frame without occluders (variant V1: visibility == medium transmittance), the
    #      pure sample+shade rate of the gather kernel that the FP32-issue roofline describes -------------------
    note("e2e done")
    secondary = None
    if not args.no_occluders and not args.no_secondary:
        sc1 = scenes.box_fog_scene(occluders=False)
        opt1 = scenes.default_options(samples_per_query=args.q, visibility=0)
        ctx.set_scene(sc1)
        ctx.upload_vrls(vd)
        n1 = ctx.build_segments(cam, opt1)
        g1 = []
        for i in range(3):
            ctx.gather(opt1); ctx.synchronize()
            g1.append(ctx.timing()["ms_gather"])
        g1 = float(np.mean(g1[1:]))
        ev1 = n1 * (hi - lo) * opt1.samples_per_query
        secondary = {"workload": "same frame, variant V1 (no occluders, no shadow traversal), gather kernel only, this rank's shard",
                     "evals_per_s_per_gpu": ev1 / (g1 * 1e-3), "ms_per_launch": g1,
                     "fp32_tflops": ev1 * FLOPS_PER_EVAL / (g1 * 1e-3) * 1e-12,
                     "frac_of_fp32_peak": ev1 * FLOPS_PER_EVAL / (g1 * 1e-3) * 1e-12 / fp32_peak if fp32_peak else None}

    # ---- BASELINE's other workloads on this box, one gather launch each (same sharding; max over ranks) ----------
    def one_launch(scene_x, cam_x, opt_x, soa_x, label, kernel):
        shard_x, vd_x = pkg.sharding.shard_desc(soa_x, rank, world)
        ctx.set_scene(scene_x)
        ctx.upload_vrls(vd_x)
        n_x = ctx.build_segments(cam_x, opt_x)
        ctx.gather(opt_x); ctx.synchronize()                   # warm-up launch (first use of this kernel instantiation)
        with torch.cuda.stream(stream):
            flush.zero_()
        ctx.build_segments(cam_x, opt_x)
        ctx.gather(opt_x); ctx.synchronize()
        ms_x = torch.tensor([ctx.timing()["ms_gather"]], device="cuda")
        if world > 1:
            dist.all_reduce(ms_x, op=dist.ReduceOp.MAX)
        ms_x = float(ms_x.item())
        ev_x = n_x * soa_x.shape[1] * opt_x.samples_per_query
        return {"workload": label, "n_segments": int(n_x), "evals_per_frame": int(ev_x), "evals_per_s": ev_x / (ms_x * 1e-3),
                "ms_per_frame_gather": ms_x, "kernel": kernel,
                "fp32_tflops": ev_x * FLOPS_PER_EVAL / (ms_x * 1e-3) * 1e-12 / world,
                "frac_of_fp32_peak": (ev_x * FLOPS_PER_EVAL / (ms_x * 1e-3) * 1e-12 / world / fp32_peak) if fp32_peak else None}

    workloads = []
    if not args.no_occluders and not args.no_secondary:
        # configs[1] shape (V3): grid medium 128 x 50 x 128, HG phase, 512 x 512, 10k VRLs, shadow rays through the BVH
        workloads.append(one_launch(scenes.grid_fog_scene(occluders=True, res=(128, 50, 128)), scenes.box_fog_camera(512, 512),
                                    scenes.default_options(visibility=1), scenes.synthetic_vrls(10000, seed=2),
                                    "box-fog V3 grid medium 128x50x128 HG, 512x512, N_vrl=10000, BVH visibility",
                                    "gather_kernel<GRID> (exactly rounded: the stochastic transmittance is a chain of accept / reject decisions)"))
        note("V3 done")
        # the north star's target (V4): graded-IOR medium, curved camera paths in straight parts, 1920 x 1080, 100k VRLs
        workloads.append(one_launch(scenes.nl_fog_scene(occluders=True, res_y=320), scenes.box_fog_camera(args.width, args.height),
                                    scenes.default_options(visibility=1, use_non_linear_camera=1), soa,
                                    f"box-fog V4 graded IOR 1.0->0.95 (320 layers), curved camera paths, {args.width}x{args.height}, N_vrl={n_total}, BVH visibility",
                                    "gather_fast_vis_kernel"))
        note("V4 done")
        # the bench frame through the VRL light cut (use_light_cut): the reference's tree built on the device, one
        # cut per camera segment.  N = 1e6 (--lightcut-1m) builds for minutes and is measured once per round (profiles/).
        if world == 1:
            for n_lc in ([min(n_total, 100000)] + ([1000000] if args.lightcut_1m else [])):
                soa_lc = soa if n_lc == n_total else scenes.synthetic_vrls(n_lc, seed=1)
                ctx.set_scene(sc)
                ctx.upload_vrls(scenes.vrlset_desc(soa_lc))
                opt_lc = scenes.default_options(visibility=1, use_light_cut=1, threshold_error=0.02)
                n_x = ctx.build_segments(cam, opt_lc)
                ctx.lightcut_build(seed_offset=1)
                ctx.clear_film(); ctx.gather(opt_lc); ctx.synchronize()
                st, ms_x = ctx.lightcut_stats(), ctx.timing()["ms_gather"]
                workloads.append({"workload": f"box-fog V2 {args.width}x{args.height}, N_vrl={n_lc}, use_light_cut threshold_error=0.02 threshold_better_dist=8",
                                  "kernel": "lightcut_gather_kernel (exactly rounded getContrib; the reference's tree and refinement order)",
                                  "n_segments": int(n_x), "tree_nodes": int(st.n_nodes), "build_ms": float(st.build_ms), "build_partner_searches": int(st.build_searches),
                                  "ms_per_frame_gather": float(ms_x), "evaluations_per_query": st.evaluations / max(st.queries, 1),
                                  "fraction_of_set_evaluated": st.evaluations / max(st.queries, 1) / n_lc, "largest_heap": int(st.max_heap),
                                  "overflowed_queries": int(st.overflows), "cut_evals_per_s": (st.evaluations + st.queries) / (ms_x * 1e-3),
                                  "full_gather_equivalent_evals_per_s": n_x * n_lc / (ms_x * 1e-3)})
                note(f"light cut N={n_lc} done")

    # ---- N > 1: is the sharded, all-reduced film the single-GPU film?  (outside every timed region) ------------------
    multi_gpu_check = None
    if world > 1 and not args.no_secondary:
        # the bench scene plus what only rank 0 may add (the films are SUMMED): an area light the camera sees directly
        # (segments.cu add_direct) and a volume photon map (the direct in-scattering term, photon_gather_impl)
        n_chk = min(8192, n_total)
        soa_chk = np.ascontiguousarray(soa[:, :n_chk])
        sc_chk = scenes.box_fog_scene(occluders=True)
        e_id = sc_chk.add_emitter((4.0, 3.0, 2.0))
        quad = np.array([[[-2, 6, -3], [2, 6, -3], [2, 9, -3]], [[-2, 6, -3], [2, 9, -3], [-2, 9, -3]]], np.float32)
        sc_chk.add_tris(quad[:, ::-1], pkg.abi.SURF_BSDF_SMOOTH | pkg.abi.SURF_EMITTER | pkg.abi.surf_emitter_id(e_id))
        ph = scenes.photonset_desc(*scenes.synthetic_photons(2000, seed=9), scale=1.0 / 2000, lookup_radius=0.6)
        ctx.set_scene(sc_chk)
        ctx.upload_photons(ph)
        _, vd_chk = pkg.sharding.shard_desc(soa_chk, rank, world)
        ctx.upload_vrls(vd_chk)
        ctx.build_segments(cam, opt); ctx.gather(opt); ctx.film_allreduce(); ctx.synchronize()
        sharded = ctx.film_rgb(cam.width, cam.height).astype(np.float64) if rank == 0 else None
        barrier()
        if rank == 0:
            ctx.upload_vrls(scenes.vrlset_desc(soa_chk))
            ctx.build_segments(cam, opt); ctx.gather(opt); ctx.synchronize()
            single = ctx.film_rgb(cam.width, cam.height).astype(np.float64)
            rel = np.abs(sharded - single) / np.maximum(np.abs(single), 1e-6 * np.abs(single).max())
            lit = single.reshape(-1, 3).max(axis=1) > 1.0     # pixels that see the light: (4, 3, 2) x the fog's transmittance (0.46), added once
            multi_gpu_check = {"n_vrls": int(n_chk), "max_rel": float(rel.max()), "pixels_above_1e-5": int((rel.max(axis=-1) > 1e-5).sum()),
                               "pixels_seeing_the_emitter": int(lit.sum()), "volume_photons": 2000,
                               "what": f"film of {world} VRL shards summed by NCCL (emitter radiance and the photon term added by rank 0 only) vs the same frame on one GPU"}
            # no assert here: the other ranks wait in the barrier below; the verdict is raised after the line is printed
            multi_gpu_check["ok"] = bool(rel.max() <= 1e-4 and lit.sum() > 100)
        ctx.clear_photons()
        barrier()

    note("secondary done")
    if rank == 0:
        g_ms = float(np.mean(gather_ms))
        shard_evals = n_segs * (hi - lo) * opt.samples_per_query
        achieved = shard_evals * FLOPS_PER_EVAL / (g_ms * 1e-3) * 1e-12
        # DRAM traffic and pipe utilisation cannot be measured without the profiler: they come from the kept ncu --set full
        # capture of this kernel on this frame (a 2000-VRL launch; the traffic -- the segment buffer read once, the film
        # written once -- does not grow with the VRL count), named in traffic_source
        traffic, kept = None, {}
        tr_file = os.path.join(ROOT, "profiles", "gather_traffic.json")
        if os.path.exists(tr_file) and not args.no_occluders:
            try:
                kept = json.load(open(tr_file))
                traffic = kept.get("dram_bytes_per_launch")
            except Exception:
                traffic, kept = None, {}
        line = {"metric": METRIC, "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": total_ms / args.steps, "ms_per_frame": total_ms / args.steps, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": bench_config(name, n_segs, frame_evals, world),
                "roofline": {"bound": "fp32", "achieved": achieved, "peak": fp32_peak, "unit": "TFLOP/s",
                             "frac": achieved / fp32_peak if fp32_peak else None, "traffic": traffic,
                             "kernel": "gather_fast_vis_kernel" if not args.no_occluders else "gather_fast_kernel", "ms_per_launch": g_ms, "flops_per_eval": FLOPS_PER_EVAL,
                             "peak_source": "FFMA microkernel measured in this run (vrlb_measure_fp32_peak); "
                                            "MEASURED_PEAKS.json has no FP32 entry",
                             "traffic_source": kept.get("source"), "pipe_fma_pct": kept.get("pipe_fma_pct"),
                             "pipe_alu_pct": kept.get("pipe_alu_pct"), "pipe_xu_pct": kept.get("pipe_xu_pct"),
                             "issue_active_pct": kept.get("issue_active_pct")},
                "clocks": clk, "gpu_launches": int(launches), "e2e": e2e, "secondary": secondary, "workloads": workloads,
                "multi_gpu_check": multi_gpu_check}
        if world == 1 and not args.no_cpu_baseline:
            rate, cores, desc, _, _ = cpu_sample(sc, cam, opt, soa, args.cpu_seconds)
            line["cpu_baseline"] = {"value": rate, "unit": "evals/s", "cores": cores, "kind": "port", "sample": desc}
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    if rank == 0 and multi_gpu_check is not None and not multi_gpu_check.get("ok", False):
        print(f"[bench] multi_gpu_check FAILED: {multi_gpu_check}", file=sys.stderr)
        return 3
    return 0


def main():
    import faulthandler
    import signal
    faulthandler.register(signal.SIGUSR1, all_threads=True)       # `kill -USR1 <pid>` dumps the Python stacks to stderr
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())

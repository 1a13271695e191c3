"""TEST-SIDE loader for the CPU oracle (oracle/libvrl_oracle.so).  Product code never imports this."""
import ctypes as C
import os
import subprocess

import numpy as np

import mitsuba_nlvrl_b200 as pkg
from mitsuba_nlvrl_b200 import abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
RNG_ADDRESSED, RNG_SEQUENTIAL = 0, 1


def _build():
    so = os.path.join(ORACLE_DIR, "libvrl_oracle.so")
    src = os.path.join(ORACLE_DIR, "vrl_oracle.cpp")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "libvrl_oracle.so"], stdout=subprocess.DEVNULL)
    return so


_lib = None


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_build())
        f3 = C.POINTER(C.c_float)
        L.vrlo_sampler_stream.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, f3, C.POINTER(C.c_uint32)]
        L.vrlo_pcg32_seed.argtypes = [C.c_uint64, C.c_uint64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.vrlo_pcg32_next_u32.argtypes = [C.POINTER(C.c_uint64), C.c_uint64]
        L.vrlo_pcg32_next_u32.restype = C.c_uint32
        L.vrlo_pcg32_advance.argtypes = [C.POINTER(C.c_uint64), C.c_uint64, C.c_int64]
        L.vrlo_spiral_block_ids.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_uint32), C.c_uint32]
        L.vrlo_spiral_block_ids.restype = C.c_uint32
        L.vrlo_morton_decode.argtypes = [C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.vrlo_find_closest_point.argtypes = [f3, f3, C.c_float, f3, f3, C.c_float, f3, f3]
        L.vrlo_sampling_vrl.argtypes = [f3, f3, C.c_float, f3, f3, C.c_float, C.c_int, C.c_float, C.c_float, f3]
        L.vrlo_sampling_vrl.restype = C.c_int
        L.vrlo_phase_eval.argtypes = [C.POINTER(abi.Medium), f3, f3]
        L.vrlo_phase_eval.restype = C.c_float
        L.vrlo_get_contrib.argtypes = [C.POINTER(abi.SceneDesc), C.POINTER(abi.Medium), f3, f3, f3, C.c_float,
                                       C.c_int, C.c_float, C.c_float, f3]
        L.vrlo_scene_create.argtypes = [C.POINTER(abi.SceneDesc)]
        L.vrlo_scene_create.restype = C.c_void_p
        L.vrlo_scene_destroy.argtypes = [C.c_void_p]
        L.vrlo_ray_intersect.argtypes = [C.c_void_p, f3, f3, C.c_float, C.c_float, f3, f3, f3]
        L.vrlo_ray_intersect.restype = C.c_int
        L.vrlo_eval_transmittance.argtypes = [C.c_void_p, f3, f3, C.c_int, f3]
        L.vrlo_grid_eval.argtypes = [C.c_void_p, C.c_int, f3]
        L.vrlo_grid_eval.restype = C.c_float
        L.vrlo_grid_transmittance.argtypes = [C.c_void_p, C.c_int, f3, f3, C.c_float, C.c_uint64]
        L.vrlo_grid_transmittance.restype = C.c_float
        L.vrlo_photon_gather_segments.argtypes = [C.c_void_p, C.POINTER(abi.PhotonSetDesc), C.c_void_p, C.c_uint64, C.POINTER(abi.Options), f3]
        L.vrlo_surface_gather_queries.argtypes = [C.POINTER(abi.PhotonSetDesc), C.c_int32, C.POINTER(abi.PhotonSetDesc), C.c_int32, C.c_void_p, C.c_uint64, f3]
        L.vrlo_build_surface_queries.argtypes = [C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Options), C.c_uint64, C.c_void_p, C.c_uint64]
        L.vrlo_build_surface_queries.restype = C.c_uint64
        L.vrlo_estimate_radiance_volume.argtypes = [C.POINTER(abi.Medium), C.POINTER(abi.PhotonSetDesc), f3, f3, C.c_float, f3]
        L.vrlo_sample_delta.argtypes = [C.c_uint32, f3, f3, f3, C.c_float, f3, f3, C.POINTER(C.c_int)]
        L.vrlo_sample_delta.restype = C.c_int
        L.vrlo_set_build_crop.argtypes = [C.c_int] * 4
        L.vrlo_set_build_crop.restype = None
        L.vrlo_nl_ior.argtypes = [C.POINTER(abi.Medium), f3]
        L.vrlo_nl_ior.restype = C.c_float
        L.vrlo_nl_march.argtypes = [C.c_void_p, C.c_int, f3, f3, C.c_float, f3]
        L.vrlo_vrls_to_long.argtypes = [C.c_void_p, f3, C.c_uint64, f3]
        L.vrlo_vrls_to_long.restype = C.c_uint64
        L.vrlo_vrls_dice.argtypes = [C.c_void_p, f3, C.POINTER(C.c_uint16), C.c_uint64, C.c_uint32, f3, C.POINTER(C.c_uint16), C.c_uint64]
        L.vrlo_vrls_dice.restype = C.c_uint64
        L.vrlo_build_segments.argtypes = [C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Options), C.c_uint64, C.c_int,
                                          C.c_void_p, C.c_uint64, f3, f3]
        L.vrlo_build_segments.restype = C.c_uint64
        L.vrlo_gather_segments.argtypes = [C.c_void_p, C.POINTER(abi.VrlSetDesc), C.c_void_p, C.c_uint64,
                                           C.POINTER(abi.Options), C.c_int, C.c_int, f3, C.POINTER(C.c_double)]
        L.vrlo_gather_segments.restype = C.c_uint64
        L.vrlo_render.argtypes = [C.c_void_p, C.POINTER(abi.Camera), C.POINTER(abi.Options), C.POINTER(abi.VrlSetDesc),
                                  C.c_int, C.c_int, C.POINTER(C.c_int32), f3, f3]
        L.vrlo_render.restype = C.c_uint64
        L.vrlo_lightcut_build.argtypes = [C.POINTER(abi.VrlSetDesc), C.c_uint64, C.c_uint64, C.c_int]
        L.vrlo_lightcut_build.restype = C.c_void_p
        L.vrlo_lightcut_destroy.argtypes = [C.c_void_p]
        L.vrlo_lightcut_n_nodes.argtypes = [C.c_void_p]
        L.vrlo_lightcut_n_nodes.restype = C.c_uint64
        L.vrlo_lightcut_build_draws.argtypes = [C.c_void_p]
        L.vrlo_lightcut_build_draws.restype = C.c_uint64
        L.vrlo_lightcut_nodes.argtypes = [C.c_void_p, f3, C.POINTER(C.c_int32)]
        L.vrlo_lc_node_dist.argtypes = [f3, f3]
        L.vrlo_lc_node_dist.restype = C.c_float
        L.vrlo_lc_merge.argtypes = [f3, f3, C.c_float, f3]
        L.vrlo_lc_box_dist.argtypes = [f3, f3, f3, f3, C.c_float]
        L.vrlo_lc_box_dist.restype = C.c_float
        L.vrlo_lc_seg_dist.argtypes = [f3, f3, f3, C.c_float]
        L.vrlo_lc_seg_dist.restype = C.c_float
        L.vrlo_lightcut_bound.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, f3, C.c_int, C.c_int]
        L.vrlo_lightcut_bound.restype = C.c_float
        L.vrlo_lightcut_query.argtypes = [C.c_void_p, C.c_void_p, C.c_float, f3, C.c_int, C.POINTER(abi.Options), C.c_int,
                                          C.c_uint64, C.c_uint64, f3, C.POINTER(C.c_double)]
        L.vrlo_lightcut_gather_segments.argtypes = [C.c_void_p, C.c_void_p, C.c_float, C.c_void_p, C.c_uint64, C.POINTER(abi.Options),
                                                    C.c_int, C.c_int, f3, f3]
        L.vrlo_lightcut_gather_segments.restype = C.c_uint64
        _lib = L
    return _lib


def f32(a):
    return np.ascontiguousarray(a, np.float32)


def fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class OracleScene:
    def __init__(self, scene):
        self.scene = scene
        self._desc = scene.desc()
        self.h = lib().vrlo_scene_create(C.byref(self._desc))

    def __del__(self):
        if getattr(self, "h", None):
            lib().vrlo_scene_destroy(self.h)
            self.h = None

    def intersect(self, o, d, mint=0.0, maxt=np.inf):
        o, d = f32(o), f32(d)
        t, u, v = (C.c_float(), C.c_float(), C.c_float())
        prim = lib().vrlo_ray_intersect(self.h, fp(o), fp(d), mint, maxt, C.byref(t), C.byref(u), C.byref(v))
        return prim, t.value, u.value, v.value

    def transmittance(self, p1, p2, medium=0):
        p1, p2, out = f32(p1), f32(p2), np.zeros(3, np.float32)
        lib().vrlo_eval_transmittance(self.h, fp(p1), fp(p2), medium, fp(out))
        return out

    def grid_eval(self, p, medium=0):
        p = f32(p)
        return lib().vrlo_grid_eval(self.h, medium, fp(p))

    def grid_transmittance(self, o, d, maxt, seed, medium=0):
        o, d = f32(o), f32(d)
        return lib().vrlo_grid_transmittance(self.h, medium, fp(o), fp(d), maxt, seed)

    def build_surface_queries(self, cam, opt, n_vrls_total):
        """Surface queries of the camera paths (vrl.cpp:671-680)."""
        cap = cam.width * cam.height * max(opt.spp, 1)
        q = np.zeros(cap, abi.SURFACE_QUERY_DTYPE)
        n = lib().vrlo_build_surface_queries(self.h, C.byref(cam), C.byref(opt), n_vrls_total, q.ctypes.data, cap)
        return q[:n]

    def photon_gather(self, photon_desc, segs, opt):
        """Direct in-scattering term per segment (vrl.cpp:523-611), throughput included."""
        segs = np.ascontiguousarray(segs)
        out = np.zeros((len(segs), 3), np.float32)
        lib().vrlo_photon_gather_segments(self.h, C.byref(photon_desc), segs.ctypes.data, len(segs), C.byref(opt), fp(out))
        return out

    def nl_march(self, o, d, medium=0, mint=0.0):
        """Fixed-step eikonal march from o along d to the next surface: (found, exit point, exit direction, length, steps)."""
        out = np.zeros(8, np.float32)
        ok = lib().vrlo_nl_march(self.h, medium, fp(f32(o)), fp(f32(d)), mint, fp(out))
        return bool(ok), out[0:3].copy(), out[3:6].copy(), float(out[6]), int(out[7])

    def vrls_to_long(self, soa10):
        """VRLMap::toLong: (new lengths, number of VRLs that see no surface)."""
        soa10 = np.ascontiguousarray(soa10, np.float32)
        n = soa10.shape[1]
        out = np.zeros(n, np.float32)
        missed = lib().vrlo_vrls_to_long(self.h, fp(soa10), n, fp(out))
        return out, int(missed)

    def vrls_dice(self, soa10, avg_number_dice, medium=None):
        """VRLMap::dicingVRL: the diced [10, m] set (+ medium ids)."""
        soa10 = np.ascontiguousarray(soa10, np.float32)
        n = soa10.shape[1]
        med = None if medium is None else np.ascontiguousarray(medium, np.uint16)
        mp = None if med is None else med.ctypes.data_as(C.POINTER(C.c_uint16))
        cap = int(lib().vrlo_vrls_dice(self.h, fp(soa10), mp, n, avg_number_dice, None, None, 0))
        out, om = np.zeros((10, max(cap, 1)), np.float32), np.zeros(max(cap, 1), np.uint16)
        m = int(lib().vrlo_vrls_dice(self.h, fp(soa10), mp, n, avg_number_dice, fp(out), om.ctypes.data_as(C.POINTER(C.c_uint16)), cap))
        assert m == cap
        return out[:, :m], om[:m]

    def build_segments(self, cam, opt, n_vrls_total, rng_mode=RNG_ADDRESSED, cap=None, crop=None):
        """crop = (x0, y0, x1, y1): only those pixels' paths (the whole film otherwise)."""
        lib().vrlo_set_build_crop(*(crop if crop is not None else (0, 0, 0, 0)))
        try:
            return self._build_segments(cam, opt, n_vrls_total, rng_mode, cap)
        finally:
            lib().vrlo_set_build_crop(0, 0, 0, 0)

    def _build_segments(self, cam, opt, n_vrls_total, rng_mode, cap):
        npix = cam.width * cam.height
        cap = cap or npix * opt.spp * 4
        segs = np.zeros(cap, abi.SEGMENT_DTYPE)
        direct = np.zeros((cam.height, cam.width, 3), np.float32)
        aw = np.zeros((cam.height, cam.width, 2), np.float32)
        n = lib().vrlo_build_segments(self.h, C.byref(cam), C.byref(opt), n_vrls_total, rng_mode,
                                      segs.ctypes.data, cap, fp(direct), fp(aw))
        assert n <= cap, "segment capacity exceeded"
        return segs[:n].copy(), direct, aw

    def gather(self, vrl_desc, segs, opt, rng_mode=RNG_ADDRESSED, n_threads=None):
        n = len(segs)
        segs = np.ascontiguousarray(segs)
        o32, o64 = np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float64)
        ev = lib().vrlo_gather_segments(self.h, C.byref(vrl_desc), segs.ctypes.data, n, C.byref(opt), rng_mode,
                                        n_threads or os.cpu_count(), fp(o32), o64.ctypes.data_as(C.POINTER(C.c_double)))
        return o32, o64, ev

    def render(self, cam, opt, vrl_desc, rng_mode=RNG_ADDRESSED, n_threads=None, crop=None):
        xyzaw = np.zeros((cam.height, cam.width, 5), np.float32)
        rgb = np.zeros((cam.height, cam.width, 3), np.float32)
        cr = (C.c_int32 * 4)(*crop) if crop is not None else None
        ev = lib().vrlo_render(self.h, C.byref(cam), C.byref(opt), C.byref(vrl_desc), rng_mode,
                               n_threads or os.cpu_count(), cr, fp(xyzaw), fp(rgb))
        return xyzaw, rgb, ev


def surface_gather(queries, global_desc=None, k_global=0, caustic_desc=None, k_caustic=0):
    """caustic + global surface estimate per query (vrl.cpp:671-680, photonmap.h:131-211), before the path weight."""
    q = np.ascontiguousarray(queries, abi.SURFACE_QUERY_DTYPE)
    out = np.zeros((len(q), 3), np.float32)
    lib().vrlo_surface_gather_queries(C.byref(global_desc) if global_desc is not None else None, k_global,
                                      C.byref(caustic_desc) if caustic_desc is not None else None, k_caustic,
                                      q.ctypes.data, len(q), fp(out))
    return out


def sample_delta(flags, prm, n, rd, sample1):
    """(wo, weight, transmitted) of the delta BSDF sample, or None when the lobe is inactive."""
    prm, n, rd = f32(prm), f32(n), f32(rd)
    wo, wt = np.zeros(3, np.float32), np.zeros(3, np.float32)
    t = C.c_int()
    if not lib().vrlo_sample_delta(flags, fp(prm), fp(n), fp(rd), sample1, fp(wo), fp(wt), C.byref(t)):
        return None
    return wo, wt, bool(t.value)


def estimate_radiance_volume(medium, photon_desc, gp, wo, radius):
    """PhotonMap::estimateRadianceVolume at one gather point (photonmap.h:216-271)."""
    gp, wo = f32(gp), f32(wo)
    out = np.zeros(3, np.float32)
    lib().vrlo_estimate_radiance_volume(C.byref(medium), C.byref(photon_desc), fp(gp), fp(wo), radius, fp(out))
    return out


class OracleLightcut:
    """The reference's light tree over a VRL set (vrl_lightcut.h:196-212, 515-627) and queries through its cut."""

    def __init__(self, vrl_desc, seed_offset=0, base_seed=0, n_threads=None):
        self._desc = vrl_desc
        self.scale = vrl_desc.scale
        self.h = lib().vrlo_lightcut_build(C.byref(vrl_desc), base_seed, seed_offset, n_threads or os.cpu_count())

    def __del__(self):
        if getattr(self, "h", None):
            lib().vrlo_lightcut_destroy(self.h)
            self.h = None

    @property
    def build_draws(self):
        return lib().vrlo_lightcut_build_draws(self.h)

    def nodes(self):
        """Preorder: (f [n,13] = flux, add_flux, lo, hi, max_len; i [n,6] = leaf, rep VRL, same child, n_leaves, id, parent id)."""
        n = lib().vrlo_lightcut_n_nodes(self.h)
        f, i = np.zeros((n, 13), np.float32), np.zeros((n, 6), np.int32)
        lib().vrlo_lightcut_nodes(self.h, fp(f), i.ctypes.data_as(C.POINTER(C.c_int32)))
        return f, i

    def bound(self, osc, node, parts, better_dist=8):
        parts = f32(parts).reshape(-1, 7)
        return lib().vrlo_lightcut_bound(osc.h, self.h, node, fp(parts), len(parts), better_dist)

    def query(self, osc, parts, opt, seed_offset, rng_mode=RNG_SEQUENTIAL, scale=1.0, base_seed=0):
        parts = f32(parts).reshape(-1, 7)
        out, st = np.zeros(3, np.float32), np.zeros(5, np.float64)
        lib().vrlo_lightcut_query(osc.h, self.h, scale, fp(parts), len(parts), C.byref(opt), rng_mode, base_seed, seed_offset,
                                  fp(out), st.ctypes.data_as(C.POINTER(C.c_double)))
        return out, dict(evals=int(st[0]), draws=int(st[1]), pops=int(st[2]), max_heap=int(st[3]), min_margin=float(st[4]))

    def gather(self, osc, segs, opt, rng_mode=RNG_ADDRESSED, n_threads=None):
        """Per segment: Li (fp32, the reference's arithmetic) and stats [evals, pops, largest heap, min decision margin]."""
        n = len(segs)
        segs = np.ascontiguousarray(segs)
        out, st = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32)
        ev = lib().vrlo_lightcut_gather_segments(osc.h, self.h, self.scale, segs.ctypes.data, n, C.byref(opt), rng_mode,
                                                 n_threads or os.cpu_count(), fp(out), fp(st))
        assert ev != 2 ** 64 - 1, "light cut over a non-homogeneous medium (the reference raises Log(Error), vrl_lightcut.h:485)"
        return out, st, ev


def lc_node_dist(a10, b10):
    a10, b10 = f32(a10), f32(b10)
    return lib().vrlo_lc_node_dist(fp(a10), fp(b10))


def lc_merge(a10, b10, v):
    a10, b10, out = f32(a10), f32(b10), np.zeros(14, np.float32)
    lib().vrlo_lc_merge(fp(a10), fp(b10), v, fp(out))
    return dict(pick=int(out[0]), flux=out[1:4], add_flux=out[4:7], lo=out[7:10], hi=out[10:13], max_len=out[13])


def lc_box_dist(lo, hi, ray7):
    lo, hi, r = f32(lo), f32(hi), f32(ray7)
    o, d = f32(r[0:3]), f32(r[3:6])
    return lib().vrlo_lc_box_dist(fp(lo), fp(hi), fp(o), fp(d), float(r[6]))


def lc_seg_dist(vrl10, ray7):
    v, r = f32(vrl10), f32(ray7)
    o, d = f32(r[0:3]), f32(r[3:6])
    return lib().vrlo_lc_seg_dist(fp(v), fp(o), fp(d), float(r[6]))

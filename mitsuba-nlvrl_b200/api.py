"""ctypes binding of libvrl_b200.so (include/vrl_b200.h).  This is the host-side mirror the tests, the
bench and the integrator front-end use; it adds no computation.  The library must exist in-tree
(csrc/libvrl_b200.so, built by `make -C mitsuba-nlvrl_b200/csrc` or __graft_entry__.build()); there is no
fallback of any kind: a missing library or a missing GPU raises."""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VRLB_LIB_PATH") or os.path.join(_HERE, "csrc", "libvrl_b200.so")   # the override is for kernel-variant experiments (tools/)

# every symbol include/vrl_b200.h declares
EXPORTS = [
    "vrlb_abi_version", "vrlb_create", "vrlb_destroy", "vrlb_last_error", "vrlb_set_stream", "vrlb_scene_set",
    "vrlb_vrls_upload", "vrlb_vrls_download", "vrlb_vrls_count", "vrlb_segments_build", "vrlb_segments_upload",
    "vrlb_segments_download", "vrlb_segments_count", "vrlb_film_clear", "vrlb_gather", "vrlb_gather_segments",
    "vrlb_film_device_ptr", "vrlb_film_download_xyzaw", "vrlb_film_download_rgb", "vrlb_render", "vrlb_cancel",
    "vrlb_get_timing", "vrlb_synchronize", "vrlb_comm_unique_id", "vrlb_comm_init", "vrlb_comm_destroy",
    "vrlb_film_allreduce", "vrlb_film_resolve", "vrlb_measure_fp32_peak", "vrlb_vrls_generate", "vrlb_vrls_scale",
    "vrlb_vrls_download_depth", "vrlb_shadow_proxies", "vrlb_bvh_depths", "vrlb_vrls_to_long", "vrlb_vrls_dice", "vrlb_vrls_dice_noalloc",
    "vrlb_surface_photons_upload", "vrlb_surface_photons_generate", "vrlb_surface_photons_count", "vrlb_surface_photons_download",
    "vrlb_surface_queries_upload", "vrlb_surface_queries_download", "vrlb_surface_gather_queries",
    "vrlb_photons_upload", "vrlb_photons_count", "vrlb_photon_gather_segments", "vrlb_photons_generate", "vrlb_photons_download",
    "vrlb_lightcut_build", "vrlb_lightcut_download", "vrlb_lightcut_get_stats", "vrlb_lightcut_gather_segments",
    "vrlb_set_bvh_builder", "vrlb_bvh_info",
]


SOLID_DTYPE = np.dtype([("lo", "<f4", 3), ("hi", "<f4", 3), ("root", "<u4"), ("flags", "<u4"), ("c", "<f4", 3), ("r_out2", "<f4"),
                        ("ci", "<f4", 3), ("r_in2", "<f4"), ("ax", "<f4", (3, 3)), ("oc", "<f4", 3), ("h_out", "<f4", 3), ("h_in", "<f4", 3)])


def shadow_proxies(tri_verts):
    """Host-only diagnostic (vrlb_shadow_proxies): (solids as a SOLID_DTYPE array, n_shells, n_main_tris)."""
    v = np.ascontiguousarray(tri_verts, np.float32).reshape(-1)
    out = np.zeros(8, SOLID_DTYPE)
    counts = (C.c_uint32 * 3)()
    rc = load_library().vrlb_shadow_proxies(v.ctypes.data_as(C.POINTER(C.c_float)), v.size // 9,
                                            out.ctypes.data_as(C.POINTER(C.c_float)), len(out), counts)
    if rc != abi.OK:
        raise VrlbError(rc, "vrlb_shadow_proxies failed")
    return out[:counts[0]], int(counts[1]), int(counts[2])


class VrlbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libvrl_b200 error {code}: {msg}")
        self.code = code


_lib = None


def load_library():
    """Loads the in-tree shared library and declares prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FileNotFoundError(f"{LIB_PATH} is missing: build it with `make -C {os.path.dirname(LIB_PATH)}` "
                                "(the CUDA extension is mandatory, there is no CPU path)")
    L = C.CDLL(LIB_PATH)
    vp, f32p = C.c_void_p, C.POINTER(C.c_float)
    L.vrlb_abi_version.restype = C.c_int
    L.vrlb_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.vrlb_destroy.argtypes = [vp]
    L.vrlb_destroy.restype = None
    L.vrlb_last_error.argtypes = [vp]
    L.vrlb_last_error.restype = C.c_char_p
    L.vrlb_set_stream.argtypes = [vp, vp]
    L.vrlb_scene_set.argtypes = [vp, C.POINTER(abi.SceneDesc)]
    L.vrlb_vrls_upload.argtypes = [vp, C.POINTER(abi.VrlSetDesc)]
    L.vrlb_vrls_download.argtypes = [vp, f32p, C.POINTER(C.c_uint16), C.c_uint64, C.POINTER(C.c_uint64)]
    L.vrlb_vrls_count.argtypes = [vp]
    L.vrlb_vrls_count.restype = C.c_uint64
    L.vrlb_segments_build.argtypes = [vp, C.POINTER(abi.Camera), C.POINTER(abi.Options)]
    L.vrlb_segments_upload.argtypes = [vp, vp, C.c_uint64, C.c_int32, C.c_int32, C.c_int32]
    L.vrlb_segments_download.argtypes = [vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
    L.vrlb_segments_count.argtypes = [vp]
    L.vrlb_segments_count.restype = C.c_uint64
    L.vrlb_film_clear.argtypes = [vp]
    L.vrlb_gather.argtypes = [vp, C.POINTER(abi.Options)]
    L.vrlb_gather_segments.argtypes = [vp, C.POINTER(abi.Options), f32p]
    L.vrlb_film_device_ptr.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_uint64)]
    L.vrlb_film_download_xyzaw.argtypes = [vp, f32p]
    L.vrlb_film_download_rgb.argtypes = [vp, f32p]
    L.vrlb_render.argtypes = [vp, C.POINTER(abi.Camera), C.POINTER(abi.Options), C.POINTER(abi.VrlSetDesc), f32p]
    L.vrlb_cancel.argtypes = [vp]
    L.vrlb_get_timing.argtypes = [vp, C.POINTER(abi.Timing)]
    L.vrlb_synchronize.argtypes = [vp]
    L.vrlb_comm_unique_id.argtypes = [vp]
    L.vrlb_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
    L.vrlb_comm_destroy.argtypes = [vp]
    L.vrlb_film_allreduce.argtypes = [vp]
    L.vrlb_film_resolve.argtypes = [vp, C.POINTER(vp)]
    L.vrlb_measure_fp32_peak.argtypes = [vp, f32p]
    L.vrlb_vrls_generate.argtypes = [vp, C.POINTER(abi.GenerateDesc), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.vrlb_vrls_to_long.argtypes = [vp]
    L.vrlb_vrls_dice.argtypes = [vp, C.c_uint32, C.POINTER(C.c_uint64)]
    L.vrlb_vrls_dice_noalloc.argtypes = [C.c_int, vp, vp, C.c_uint64, C.c_float, C.c_uint32, vp, C.c_uint64, vp]
    L.vrlb_photons_upload.argtypes = [vp, C.POINTER(abi.PhotonSetDesc)]
    L.vrlb_photons_count.argtypes = [vp]
    L.vrlb_photons_generate.argtypes = [vp, C.POINTER(abi.GenerateDesc), C.c_uint64, C.c_float, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.vrlb_photons_download.argtypes = [vp, f32p, f32p, f32p, C.c_uint64, f32p]
    L.vrlb_photons_count.restype = C.c_uint64
    L.vrlb_photon_gather_segments.argtypes = [vp, C.POINTER(abi.Options), f32p]
    L.vrlb_vrls_scale.argtypes = [vp]
    L.vrlb_vrls_scale.restype = C.c_float
    L.vrlb_vrls_download_depth.argtypes = [vp, C.POINTER(C.c_int32), C.c_uint64]
    L.vrlb_surface_photons_upload.argtypes = [vp, C.c_int, C.POINTER(abi.PhotonSetDesc), C.c_int32]
    L.vrlb_surface_photons_generate.argtypes = [vp, C.c_int, C.POINTER(abi.GenerateDesc), C.c_uint64, C.c_float, C.c_int32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.vrlb_surface_photons_count.argtypes = [vp, C.c_int]
    L.vrlb_surface_photons_count.restype = C.c_uint64
    L.vrlb_surface_photons_download.argtypes = [vp, C.c_int, f32p, f32p, f32p, C.c_uint64, f32p]
    L.vrlb_surface_queries_upload.argtypes = [vp, vp, C.c_uint64, C.c_int32, C.c_int32]
    L.vrlb_surface_queries_download.argtypes = [vp, vp, C.c_uint64, C.POINTER(C.c_uint64)]
    L.vrlb_surface_gather_queries.argtypes = [vp, f32p]
    L.vrlb_lightcut_build.argtypes = [vp, C.c_uint64, C.c_uint64]
    L.vrlb_lightcut_download.argtypes = [vp, f32p, C.POINTER(C.c_int32), C.c_uint64, C.POINTER(C.c_uint64)]
    L.vrlb_lightcut_get_stats.argtypes = [vp, C.POINTER(abi.LightcutStats)]
    L.vrlb_lightcut_gather_segments.argtypes = [vp, C.POINTER(abi.Options), f32p, f32p]
    L.vrlb_set_bvh_builder.argtypes = [vp, C.c_int]
    L.vrlb_bvh_info.argtypes = [vp, C.POINTER(C.c_uint32)]
    L.vrlb_bvh_depths.argtypes = [C.POINTER(C.c_float), C.c_uint32, C.POINTER(C.c_uint32)]
    L.vrlb_shadow_proxies.argtypes = [C.POINTER(C.c_float), C.c_uint32, C.POINTER(C.c_float), C.c_uint32, C.POINTER(C.c_uint32)]
    _lib = L
    return L


class Context:
    """One GPU context (one per process/GPU).  Thin object wrapper over the C handle."""

    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        rc = self.lib.vrlb_create(device, C.byref(h))
        if rc != abi.OK:
            raise VrlbError(rc, self.lib.vrlb_last_error(None).decode())
        self.h = h
        self.device = device
        self._keep = []

    def close(self):
        if getattr(self, "h", None):
            self.lib.vrlb_destroy(self.h)
            self.h = None

    __del__ = close

    def _check(self, rc):
        if rc != abi.OK:
            raise VrlbError(rc, self.lib.vrlb_last_error(self.h).decode())

    def set_stream(self, cuda_stream_ptr):
        self._check(self.lib.vrlb_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def set_scene(self, scene):
        d = scene.desc()
        self._check(self.lib.vrlb_scene_set(self.h, C.byref(d)))

    def upload_vrls(self, vrl_desc):
        self._check(self.lib.vrlb_vrls_upload(self.h, C.byref(vrl_desc)))

    def download_vrls(self):
        n = int(self.lib.vrlb_vrls_count(self.h))
        soa = np.zeros((10, max(n, 1)), np.float32)
        med = np.zeros(max(n, 1), np.uint16)
        got = C.c_uint64()
        self._check(self.lib.vrlb_vrls_download(self.h, abi.fptr(soa), med.ctypes.data_as(C.POINTER(C.c_uint16)), n, C.byref(got)))
        return soa[:, :n], med[:n]

    def generate_vrls(self, gen):
        """Light tracing on the device (vrlb_vrls_generate); returns (n_vrls, n_particles)."""
        n, pcount = C.c_uint64(), C.c_uint64()
        self._check(self.lib.vrlb_vrls_generate(self.h, C.byref(gen), C.byref(n), C.byref(pcount)))
        return int(n.value), int(pcount.value)

    def vrls_to_long(self):
        """`long_vrl`: extend every resident VRL to the next surface (VRLMap::toLong)."""
        self._check(self.lib.vrlb_vrls_to_long(self.h))

    def dice_vrls(self, avg_number_dice):
        """`dice_vrl`: cut the resident VRLs into pieces of mean length / avg_number_dice (VRLMap::dicingVRL); returns the new count."""
        n = C.c_uint64()
        self._check(self.lib.vrlb_vrls_dice(self.h, int(avg_number_dice), C.byref(n)))
        return int(n.value)

    def upload_photons(self, photon_desc):
        """Volume photon set (scenes.photonset_desc); a resident set switches the direct in-scattering term on, n = 0 off."""
        self._check(self.lib.vrlb_photons_upload(self.h, C.byref(photon_desc)))

    def generate_photons(self, gen, capacity, lookup_radius):
        """Volume photons by light tracing on the device (vrlb_photons_generate); returns (n_photons, n_particles)."""
        n, pcount = C.c_uint64(), C.c_uint64()
        self._check(self.lib.vrlb_photons_generate(self.h, C.byref(gen), int(capacity), float(lookup_radius), C.byref(n), C.byref(pcount)))
        return int(n.value), int(pcount.value)

    def download_photons(self):
        """(pos, dir, power, scale) of the resident photon set, in grid order."""
        n = self.photons_count()
        pos, direction, power = (np.zeros((max(n, 1), 3), np.float32) for _ in range(3))
        scale = C.c_float()
        self._check(self.lib.vrlb_photons_download(self.h, abi.fptr(pos), abi.fptr(direction), abi.fptr(power), n, C.byref(scale)))
        return pos[:n], direction[:n], power[:n], float(scale.value)

    def clear_photons(self):
        d = abi.PhotonSetDesc()
        self._check(self.lib.vrlb_photons_upload(self.h, C.byref(d)))

    def photons_count(self):
        return int(self.lib.vrlb_photons_count(self.h))

    def photon_gather_segments(self, opt):
        """The direct term per resident camera segment (throughput included)."""
        n = int(self.lib.vrlb_segments_count(self.h))
        out = np.zeros((max(n, 1), 3), np.float32)
        self._check(self.lib.vrlb_photon_gather_segments(self.h, C.byref(opt), abi.fptr(out)))
        return out[:n]

    # ---- surface photon maps (vrl.cpp:671-680) ----
    def upload_surface_photons(self, which, photon_desc, lookup_size):
        self._check(self.lib.vrlb_surface_photons_upload(self.h, which, C.byref(photon_desc), lookup_size))

    def generate_surface_photons(self, gen, capacity, lookup_radius, lookup_size, which=abi.PHOTONMAP_GLOBAL):
        """Global / caustic surface photons by light tracing on the device; returns (n_photons, n_particles)."""
        n, pcount = C.c_uint64(), C.c_uint64()
        self._check(self.lib.vrlb_surface_photons_generate(self.h, which, C.byref(gen), int(capacity), float(lookup_radius), int(lookup_size),
                                                           C.byref(n), C.byref(pcount)))
        return int(n.value), int(pcount.value)

    def clear_surface_photons(self, which):
        self._check(self.lib.vrlb_surface_photons_upload(self.h, which, C.byref(abi.PhotonSetDesc()), 0))

    def surface_photons_count(self, which):
        return int(self.lib.vrlb_surface_photons_count(self.h, which))

    def download_surface_photons(self, which):
        n = self.surface_photons_count(which)
        pos, d, pw = (np.zeros((n, 3), np.float32) for _ in range(3))
        scale = C.c_float()
        self._check(self.lib.vrlb_surface_photons_download(self.h, which, abi.fptr(pos), abi.fptr(d), abi.fptr(pw), n, C.byref(scale)))
        return pos, d, pw, scale.value

    def upload_surface_queries(self, queries, width, height):
        q = np.ascontiguousarray(queries, abi.SURFACE_QUERY_DTYPE)
        self._check(self.lib.vrlb_surface_queries_upload(self.h, q.ctypes.data, len(q), width, height))

    def download_surface_queries(self):
        n = C.c_uint64()
        self._check(self.lib.vrlb_surface_queries_download(self.h, None, 0, C.byref(n)))
        q = np.zeros(n.value, abi.SURFACE_QUERY_DTYPE)
        self._check(self.lib.vrlb_surface_queries_download(self.h, q.ctypes.data, len(q), C.byref(n)))
        return q

    def surface_gather_queries(self):
        """caustic + global estimate per resident surface query, before the weight."""
        n = C.c_uint64()
        self._check(self.lib.vrlb_surface_queries_download(self.h, None, 0, C.byref(n)))
        out = np.zeros((n.value, 3), np.float32)
        self._check(self.lib.vrlb_surface_gather_queries(self.h, abi.fptr(out)))
        return out

    def vrls_scale(self):
        return float(self.lib.vrlb_vrls_scale(self.h))

    def download_vrl_depths(self):
        n = int(self.lib.vrlb_vrls_count(self.h))
        d = np.zeros(max(n, 1), np.int32)
        self._check(self.lib.vrlb_vrls_download_depth(self.h, d.ctypes.data_as(C.POINTER(C.c_int32)), n))
        return d[:n]

    def build_segments(self, cam, opt):
        self._check(self.lib.vrlb_segments_build(self.h, C.byref(cam), C.byref(opt)))
        return int(self.lib.vrlb_segments_count(self.h))

    def upload_segments(self, segs, width, height, spp=1):
        segs = np.ascontiguousarray(segs, abi.SEGMENT_DTYPE)
        self._check(self.lib.vrlb_segments_upload(self.h, segs.ctypes.data, len(segs), width, height, spp))

    def download_segments(self):
        n = int(self.lib.vrlb_segments_count(self.h))
        segs = np.zeros(max(n, 1), abi.SEGMENT_DTYPE)
        got = C.c_uint64()
        self._check(self.lib.vrlb_segments_download(self.h, segs.ctypes.data, n, C.byref(got)))
        return segs[:n]

    def clear_film(self):
        self._check(self.lib.vrlb_film_clear(self.h))

    def gather(self, opt):
        self._check(self.lib.vrlb_gather(self.h, C.byref(opt)))

    def set_bvh_builder(self, builder):
        """0 = host binned SAH (default), 1 = device LBVH, for the closest-hit tree of the next set_scene."""
        self._check(self.lib.vrlb_set_bvh_builder(self.h, builder))

    def bvh_info(self):
        info = (C.c_uint32 * 6)()
        self._check(self.lib.vrlb_bvh_info(self.h, info))
        return dict(builder=int(info[0]), depth=int(info[1]), nodes=int(info[2]), build_us=int(info[3]),
                    shadow_trees_on_device=int(info[4]), shadow_depth=int(info[5]))

    def lightcut_build(self, seed_offset=0, base_seed=0):
        """VRLLightCut constructor (vrl_lightcut.h:196-212) over the resident VRL set; the sampler is sampler.seed(base_seed, seed_offset)."""
        self._check(self.lib.vrlb_lightcut_build(self.h, base_seed, seed_offset))

    def lightcut_download(self):
        """The tree in preorder: (f [n,13] = flux, add_flux, lo, hi, max_len; i [n,6] = leaf, rep VRL, same child, n_leaves, id, parent id)."""
        n = C.c_uint64()
        self._check(self.lib.vrlb_lightcut_download(self.h, None, None, 0, C.byref(n)))
        f, i = np.zeros((n.value, 13), np.float32), np.zeros((n.value, 6), np.int32)
        self._check(self.lib.vrlb_lightcut_download(self.h, f.ctypes.data_as(C.POINTER(C.c_float)), i.ctypes.data_as(C.POINTER(C.c_int32)), n.value, C.byref(n)))
        return f, i

    def lightcut_stats(self):
        st = abi.LightcutStats()
        self._check(self.lib.vrlb_lightcut_get_stats(self.h, C.byref(st)))
        return st

    def lightcut_gather_segments(self, opt):
        """Li per resident segment through the cut + per-segment [fresh evaluations, refinements, largest heap, overflowed]."""
        n = self.lib.vrlb_segments_count(self.h)
        out, st = np.zeros((n, 3), np.float32), np.zeros((n, 4), np.float32)
        self._check(self.lib.vrlb_lightcut_gather_segments(self.h, C.byref(opt), out.ctypes.data_as(C.POINTER(C.c_float)),
                                                           st.ctypes.data_as(C.POINTER(C.c_float))))
        return out, st

    def gather_segments(self, opt):
        n = int(self.lib.vrlb_segments_count(self.h))
        out = np.zeros((max(n, 1), 3), np.float32)
        self._check(self.lib.vrlb_gather_segments(self.h, C.byref(opt), abi.fptr(out)))
        return out[:n]

    def film_device_ptr(self):
        p, n = C.c_void_p(), C.c_uint64()
        self._check(self.lib.vrlb_film_device_ptr(self.h, C.byref(p), C.byref(n)))
        return p.value, int(n.value)

    def film_xyzaw(self, width, height):
        out = np.zeros((height, width, 5), np.float32)
        self._check(self.lib.vrlb_film_download_xyzaw(self.h, abi.fptr(out)))
        return out

    def film_rgb(self, width, height):
        out = np.zeros((height, width, 3), np.float32)
        self._check(self.lib.vrlb_film_download_rgb(self.h, abi.fptr(out)))
        return out

    def render(self, cam, opt, vrl_desc=None, out=None):
        if out is None:
            out = np.zeros((cam.height, cam.width, 5), np.float32)
        self._check(self.lib.vrlb_render(self.h, C.byref(cam), C.byref(opt), C.byref(vrl_desc) if vrl_desc is not None else None,
                                         abi.fptr(out) if out is not False else None))
        return out

    # ---- multi-GPU film reduce (one Context per rank) ----
    @staticmethod
    def comm_unique_id():
        buf = C.create_string_buffer(abi.COMM_ID_BYTES)
        rc = load_library().vrlb_comm_unique_id(buf)
        if rc != abi.OK:
            raise VrlbError(rc, load_library().vrlb_last_error(None).decode())
        return buf.raw

    def comm_init(self, unique_id, rank, n_ranks):
        buf = C.create_string_buffer(bytes(unique_id), abi.COMM_ID_BYTES)
        self._check(self.lib.vrlb_comm_init(self.h, buf, rank, n_ranks))

    def comm_destroy(self):
        self._check(self.lib.vrlb_comm_destroy(self.h))

    def film_allreduce(self):
        self._check(self.lib.vrlb_film_allreduce(self.h))

    def film_resolve(self):
        p = C.c_void_p()
        self._check(self.lib.vrlb_film_resolve(self.h, C.byref(p)))
        return p.value

    def measure_fp32_peak(self):
        t = C.c_float()
        self._check(self.lib.vrlb_measure_fp32_peak(self.h, C.byref(t)))
        return float(t.value)

    def cancel(self):
        self._check(self.lib.vrlb_cancel(self.h))

    def synchronize(self):
        self._check(self.lib.vrlb_synchronize(self.h))

    def timing(self):
        t = abi.Timing()
        self._check(self.lib.vrlb_get_timing(self.h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in abi.Timing._fields_}

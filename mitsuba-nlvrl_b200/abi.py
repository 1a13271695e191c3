"""ctypes mirror of include/vrl_b200.h (the C ABI PODs).  Field order and types must match the header."""
import ctypes as C

COMM_ID_BYTES = 128
NOALLOC_MAX_VRLS = 65536
OK, ERR_INVALID, ERR_CUDA, ERR_STATE, ERR_CANCELLED, ERR_NOMEM, ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6

SURF_BSDF_SMOOTH, SURF_BSDF_NULL, SURF_EMITTER, SURF_TWOSIDED = 0x0, 0x1, 0x4, 0x8
SURF_BSDF_MIRROR, SURF_BSDF_DIELECTRIC = 0x2, 0x3
MEDIUM_HOMOGENEOUS, MEDIUM_GRID, MEDIUM_NONLINEAR = 0, 1, 2
PHASE_ISOTROPIC, PHASE_HG = 0, 1


def surf_interior(m): return (m + 1) << 8
def surf_exterior(m): return (m + 1) << 16
def surf_emitter_id(e): return e << 24


class Medium(C.Structure):
    _fields_ = [("kind", C.c_int32), ("phase", C.c_int32), ("g", C.c_float), ("scale", C.c_float),
                ("sigma_t", C.c_float * 3), ("albedo", C.c_float * 3),
                ("grid", C.POINTER(C.c_float)), ("grid_res", C.c_int32 * 3),
                ("grid_bbox_min", C.c_float * 3), ("grid_bbox_max", C.c_float * 3), ("grid_max", C.c_float),
                ("nl_res", C.c_int32 * 3), ("nl_bbox_min", C.c_float * 3), ("nl_bbox_max", C.c_float * 3),
                ("top_ior", C.c_float), ("bottom_ior", C.c_float), ("ior_method", C.c_int32), ("from_bottom", C.c_int32)]


class Emitter(C.Structure):
    _fields_ = [("radiance", C.c_float * 3), ("pad", C.c_float)]


class SceneDesc(C.Structure):
    _fields_ = [("tri_verts", C.POINTER(C.c_float)), ("tri_flags", C.POINTER(C.c_uint32)), ("n_tris", C.c_uint32),
                ("media", C.POINTER(Medium)), ("n_media", C.c_uint32),
                ("emitters", C.POINTER(Emitter)), ("n_emitters", C.c_uint32),
                ("tri_reflectance", C.POINTER(C.c_float))]


class Camera(C.Structure):
    _fields_ = [("sample_to_camera", C.c_float * 16), ("to_world", C.c_float * 16),
                ("near_clip", C.c_float), ("far_clip", C.c_float), ("width", C.c_int32), ("height", C.c_int32),
                ("medium", C.c_int32), ("pad", C.c_int32)]


class Options(C.Structure):
    _fields_ = [("spp", C.c_int32), ("block_size", C.c_int32), ("base_seed", C.c_uint64),
                ("samples_per_query", C.c_int32), ("use_uniform_sampling", C.c_int32), ("rr_vrl", C.c_int32),
                ("scale_rr", C.c_float), ("max_depth", C.c_int32), ("use_non_linear_camera", C.c_int32),
                ("use_nl_atomic_query", C.c_int32), ("visibility", C.c_int32), ("hide_emitters", C.c_int32),
                ("math_mode", C.c_int32), ("reserved", C.c_int32 * 2),
                ("use_light_cut", C.c_int32), ("threshold_error", C.c_float), ("threshold_better_dist", C.c_int32), ("lightcut_heap", C.c_int32)]


class VrlSetDesc(C.Structure):
    _fields_ = [("soa10", C.POINTER(C.c_float)), ("medium", C.POINTER(C.c_uint16)), ("n", C.c_uint64),
                ("first_index", C.c_uint64), ("total", C.c_uint64), ("scale", C.c_float), ("pad", C.c_float)]


class PhotonSetDesc(C.Structure):
    _fields_ = [("pos", C.POINTER(C.c_float)), ("dir", C.POINTER(C.c_float)), ("power", C.POINTER(C.c_float)), ("n", C.c_uint64),
                ("scale", C.c_float), ("lookup_radius", C.c_float)]


class GenerateDesc(C.Structure):
    _fields_ = [("target_vrls", C.c_uint64), ("seed", C.c_uint64), ("max_depth", C.c_int32), ("rr_depth", C.c_int32),
                ("min_vrl_length", C.c_float), ("use_non_linear", C.c_int32), ("use_laser", C.c_int32),
                ("laser_origin", C.c_float * 3), ("laser_direction", C.c_float * 3), ("laser_power", C.c_float * 3),
                ("start_medium", C.c_int32), ("max_steps", C.c_int32), ("reserved", C.c_int32 * 2)]


class Segment(C.Structure):
    _fields_ = [("o", C.c_float * 3), ("maxt", C.c_float), ("d", C.c_float * 3), ("pixel", C.c_uint32),
                ("weight", C.c_float * 3), ("medium", C.c_int32), ("rng_state", C.c_uint64),
                ("sample", C.c_uint32), ("part", C.c_uint32)]


class SurfaceQuery(C.Structure):
    _fields_ = [("p", C.c_float * 3), ("pixel", C.c_uint32), ("n", C.c_float * 3), ("flags", C.c_uint32),
                ("wi", C.c_float * 3), ("pad0", C.c_float), ("reflectance", C.c_float * 3), ("pad1", C.c_float),
                ("weight", C.c_float * 3), ("pad2", C.c_float)]


PHOTONMAP_GLOBAL, PHOTONMAP_CAUSTIC = 0, 1


class Timing(C.Structure):
    _fields_ = [("ms_upload", C.c_float), ("ms_segments", C.c_float), ("ms_gather", C.c_float), ("ms_film", C.c_float),
                ("ms_download", C.c_float), ("gather_launches", C.c_uint32), ("total_launches", C.c_uint32),
                ("n_evals", C.c_uint64), ("n_segments", C.c_uint64),
                ("launches_lifetime", C.c_uint64), ("ms_reduce", C.c_float), ("pad", C.c_float)]


class LightcutStats(C.Structure):
    _fields_ = [("n_nodes", C.c_uint64), ("build_searches", C.c_uint64), ("build_ms", C.c_float), ("pad", C.c_float),
                ("queries", C.c_uint64), ("evaluations", C.c_uint64), ("refinements", C.c_uint64), ("overflows", C.c_uint64),
                ("max_heap", C.c_uint64)]


assert C.sizeof(Segment) == 64

import numpy as _np
SEGMENT_DTYPE = _np.dtype([("o", "<f4", 3), ("maxt", "<f4"), ("d", "<f4", 3), ("pixel", "<u4"),
                           ("weight", "<f4", 3), ("medium", "<i4"), ("rng_state", "<u8"),
                           ("sample", "<u4"), ("part", "<u4")])
assert SEGMENT_DTYPE.itemsize == 64
SURFACE_QUERY_DTYPE = _np.dtype([("p", "<f4", 3), ("pixel", "<u4"), ("n", "<f4", 3), ("flags", "<u4"), ("wi", "<f4", 3), ("pad0", "<f4"),
                                 ("reflectance", "<f4", 3), ("pad1", "<f4"), ("weight", "<f4", 3), ("pad2", "<f4")])
assert SURFACE_QUERY_DTYPE.itemsize == C.sizeof(SurfaceQuery) == 80


def fptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))

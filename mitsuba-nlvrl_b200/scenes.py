"""Host-side scene description for the VRL gather: sensor matrices, triangle soups, media and seeded
synthetic VRL sets of the shape SURVEY.md 8(d) names ("box-fog").  Pure numpy; produces the PODs of
include/vrl_b200.h.  Matrix construction restates the reference's sensor set-up
(include/mitsuba/render/sensor.h:195-231, src/librender/sensor.cpp parse_fov,
include/mitsuba/core/transform.h look_at / perspective) in float64 and rounds once to fp32 -- the
oracle and the CUDA path both consume the resulting fp32 matrices, so parity does not depend on how
they were built."""
import ctypes as C
import math

import numpy as np

from . import abi


# --------------------------------------------------------------------------- sensor
def look_at(origin, target, up):
    """Transform::look_at (transform.h): columns left, new_up, dir, origin."""
    o, t, u = (np.asarray(v, np.float64) for v in (origin, target, up))
    d = (t - o) / np.linalg.norm(t - o)
    left = np.cross(u, d)
    left /= np.linalg.norm(left)
    new_up = np.cross(d, left)
    m = np.eye(4)
    m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = left, new_up, d, o
    return m


def _perspective(fov_deg, near, far):
    recip = 1.0 / (far - near)
    cot = 1.0 / math.tan(math.radians(fov_deg * 0.5))
    m = np.zeros((4, 4))
    m[0, 0] = cot
    m[1, 1] = cot
    m[2, 2] = far * recip
    m[2, 3] = -near * far * recip
    m[3, 2] = 1.0
    return m


def _scale(x, y, z):
    return np.diag([x, y, z, 1.0])


def _translate(x, y, z):
    m = np.eye(4)
    m[:3, 3] = (x, y, z)
    return m


def parse_fov(fov, fov_axis, aspect):
    """parse_fov (src/librender/sensor.cpp): returns the x field of view in degrees."""
    if fov_axis == "smaller":
        fov_axis = "y" if aspect > 1 else "x"
    elif fov_axis == "larger":
        fov_axis = "x" if aspect > 1 else "y"
    if fov_axis == "x":
        return fov
    if fov_axis == "y":
        return math.degrees(2 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
    if fov_axis == "diagonal":
        diag = 2 * math.tan(0.5 * math.radians(fov))
        width = diag / math.sqrt(1 + 1 / (aspect * aspect))
        return math.degrees(2 * math.atan(width * 0.5))
    raise ValueError("fov_axis must be x, y, diagonal, smaller or larger")


def make_camera(width, height, origin, target, up=(0, 1, 0), fov=39.3077, fov_axis="x",
                near_clip=1e-2, far_clip=1e4, medium=-1):
    aspect = width / float(height)
    x_fov = parse_fov(fov, fov_axis, aspect)
    # perspective_projection (sensor.h:195-231) with crop == film
    c2s = _scale(-0.5, -0.5 * aspect, 1.0) @ _translate(-1.0, -1.0 / aspect, 0.0) @ _perspective(x_fov, near_clip, far_clip)
    s2c = np.linalg.inv(c2s)
    cam = abi.Camera()
    cam.sample_to_camera[:] = s2c.astype(np.float32).reshape(-1).tolist()
    cam.to_world[:] = look_at(origin, target, up).astype(np.float32).reshape(-1).tolist()
    cam.near_clip, cam.far_clip = near_clip, far_clip
    cam.width, cam.height, cam.medium = width, height, medium
    return cam


# --------------------------------------------------------------------------- geometry helpers
def box_tris(lo, hi, faces="xXyYzZ", inward=False):
    """Axis-aligned box as triangles (outward normals unless inward). faces picks which sides."""
    lo, hi = np.asarray(lo, np.float32), np.asarray(hi, np.float32)
    c = [np.array([lo[0] if i & 1 == 0 else hi[0], lo[1] if i & 2 == 0 else hi[1], lo[2] if i & 4 == 0 else hi[2]], np.float32)
         for i in range(8)]
    quads = {"x": (0, 4, 6, 2), "X": (1, 3, 7, 5), "y": (0, 1, 5, 4), "Y": (2, 6, 7, 3), "z": (0, 2, 3, 1), "Z": (4, 5, 7, 6)}
    tris = []
    for f in faces:
        a, b, cc, d = quads[f]
        q = [(a, b, cc), (a, cc, d)]
        for t in q:
            t = t[::-1] if inward else t
            tris.append(np.stack([c[t[0]], c[t[1]], c[t[2]]]))
    return np.stack(tris).astype(np.float32)


def transform_tris(tris, m):
    m = np.asarray(m, np.float64)
    p = tris.reshape(-1, 3).astype(np.float64)
    p = p @ m[:3, :3].T + m[:3, 3]
    return p.reshape(-1, 3, 3).astype(np.float32)


def rotation(axis, deg):
    a = np.asarray(axis, np.float64)
    a /= np.linalg.norm(a)
    t = math.radians(deg)
    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
    m = np.eye(4)
    m[:3, :3] = np.eye(3) + math.sin(t) * K + (1 - math.cos(t)) * (K @ K)
    return m


def sphere_tris(center, radius, n_lat=20, n_lon=32):
    """UV sphere: 2*n_lon*(n_lat-1) triangles (20x32 -> 1216; SURVEY's 1280 is an icosphere; same order of size)."""
    c = np.asarray(center, np.float64)
    tris = []

    def p(i, j):
        th = math.pi * i / n_lat
        ph = 2 * math.pi * j / n_lon
        return c + radius * np.array([math.sin(th) * math.cos(ph), math.cos(th), math.sin(th) * math.sin(ph)])
    for i in range(n_lat):
        for j in range(n_lon):
            a, b, cc, d = p(i, j), p(i + 1, j), p(i + 1, j + 1), p(i, j + 1)
            if i != 0:
                tris.append(np.stack([a, b, d]))
            if i != n_lat - 1:
                tris.append(np.stack([b, cc, d]))
    return np.stack(tris).astype(np.float32)


# --------------------------------------------------------------------------- scene container
def make_medium(kind=abi.MEDIUM_HOMOGENEOUS, sigma_t=(1, 1, 1), albedo=(0.75, 0.75, 0.75), scale=1.0,
                phase=abi.PHASE_ISOTROPIC, g=0.8):
    m = abi.Medium()
    m.kind, m.phase, m.g, m.scale = kind, phase, g, scale
    m.sigma_t[:] = list(sigma_t)
    m.albedo[:] = list(albedo)
    return m


class Scene:
    """Owns the numpy buffers behind a vrlb_scene_desc."""

    def __init__(self):
        self._tris, self._flags, self.media, self.emitters, self._grids = [], [], [], [], []
        self._refl, self._any_refl = [], False

    def add_tris(self, tris, flags, reflectance=None):
        """`reflectance` (grey or RGB): diffuse reflectance of a smooth surface, read by the light tracer only
        (vrlb_scene_desc::tri_reflectance); None = 0, particles end on the surface."""
        tris = np.asarray(tris, np.float32).reshape(-1, 3, 3)
        self._tris.append(tris)
        self._flags.append(np.full(len(tris), flags, np.uint32))
        r = np.zeros(3, np.float32) if reflectance is None else np.broadcast_to(np.asarray(reflectance, np.float32), (3,))
        self._refl.append(np.tile(r, (len(tris), 1)))
        self._any_refl = self._any_refl or reflectance is not None
        return self

    def add_medium(self, medium, grid=None):
        if grid is not None:
            g = np.ascontiguousarray(grid, np.float32)   # [nz, ny, nx]
            self._grids.append(g)
            medium.grid = abi.fptr(g)
            medium.grid_res[:] = [g.shape[2], g.shape[1], g.shape[0]]
            medium.grid_max = float(g.max())
        self.media.append(medium)
        return len(self.media) - 1

    def add_emitter(self, radiance):
        e = abi.Emitter()
        e.radiance[:] = list(radiance)
        self.emitters.append(e)
        return len(self.emitters) - 1

    @property
    def tri_verts(self):
        return np.ascontiguousarray(np.concatenate(self._tris), np.float32) if self._tris else np.zeros((0, 3, 3), np.float32)

    @property
    def tri_flags(self):
        return np.ascontiguousarray(np.concatenate(self._flags), np.uint32) if self._flags else np.zeros(0, np.uint32)

    def desc(self):
        self._v, self._f = self.tri_verts, self.tri_flags
        self._m = (abi.Medium * max(len(self.media), 1))(*self.media)
        self._e = (abi.Emitter * max(len(self.emitters), 1))(*self.emitters)
        d = abi.SceneDesc()
        d.tri_verts = abi.fptr(self._v)
        d.tri_flags = self._f.ctypes.data_as(C.POINTER(C.c_uint32))
        d.n_tris = len(self._f)
        d.media, d.n_media = self._m, len(self.media)
        d.emitters, d.n_emitters = self._e, len(self.emitters)
        if self._any_refl:
            self._r = np.ascontiguousarray(np.concatenate(self._refl), np.float32)
            d.tri_reflectance = abi.fptr(self._r)
        return d


def default_options(**kw):
    o = abi.Options()
    o.spp, o.block_size, o.base_seed = 1, 32, 0
    o.samples_per_query, o.use_uniform_sampling, o.rr_vrl, o.scale_rr = 1, 0, 0, 0.5
    o.max_depth, o.use_non_linear_camera, o.use_nl_atomic_query = 512, 1, 0
    o.visibility, o.hide_emitters, o.math_mode = 1, 0, 1
    o.use_light_cut, o.threshold_error, o.threshold_better_dist, o.lightcut_heap = 0, 0.02, 8, 0      # vrl.cpp:66-72
    for k, v in kw.items():
        if not hasattr(o, k):
            raise AttributeError(k)
        setattr(o, k, v)
    return o


def generate_desc(target_vrls, seed=0, max_depth=512, rr_depth=5, min_vrl_length=5.0, use_non_linear=1, laser=None,
                  start_medium=-1, max_steps=0):
    """vrlb_generate_desc from the integrator properties of vrl.cpp:33-73; laser = (origin, direction, power) or None."""
    g = abi.GenerateDesc()
    g.target_vrls, g.seed, g.max_depth, g.rr_depth = target_vrls, seed, max_depth, rr_depth
    g.min_vrl_length, g.use_non_linear, g.start_medium, g.max_steps = min_vrl_length, use_non_linear, start_medium, max_steps
    if laser is not None:
        g.use_laser = 1
        g.laser_origin[:], g.laser_direction[:], g.laser_power[:] = list(laser[0]), list(laser[1]), list(laser[2])
    return g


def make_nonlinear_medium(lo, hi, res_y, bottom_ior=1.0, top_ior=0.95, **kw):
    """`nonlinear` medium (src/media/nonlinear.cpp:40-60): homogeneous coefficients + IOR linear in height over [lo, hi]."""
    m = make_medium(kind=abi.MEDIUM_NONLINEAR, **kw)
    m.nl_bbox_min[:], m.nl_bbox_max[:] = list(lo), list(hi)
    m.nl_res[:] = [1, res_y, 1]
    m.bottom_ior, m.top_ior, m.ior_method, m.from_bottom = bottom_ior, top_ior, 0, 1
    return m


# --------------------------------------------------------------------------- the synthetic "box-fog" scene
BOX = 10.0


def box_fog_scene(occluders=True, phase=abi.PHASE_ISOTROPIC, g=0.8, sigma_s=0.1, sigma_a=0.01, wall_reflectance=None):
    """Fog box [-10.01,10.01]^3 (null BSDF, interior medium 0) around five diffuse walls at +-10 (front
    face open towards the camera), optional sphere r=4 and rotated cuboid as occluders inside."""
    sc = Scene()
    st = sigma_s + sigma_a
    med = sc.add_medium(make_medium(sigma_t=(st,) * 3, albedo=(sigma_s / st,) * 3, scale=1.0, phase=phase, g=g))
    e = BOX + 0.01
    sc.add_tris(box_tris((-e,) * 3, (e,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(med))
    sc.add_tris(box_tris((-BOX,) * 3, (BOX,) * 3, faces="xXyYZ", inward=True), abi.SURF_BSDF_SMOOTH, reflectance=wall_reflectance)
    if occluders:
        sc.add_tris(sphere_tris((4.5, -2.0, -1.0), 4.0), abi.SURF_BSDF_SMOOTH)
        cub = box_tris((-2.0, -3.0, -1.5), (2.0, 3.0, 1.5))
        m = _translate(-4.0, -1.0, 2.0) @ rotation((0.3, 1.0, 0.2), 35.0)
        sc.add_tris(transform_tris(cub, m), abi.SURF_BSDF_SMOOTH)
    return sc


def specular_fog_scene(mirror=(0.9, 0.8, 0.7), eta=1.5, wall_reflectance=(0.7, 0.6, 0.5)):
    """box-fog with a mirror sphere (smooth conductor) and a glass cuboid (smooth dielectric, fog on both sides of its
    faces): camera paths carry on after them (specular chains, vrl.cpp:682-716)."""
    sc = Scene()
    med = sc.add_medium(make_medium(sigma_t=(0.11,) * 3, albedo=(0.1 / 0.11,) * 3, scale=1.0))
    e = BOX + 0.01
    sc.add_tris(box_tris((-e,) * 3, (e,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(med))
    sc.add_tris(box_tris((-BOX,) * 3, (BOX,) * 3, faces="xXyYZ", inward=True), abi.SURF_BSDF_SMOOTH, reflectance=wall_reflectance)
    sc.add_tris(sphere_tris((4.5, -2.0, -1.0), 4.0)[:, ::-1], abi.SURF_BSDF_MIRROR, reflectance=mirror)   # outward normals: a conductor is one-sided
    cub = box_tris((-2.0, -3.0, -1.5), (2.0, 3.0, 1.5))
    m = _translate(-4.0, -1.0, 2.0) @ rotation((0.3, 1.0, 0.2), 35.0)
    sc.add_tris(transform_tris(cub, m), abi.SURF_BSDF_DIELECTRIC, reflectance=(eta, 1.0, 1.0))
    return sc


def nl_fog_scene(occluders=True, res_y=320, bottom_ior=1.0, top_ior=0.95, sigma_s=0.1, sigma_a=0.01):
    """box-fog variant V4: the fog is a graded-IOR `nonlinear` medium, IOR(y) = lerp(bottom, top) over the
    box height in res_y layers (cbox-nonlinear-homo.xml uses 640), so camera paths are curved."""
    sc = Scene()
    st = sigma_s + sigma_a
    e = BOX + 0.01
    med = sc.add_medium(make_nonlinear_medium((-e,) * 3, (e,) * 3, res_y, bottom_ior, top_ior,
                                              sigma_t=(st,) * 3, albedo=(sigma_s / st,) * 3))
    sc.add_tris(box_tris((-e,) * 3, (e,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(med))
    sc.add_tris(box_tris((-BOX,) * 3, (BOX,) * 3, faces="xXyYZ", inward=True), abi.SURF_BSDF_SMOOTH)
    if occluders:
        sc.add_tris(sphere_tris((4.5, -2.0, -1.0), 4.0), abi.SURF_BSDF_SMOOTH)
    return sc


def smoke_grid(res=(64, 50, 64), seed=7, blobs=24):
    """Synthetic density grid [nz, ny, nx] in [0, 1] with the look of data/hetvol/smoke.vol (a few soft plumes over a
    thin background): the sum of seeded Gaussian blobs.  Deterministic (numpy PCG64 with a fixed seed)."""
    nx, ny, nz = res
    rng = np.random.Generator(np.random.PCG64(seed))
    z, y, x = np.meshgrid((np.arange(nz) + .5) / nz, (np.arange(ny) + .5) / ny, (np.arange(nx) + .5) / nx, indexing="ij")
    g = np.full((nz, ny, nx), 0.02)
    for _ in range(blobs):
        c = rng.uniform(0.15, 0.85, 3)
        s = rng.uniform(0.05, 0.18)
        a = rng.uniform(0.3, 1.0)
        g += a * np.exp(-((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) / (2 * s * s))
    return (g / g.max()).astype(np.float32)


def grid_fog_scene(occluders=True, res=(64, 50, 64), majorant=0.4, albedo=0.909, phase=abi.PHASE_HG, g=0.8):
    """box-fog variant V3: the fog is a grid (`heterogeneous`) medium; `scale` is chosen so that the
    majorant scale * max(grid) is `majorant` (a handful of free-flight steps per transmittance at this box size)."""
    sc = Scene()
    e = BOX + 0.01
    grid = smoke_grid(res)
    m = make_medium(kind=abi.MEDIUM_GRID, sigma_t=(1, 1, 1), albedo=(albedo,) * 3, scale=majorant / float(grid.max()), phase=phase, g=g)
    m.grid_bbox_min[:] = [-e] * 3
    m.grid_bbox_max[:] = [e] * 3
    med = sc.add_medium(m, grid=grid)
    sc.add_tris(box_tris((-e,) * 3, (e,) * 3), abi.SURF_BSDF_NULL | abi.surf_interior(med))
    sc.add_tris(box_tris((-BOX,) * 3, (BOX,) * 3, faces="xXyYZ", inward=True), abi.SURF_BSDF_SMOOTH)
    if occluders:
        sc.add_tris(sphere_tris((4.5, -2.0, -1.0), 4.0), abi.SURF_BSDF_SMOOTH)
    return sc


def box_fog_camera(width=1920, height=1080):
    return make_camera(width, height, origin=(0, 0, -40), target=(0, 0, -39), up=(0, 1, 0), fov=36.0, fov_axis="x",
                       near_clip=0.1, far_clip=1000.0)


def _seeded_floats(seed, n):
    """n deterministic uniforms in [0,1) from a counter-based splitmix64 (VRL sets are synthetic inputs: any
    seeded generator will do; parity only needs the oracle and the GPU to read the SAME exported set)."""
    with np.errstate(over="ignore"):
        x = np.arange(n, dtype=np.uint64) + np.uint64((int(seed) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)
        x = (x + np.uint64(0x9E3779B97F4A7C15))
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return ((z >> np.uint64(40)).astype(np.float64) / float(1 << 24)).astype(np.float32)


def synthetic_vrls(n, seed=1, sigma_t=0.11, albedo=0.909, box=BOX):
    """Seeded VRL set of SURVEY 8(d): origins uniform in the box, directions uniform on the sphere, lengths
    ~Exp(mean 1/sigma_t) clipped to the box, flux (1,1,1)*albedo^k with k~Geom(0.5).  Returns soa10 [10, n] fp32."""
    u = _seeded_floats(seed, 7 * n).reshape(7, n).astype(np.float64)
    o = (u[0:3] * 2 - 1) * (box * 0.999)
    z = 1 - 2 * u[3]
    r = np.sqrt(np.maximum(0, 1 - z * z))
    ph = 2 * np.pi * u[4]
    d = np.stack([r * np.cos(ph), r * np.sin(ph), z])
    d32 = d.astype(np.float32)
    d32 /= np.sqrt((d32 * d32).sum(0, dtype=np.float32))          # unit to fp32 rounding
    d = d32.astype(np.float64)
    length = -np.log(1 - u[5]) / sigma_t
    # clip to the box: largest t with o + t d inside
    with np.errstate(divide="ignore", invalid="ignore"):
        t1 = (box - o) / d
        t2 = (-box - o) / d
    tmax = np.where(d > 0, t1, np.where(d < 0, t2, np.inf)).min(0)
    length = np.minimum(length, 0.999 * tmax)
    k = np.floor(np.log2(1.0 / np.maximum(1 - u[6], 2.0 ** -24)))  # Geom(0.5)
    flux = albedo ** k
    soa = np.empty((10, n), np.float32)
    soa[0:3], soa[3:6], soa[6], soa[7:10] = o, d32, length, flux
    return np.ascontiguousarray(soa)


def vrlset_desc(soa10, scale=None, first_index=0, total=None, medium=None):
    """Builds a vrlb_vrlset_desc over a [10, n] fp32 array; the descriptor keeps the arrays alive."""
    assert soa10.dtype == np.float32 and soa10.flags["C_CONTIGUOUS"] and soa10.shape[0] == 10
    n = soa10.shape[1]
    d = abi.VrlSetDesc()
    d.soa10 = abi.fptr(soa10)
    d.medium = medium.ctypes.data_as(C.POINTER(C.c_uint16)) if medium is not None else None
    d.n, d.first_index, d.total = n, first_index, (total if total is not None else n)
    d.scale = (1.0 / d.total) if scale is None else scale
    d._keep = (soa10, medium)          # the C struct only holds raw pointers
    return d


def photonset_desc(pos, direction, power, scale, lookup_radius):
    """Builds a vrlb_photonset_desc over three [n, 3] fp32 arrays; the descriptor keeps them alive."""
    pos, direction, power = (np.ascontiguousarray(a, np.float32).reshape(-1, 3) for a in (pos, direction, power))
    assert len(pos) == len(direction) == len(power)
    d = abi.PhotonSetDesc()
    d.pos, d.dir, d.power = abi.fptr(pos), abi.fptr(direction), abi.fptr(power)
    d.n, d.scale, d.lookup_radius = len(pos), scale, lookup_radius
    d._keep = (pos, direction, power)
    return d


def synthetic_photons(n, seed=9, box=BOX, albedo=0.909):
    """Seeded volume photons for the box-fog scene: positions uniform in the box, unit directions, grey power."""
    u = _seeded_floats(seed, n * 7).reshape(n, 7)
    pos = ((u[:, 0:3] * 2.0 - 1.0) * box).astype(np.float32)
    z = u[:, 3] * 2.0 - 1.0
    phi = u[:, 4] * (2.0 * math.pi)
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    direction = np.stack([r * np.cos(phi), r * np.sin(phi), z], 1).astype(np.float32)
    power = (albedo * (0.5 + u[:, 5:6]) * np.array([[1.0, 0.9, 0.8]])).astype(np.float32)
    return pos, direction, power


def synthetic_surface_photons(n, seed=21, half=5.0, z=0.0):
    """Seeded surface photons on the plane z = const over [-half, half]^2: three in four arrive from above (direction
    z < 0), every fourth from below; RGB power."""
    u = _seeded_floats(seed, n * 6).reshape(n, 6)
    pos = np.stack([(u[:, 0] * 2.0 - 1.0) * half, (u[:, 1] * 2.0 - 1.0) * half, np.full(n, z)], 1).astype(np.float32)
    cz = -np.sqrt(u[:, 2])                                   # cosine-distributed arrival
    r = np.sqrt(np.maximum(0.0, 1.0 - cz * cz))
    phi = u[:, 3] * (2.0 * math.pi)
    cz[::4] *= -1.0
    direction = np.stack([r * np.cos(phi), r * np.sin(phi), cz], 1).astype(np.float32)
    power = ((0.5 + u[:, 4:5]) * np.array([[1.0, 0.8, 0.6]])).astype(np.float32)
    return pos, direction, power


def synthetic_surface_queries(n, width, height, seed=22, half=4.0, z=0.0, twosided_every=0):
    """Seeded surface queries on the same plane (normal +z): random camera-side directions, most above the plane."""
    u = _seeded_floats(seed, n * 8).reshape(n, 8)
    q = np.zeros(n, abi.SURFACE_QUERY_DTYPE)
    q["p"] = np.stack([(u[:, 0] * 2.0 - 1.0) * half, (u[:, 1] * 2.0 - 1.0) * half, np.full(n, z)], 1)
    q["n"] = (0.0, 0.0, 1.0)
    cz = u[:, 2] * 1.2 - 0.2                                 # one in six looks at the back of the surface
    r = np.sqrt(np.maximum(0.0, 1.0 - cz * cz))
    phi = u[:, 3] * (2.0 * math.pi)
    q["wi"] = np.stack([r * np.cos(phi), r * np.sin(phi), cz], 1)
    q["reflectance"] = 0.2 + 0.7 * u[:, 4:7]
    q["weight"] = (0.25 + 0.75 * u[:, 7:8]) * np.array([[1.0, 1.0, 1.0]])
    q["pixel"] = (np.arange(n) * 7919) % (width * height)
    if twosided_every:
        q["flags"][::twosided_every] = abi.SURF_TWOSIDED
    return q


# --------------------------------------------------------------------------- files (INTEGRATION.md "File formats")
def save_vrlb(path, soa10, scale=None, medium=None):
    """Exported VRL set: "VRLB" | u32 version | u64 n | f32 scale | u32 pad | 10*n f32 (SoA) | n u16 medium ids."""
    import struct
    n = soa10.shape[1]
    scale = (1.0 / n) if scale is None else scale
    med = np.zeros(n, np.uint16) if medium is None else np.ascontiguousarray(medium, np.uint16)
    with open(path, "wb") as f:
        f.write(b"VRLB" + struct.pack("<IQfI", 1, n, scale, 0))
        f.write(np.ascontiguousarray(soa10, np.float32).tobytes())
        f.write(med.tobytes())


def load_vrlb(path):
    import struct
    with open(path, "rb") as f:
        hdr = f.read(24)
        if hdr[:4] != b"VRLB":
            raise ValueError(f"{path}: not a VRLB file")
        ver, n, scale, _ = struct.unpack("<IQfI", hdr[4:])
        if ver != 1:
            raise ValueError(f"{path}: VRLB version {ver}")
        soa = np.frombuffer(f.read(40 * n), np.float32).reshape(10, n).copy()
        med = np.frombuffer(f.read(2 * n), np.uint16).copy()
    return soa, scale, med


def load_vol(path):
    """Mitsuba binary grid volume (src/textures/volume_data.h:42-107) -> (grid [nz, ny, nx(, c)], bbox_min, bbox_max)."""
    import struct
    with open(path, "rb") as f:
        hdr = f.read(4)
        if hdr[:3] != b"VOL":
            raise ValueError(f"Invalid volume file {path}")
        if hdr[3] != 3:
            raise ValueError(f"Invalid version, currently only version 3 is supported (found {hdr[3]})")
        typ, nx, ny, nz, ch = struct.unpack("<5i", f.read(20))
        if typ != 1:
            raise ValueError(f"Wrong type, currently only type == 1 (Float32) data is supported (found type = {typ})")
        bbox = struct.unpack("<6f", f.read(24))
        data = np.frombuffer(f.read(4 * nx * ny * nz * ch), "<f4")
    if data.size != nx * ny * nz * ch:
        raise ValueError(f"Unexpected end of volume file {path}")
    grid = data.reshape((nz, ny, nx, ch)) if ch > 1 else data.reshape((nz, ny, nx))
    return grid.astype(np.float32), np.array(bbox[:3], np.float32), np.array(bbox[3:], np.float32)


def save_vol(path, grid, bbox_min, bbox_max):
    import struct
    g = np.ascontiguousarray(grid, np.float32)
    ch = g.shape[3] if g.ndim == 4 else 1
    with open(path, "wb") as f:
        f.write(b"VOL\x03" + struct.pack("<5i", 1, g.shape[2], g.shape[1], g.shape[0], ch))
        f.write(struct.pack("<6f", *[float(x) for x in bbox_min], *[float(x) for x in bbox_max]))
        f.write(g.tobytes())


def load_obj(path, to_world=None):
    """Wavefront OBJ -> [n, 3, 3] triangles; polygons are fanned (v0, v_{k-1}, v_k) like src/shapes/obj.cpp:256-262."""
    pos, tris = [], []
    with open(path) as f:
        for line in f:
            s = line.split()
            if not s:
                continue
            if s[0] == "v":
                pos.append([float(s[1]), float(s[2]), float(s[3])])
            elif s[0] == "f":
                idx = [int(tok.split("/")[0]) for tok in s[1:]]
                if any(i < 1 or i > len(pos) for i in idx):
                    raise ValueError(f"{path}: reference to invalid vertex")
                for k in range(2, len(idx)):
                    tris.append([pos[idx[0] - 1], pos[idx[k - 1] - 1], pos[idx[k] - 1]])
    t = np.asarray(tris, np.float32).reshape(-1, 3, 3)
    return transform_tris(t, to_world) if to_world is not None else t


def save_scene_dump(path, scene, cam, spp=1, seed=0):
    """Test-harness scene dump read by host/vrl_host (vrl_host_main.cpp)."""
    import struct
    v, fl = scene.tri_verts, scene.tri_flags
    with open(path, "wb") as f:
        f.write(struct.pack("<I", len(fl)) + v.tobytes() + fl.tobytes())
        f.write(struct.pack("<I", len(scene.media)) + b"".join(bytes(m) for m in scene.media))
        f.write(struct.pack("<I", len(scene.emitters)) + b"".join(bytes(e) for e in scene.emitters))
        f.write(bytes(cam) + struct.pack("<iQ", spp, seed))


def load_scene_dump(path):
    """The scene dump `vrl_host dump-xml scene.xml out` writes (host/vrl_host_main.cpp): the scene as the C++ XML loader
    hands it to the plugin.  Returns (Scene, Camera, spp, sampler seed, emission medium)."""
    import ctypes as C
    import struct
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:8] != b"VRLD2\0\0\0":
        raise ValueError(f"{path}: not a VRLD2 scene dump")
    off = 8

    def take(n):
        nonlocal off
        b = raw[off:off + n]
        if len(b) != n:
            raise ValueError(f"{path}: truncated")
        off += n
        return b
    n_tris, = struct.unpack("<I", take(4))
    verts = np.frombuffer(take(n_tris * 36), np.float32).reshape(-1, 3, 3).copy()
    flags = np.frombuffer(take(n_tris * 4), np.uint32).copy()
    n_media, = struct.unpack("<I", take(4))
    media = [abi.Medium.from_buffer_copy(take(C.sizeof(abi.Medium))) for _ in range(n_media)]
    n_em, = struct.unpack("<I", take(4))
    emitters = [abi.Emitter.from_buffer_copy(take(C.sizeof(abi.Emitter))) for _ in range(n_em)]
    cam = abi.Camera.from_buffer_copy(take(C.sizeof(abi.Camera)))
    spp, seed = struct.unpack("<iQ", take(12))
    has_refl, = struct.unpack("<I", take(4))
    refl = np.frombuffer(take(n_tris * 12), np.float32).reshape(-1, 3).copy() if has_refl else None
    sc = Scene()
    for m in media:
        grid, gmax = None, m.grid_max
        if m.kind == abi.MEDIUM_GRID:
            nx, ny, nz = m.grid_res
            grid = np.frombuffer(take(nx * ny * nz * 4), np.float32).reshape(nz, ny, nx).copy()
        m.grid = None
        sc.add_medium(m, grid=grid)
        m.grid_max = gmax                       # the loader's majorant (m_sigmat->max()), not recomputed here
    for e in emitters:
        sc.add_emitter(tuple(e.radiance))
    sc._tris.append(verts)
    sc._flags.append(flags)
    sc._refl.append(refl if refl is not None else np.zeros((n_tris, 3), np.float32))
    sc._any_refl = refl is not None
    emission_medium, = struct.unpack("<i", take(4))
    return sc, cam, spp, seed, emission_medium

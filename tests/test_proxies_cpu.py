"""Host-side checks of the occluder proxies (mitsuba-nlvrl_b200/csrc/occluder_proxies.cpp) through the host-only
diagnostic vrlb_shadow_proxies: outer proxies contain every vertex, inner proxies lie strictly inside the closed
mesh (brute-force parity ray casting in numpy), open meshes get no inner proxy, shells are recognised."""
import numpy as np

from mitsuba_nlvrl_b200 import api, scenes


def _crossings(tris, p, d):
    """Moeller-Trumbore parity count of the ray p + t d, t > 0, in float64."""
    a, e1, e2 = tris[:, 0], tris[:, 1] - tris[:, 0], tris[:, 2] - tris[:, 0]
    pv = np.cross(d, e2)
    det = (e1 * pv).sum(1)
    ok = np.abs(det) > 1e-300
    inv = np.where(ok, 1.0 / np.where(ok, det, 1.0), 0.0)
    tv = p - a
    u = (tv * pv).sum(1) * inv
    qv = np.cross(tv, e1)
    v = (qv * d).sum(1) * inv
    t = (e2 * qv).sum(1) * inv
    return int((ok & (u >= 0) & (u <= 1) & (v >= 0) & (u + v <= 1) & (t > 0)).sum())


def _inside(tris, p):
    dirs = np.array([[0.5773, 0.3015, 0.7589], [-0.4242, 0.8111, 0.4028], [0.1237, -0.6419, -0.7567]])
    return all(_crossings(tris, p, d) % 2 == 1 for d in dirs)


def _component(tris, solid):
    c = tris.reshape(-1, 3, 3).astype(np.float64)
    inside_box = np.all((c >= solid["lo"] - 1e-6) & (c <= solid["hi"] + 1e-6), axis=(1, 2))
    return c[inside_box]


def test_box_fog_proxies_are_conservative():
    sc = scenes.box_fog_scene(occluders=True)
    solids, n_shells, n_main = api.shadow_proxies(sc.tri_verts)
    assert len(solids) == 2 and n_shells == 2 and n_main == 0      # sphere + cuboid; fog container + walls are shells
    rng = np.random.default_rng(5)
    for s in solids:
        tris = _component(np.asarray(sc.tri_verts, np.float32), s)
        assert len(tris) in (1216, 12)          # 20 x 32 lat-long sphere (fans at the poles), cuboid
        verts = tris.reshape(-1, 3)
        # outer proxies contain every vertex
        assert np.all(((verts - s["c"]) ** 2).sum(1) <= s["r_out2"])
        local = (verts - s["oc"]) @ s["ax"].astype(np.float64).T
        assert np.all(np.abs(local) <= s["h_out"] + 1e-6)
        assert np.all(verts >= s["lo"]) and np.all(verts <= s["hi"])
        # inner proxies: random points ON their surface are inside the closed mesh
        assert s["flags"] & 6                   # closed components get at least one inner proxy
        dirs = rng.normal(size=(40, 3)); dirs /= np.linalg.norm(dirs, axis=1)[:, None]
        for d in dirs:                          # (the inscribed sphere is computed even where its flag was dropped)
            assert _inside(tris, s["ci"] + d * np.sqrt(s["r_in2"]))
        if s["flags"] & 4:
            for _ in range(40):
                q = rng.uniform(-1, 1, 3); q[rng.integers(3)] = rng.choice([-1.0, 1.0])     # a point on the box surface
                assert _inside(tris, s["oc"] + (q * s["h_in"]) @ s["ax"].astype(np.float64))
    # the tessellated sphere: inscribed radius within 2 % of the true radius; the rotated cuboid: its own box
    sph = solids[np.argmax(solids["r_in2"])]
    assert 0.98 * 4.0 <= np.sqrt(sph["r_in2"]) < 4.0 <= np.sqrt(sph["r_out2"]) <= 1.01 * 4.0
    # flag policy (occluder_proxies.cpp, end of build): the sphere-like solid keeps only its spheres, the box-like solid
    # only its oriented boxes -- a warp pays for the longest proxy chain any of its lanes walks
    assert (sph["flags"] & 7) == 2
    cub = solids[np.argmin(solids["r_in2"])]
    assert (cub["flags"] & 7) == 5
    assert np.allclose(np.sort(cub["h_out"]), [1.5, 2.0, 3.0], atol=0.02) and np.allclose(np.sort(cub["h_in"]), [1.5, 2.0, 3.0], atol=0.02)
    assert np.all(np.sort(cub["h_in"]) < [1.5, 2.0, 3.0]) and np.all(np.sort(cub["h_out"]) > [1.5, 2.0, 3.0])


def test_open_mesh_gets_no_inner_proxy():
    """An open surface (a cube without its lid) must not claim to enclose anything."""
    tris = scenes.box_tris((-1, -1, -1), (1, 1, 1), faces="xXyzZ")
    rot = scenes.rotation((0.2, 1.0, 0.4), 25.0)
    tris = scenes.transform_tris(tris, rot)
    # 10 triangles are below the 12-triangle threshold for a solid: pad with a second open box sharing a vertex set
    more = scenes.transform_tris(scenes.box_tris((-1, 1, -1), (1, 3, 1), faces="xXzZ"), rot)
    allt = np.concatenate([np.asarray(tris, np.float32).reshape(-1, 9), np.asarray(more, np.float32).reshape(-1, 9)])
    solids, n_shells, n_main = api.shadow_proxies(allt)
    assert len(solids) == 1 and (solids[0]["flags"] & 6) == 0


def test_bvh_depth_is_bounded_for_adversarial_meshes():
    """A strip of triangles whose sizes form a geometric progression makes a binned-SAH build peel one triangle per level;
    the device traversals have fixed stacks (48 / 32 entries), so the host build must stay below its depth bound (28)."""
    import ctypes as C
    n = 400
    x = np.cumsum(1.5 ** (np.arange(n + 1) * 0.25)) * 1e-6          # 1e-6 .. ~1e11: sizes over 17 decades
    tris = np.zeros((n, 3, 3), np.float32)
    tris[:, 0, 0] = x[:-1]; tris[:, 1, 0] = x[1:]; tris[:, 2, 0] = x[:-1]
    tris[:, 2, 1] = (x[1:] - x[:-1])
    tris[:, :, 2] = np.linspace(0, 1, n)[:, None]                     # not coplanar: an ordinary component, not a shell
    lib = api.load_library()
    d = (C.c_uint32 * 2)()
    v = np.ascontiguousarray(tris.reshape(-1))
    assert lib.vrlb_bvh_depths(v.ctypes.data_as(C.POINTER(C.c_float)), n, d) == 0
    assert 8 <= d[0] <= 28 and d[1] <= 28, list(d)
    # an ordinary mesh stays shallow
    sph = np.ascontiguousarray(scenes.sphere_tris((0, 0, 0), 1.0).reshape(-1), np.float32)
    assert lib.vrlb_bvh_depths(sph.ctypes.data_as(C.POINTER(C.c_float)), len(sph) // 9, d) == 0
    assert d[0] <= 16 and d[1] <= 16

"""C++ host mirror of the `vrl` plugin (host/): property surface and error behaviour on the CPU, a frame through
VRLIntegrator::render on the GPU compared with the oracle."""
import os
import subprocess

import numpy as np
import pytest

import oracle
from mitsuba_nlvrl_b200 import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")


def _build():
    subprocess.check_call(["make", "-C", HOST], stdout=subprocess.DEVNULL)
    return os.path.join(HOST, "vrl_host")


def test_host_selftest_properties_and_errors():
    out = subprocess.run([_build(), "selftest"], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "SELFTEST PASSED" in out.stdout
    # the reference's loader rejects unknown properties (xml.cpp:1050-1068); FBD15-style names are unknown to this plugin
    assert 'unreferenced property "vrlTargetNum"' in out.stdout


def test_plugin_symbols():
    import ctypes as C
    _build()
    L = C.CDLL(os.path.join(HOST, "libvrl_plugin.so"))
    L.plugin_name.restype = C.c_char_p
    assert L.plugin_name() == b"VRLIntegrator"            # plugin.cpp:84-118 looks these two symbols up
    L.plugin_descr.restype = C.c_char_p
    assert L.plugin_descr()


def test_vrlb_file_round_trip(tmp_path):
    soa = scenes.synthetic_vrls(100, seed=4)
    p = str(tmp_path / "a.vrlb")
    scenes.save_vrlb(p, soa, scale=0.125)
    got, scale, med = scenes.load_vrlb(p)
    assert np.array_equal(got, soa) and scale == 0.125 and med.shape == (100,)


def test_vol_and_obj_loaders_agree_between_host_and_python(tmp_path):
    """`.vol` (volume_data.h:42-107) and OBJ (obj.cpp) readers: C++ host (vrl_io.cpp) against the Python mirror."""
    grid = scenes.smoke_grid((12, 7, 9))
    scenes.save_vol(tmp_path / "g.vol", grid, (-1, -2, -3), (1, 2, 3))
    g2, lo, hi = scenes.load_vol(tmp_path / "g.vol")
    assert np.array_equal(g2, grid) and list(lo) == [-1, -2, -3] and list(hi) == [1, 2, 3]
    out = subprocess.run([_build(), "vol-info", str(tmp_path / "g.vol")], capture_output=True, text=True).stdout.split()
    assert out[1:4] == ["12", "7", "9"] and out[5] == "1"
    assert abs(float(out[out.index("max") + 1]) - grid.max()) < 1e-6 and abs(float(out[out.index("mean") + 1]) - grid.astype(np.float64).mean()) < 1e-6
    (tmp_path / "q.obj").write_text("# cube face\nv 0 0 0\nv 2 0 0\nv 2 2 0\nv 0 2 0\nv 1 3 0\nvt 0 0\nf 1/1 2/1 3/1 4/1 5/1\n")
    tris = scenes.load_obj(tmp_path / "q.obj")
    assert tris.shape == (3, 3, 3) and np.array_equal(tris[2], [[0, 0, 0], [0, 2, 0], [1, 3, 0]])
    out = subprocess.run([_build(), "obj-info", str(tmp_path / "q.obj")], capture_output=True, text=True).stdout.split()
    assert out[1] == "3" and abs(float(out[out.index("checksum") + 1]) - tris.astype(np.float64).sum()) < 1e-6
    with pytest.raises(ValueError):
        (tmp_path / "bad.vol").write_bytes(b"VOL\x02" + b"\0" * 64)
        scenes.load_vol(tmp_path / "bad.vol")


REF_DATA = "/root/reference/data"


@pytest.mark.skipif(not os.path.exists(REF_DATA + "/hetvol/smoke.vol"), reason="reference fixtures only exist in the build container")
def test_loaders_read_the_reference_fixtures():
    """data/hetvol/smoke.vol is the 128 x 128 x 50 grid SURVEY 8d names; bounds.obj its bounding box."""
    grid, lo, hi = scenes.load_vol(REF_DATA + "/hetvol/smoke.vol")
    assert grid.shape == (50, 128, 128) and grid.max() == 1.0
    out = subprocess.run([_build(), "vol-info", REF_DATA + "/hetvol/smoke.vol"], capture_output=True, text=True).stdout.split()
    assert out[1:4] == ["128", "128", "50"] and abs(float(out[out.index("mean") + 1]) - grid.astype(np.float64).mean()) < 1e-8
    tris = scenes.load_obj(REF_DATA + "/hetvol/bounds.obj")
    assert tris.shape == (12, 3, 3) and np.allclose(tris.reshape(-1, 3).min(0), lo, atol=1e-5) and np.allclose(tris.reshape(-1, 3).max(0), hi, atol=1e-5)
    # the grid as a medium of the C ABI: majorant = scale * max (heterogeneous.cpp:29)
    m = scenes.make_medium(kind=scenes.abi.MEDIUM_GRID, scale=100.0)
    sc = scenes.Scene()
    m.grid_bbox_min[:], m.grid_bbox_max[:] = list(lo), list(hi)
    sc.add_medium(m, grid=grid)
    assert sc.media[0].grid_max == 1.0 and list(sc.media[0].grid_res) == [128, 128, 50]


def test_host_render_without_gpu_fails_loudly(tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    sc = scenes.box_fog_scene(occluders=False)
    cam = scenes.box_fog_camera(8, 4)
    scenes.save_scene_dump(str(tmp_path / "s.vrls"), sc, cam)
    scenes.save_vrlb(str(tmp_path / "v.vrlb"), scenes.synthetic_vrls(16, seed=1))
    out = subprocess.run([_build(), "render", str(tmp_path / "s.vrls"), str(tmp_path / "v.vrlb"), str(tmp_path / "o.f32")],
                         capture_output=True, text=True)
    assert out.returncode == 1 and "no CPU fallback" in out.stderr


@pytest.mark.gpu
def test_host_render_matches_oracle(tmp_path):
    sc = scenes.box_fog_scene(occluders=True)
    cam = scenes.box_fog_camera(48, 27)
    soa = scenes.synthetic_vrls(800, seed=6)
    scenes.save_scene_dump(str(tmp_path / "s.vrls"), sc, cam)
    scenes.save_vrlb(str(tmp_path / "v.vrlb"), soa)
    out = subprocess.run([_build(), "render", str(tmp_path / "s.vrls"), str(tmp_path / "v.vrlb"), str(tmp_path / "o.f32"),
                          "samples_per_query=1", "block_size=32"], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    film = np.fromfile(str(tmp_path / "o.f32"), np.float32).reshape(27, 48, 5)
    opt = scenes.default_options(samples_per_query=1)
    ref, _, _ = oracle.OracleScene(sc).render(cam, opt, scenes.vrlset_desc(soa))
    assert np.array_equal(film[..., 3:], ref[..., 3:])
    a, b = film[..., :3].astype(np.float64), ref[..., :3].astype(np.float64)
    err = np.abs(a - b) / np.maximum(np.abs(b), 1e-6 * np.abs(b).max())
    assert err.max() <= 2e-4, err.max()       # fp32-sequential oracle film vs compensated GPU sums


def _read_exr(path):
    """Minimal reader for the files write_exr produces: single-part, scan-line, uncompressed, FLOAT channels."""
    import struct
    b = open(path, "rb").read()
    assert struct.unpack_from("<ii", b, 0) == (20000630, 2)
    pos, attrs = 8, {}
    while b[pos] != 0:
        e = b.index(b"\0", pos); name = b[pos:e].decode(); pos = e + 1
        e = b.index(b"\0", pos); typ = b[pos:e].decode(); pos = e + 1
        n, = struct.unpack_from("<i", b, pos); pos += 4
        attrs[name] = (typ, b[pos:pos + n]); pos += n
    pos += 1
    chans, v, p = [], attrs["channels"][1], 0
    while v[p] != 0:
        e = v.index(b"\0", p); chans.append(v[p:e].decode()); p = e + 1
        ptype, _, xs, ys = struct.unpack_from("<i4sii", v, p); p += 16
        assert (ptype, xs, ys) == (2, 1, 1)
    x0, y0, x1, y1 = struct.unpack("<4i", attrs["dataWindow"][1])
    w, h = x1 - x0 + 1, y1 - y0 + 1
    assert attrs["compression"][1] == b"\0" and attrs["lineOrder"][1] == b"\0" and chans == sorted(chans)
    offsets = struct.unpack_from(f"<{h}Q", b, pos)
    img = np.zeros((h, w, len(chans)), np.float32)
    for row, off in enumerate(offsets):
        y, nbytes = struct.unpack_from("<ii", b, off)
        assert y == row and nbytes == w * len(chans) * 4
        img[row] = np.frombuffer(b, np.float32, w * len(chans), off + 8).reshape(len(chans), w).T
    assert offsets[-1] + 8 + w * len(chans) * 4 == len(b)
    return chans, img, attrs


def test_developed_film_as_exr_and_pfm(tmp_path):
    """HDRFilm::develop (hdrfilm.cpp:254-349): XYZ / W -> linear sRGB, alpha = A / W, written as an OpenEXR scan-line file
    (uncompressed float channels A, B, G, R) or a PFM; both from the same raw XYZAW film."""
    host = _build()
    rng = np.random.default_rng(5)
    w, h = 7, 5
    film = rng.uniform(0.1, 2.0, (h, w, 5)).astype(np.float32)
    film[..., 4] = rng.integers(1, 4, (h, w))
    film[0, 0] = 0                                                       # a pixel without samples develops to black
    raw = os.path.join(str(tmp_path), "film.f32")
    film.tofile(raw)
    exr, pfm = os.path.join(str(tmp_path), "film.exr"), os.path.join(str(tmp_path), "film.pfm")
    for out in (exr, pfm):
        subprocess.check_call([host, "convert-film", raw, str(w), str(h), out])
    chans, img, attrs = _read_exr(exr)
    assert chans == ["A", "B", "G", "R"] and img.shape == (h, w, 4)
    W = np.where(film[..., 4] > 0, film[..., 4], 1.0)
    xyz = film[..., :3] / W[..., None]
    m = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]], np.float32)
    rgb = xyz @ m.T
    assert np.allclose(img[..., 3], rgb[..., 0], rtol=1e-5, atol=1e-6) and np.allclose(img[..., 2], rgb[..., 1], rtol=1e-5, atol=1e-6)
    assert np.allclose(img[..., 1], rgb[..., 2], rtol=1e-5, atol=1e-6) and np.allclose(img[..., 0], film[..., 3] / W, rtol=1e-6)
    assert np.all(img[0, 0] == 0)
    assert struct_box(attrs["displayWindow"][1]) == (0, 0, w - 1, h - 1)
    # an independent decoder where the image has one: OpenCV's OpenEXR reader (BGRA)
    os.environ.setdefault("OPENCV_IO_ENABLE_OPENEXR", "1")
    try:
        import cv2
        dec = cv2.imread(exr, cv2.IMREAD_UNCHANGED)
    except Exception:
        dec = None
    if dec is not None:
        assert dec.shape == (h, w, 4) and dec.dtype == np.float32
        assert np.array_equal(dec[..., 2], img[..., 3]) and np.array_equal(dec[..., 1], img[..., 2])
        assert np.array_equal(dec[..., 0], img[..., 1]) and np.array_equal(dec[..., 3], img[..., 0])
    # component_format float16 (the film's default): HALF channels, round to nearest even like numpy's float16
    exr16 = os.path.join(str(tmp_path), "film16.exr")
    subprocess.check_call([host, "convert-film", raw, str(w), str(h), exr16, "float16"])
    b16 = open(exr16, "rb").read()
    assert len(b16) < len(open(exr, "rb").read())
    if dec is not None:
        dec16 = cv2.imread(exr16, cv2.IMREAD_UNCHANGED)
        assert np.array_equal(dec16, dec.astype(np.float16).astype(np.float32))
    # the PFM holds the same RGB, bottom row first
    b = open(pfm, "rb").read()
    head = b"PF\n%d %d\n-1.0\n" % (w, h)
    assert b.startswith(head)
    p = np.frombuffer(b, np.float32, w * h * 3, len(head)).reshape(h, w, 3)[::-1]
    assert np.array_equal(p[..., 0], img[..., 3]) and np.array_equal(p[..., 2], img[..., 1])
    # wrong size: loud
    r = subprocess.run([host, "convert-film", raw, str(w + 1), str(h), exr], capture_output=True, text=True)
    assert r.returncode != 0 and "cannot read" in r.stderr


def struct_box(b):
    import struct
    return struct.unpack("<4i", b)

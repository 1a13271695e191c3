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

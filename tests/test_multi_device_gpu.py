"""GPU suite, two or more devices in one box (skips on a single-GPU box): the VRL-sharded render IS the single-GPU render.

Two contexts in one process, one per device, each holding half of the VRL set with global stream addressing
(first_index / total); their films are added on the host.  The torchrun path (NCCL all-reduce of the films,
bench.py --gpus N `multi_gpu_check`) does the same sum on the devices; tests/test_multi_gpu_host.py covers its host logic
with gloo."""
import numpy as np
import pytest

from mitsuba_nlvrl_b200 import api, scenes, sharding
from parity_tools import REL_TOL, rel_err

pytestmark = pytest.mark.gpu


def _n_devices():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_n_devices() < 2, reason="needs two CUDA devices")
@pytest.mark.parametrize("math_mode", [0, 1])
def test_two_devices_add_up_to_one(math_mode):
    sc, cam = scenes.box_fog_scene(occluders=True), scenes.box_fog_camera(320, 180)
    opt = scenes.default_options(visibility=1, math_mode=math_mode)
    n = 8192
    soa = scenes.synthetic_vrls(n, seed=4)
    whole_ctx = api.Context(0)
    whole_ctx.set_scene(sc)
    whole_ctx.upload_vrls(scenes.vrlset_desc(soa))
    whole_ctx.build_segments(cam, opt)
    whole_ctx.clear_film(); whole_ctx.gather(opt)
    whole = whole_ctx.film_rgb(cam.width, cam.height).astype(np.float64)
    whole_ctx.close()
    total = np.zeros_like(whole)
    for rank in range(2):
        ctx = api.Context(rank)                      # device `rank`
        shard, vd = sharding.shard_desc(soa, rank, 2)
        ctx.set_scene(sc)
        ctx.upload_vrls(vd)
        ctx.build_segments(cam, opt)
        ctx.clear_film(); ctx.gather(opt)
        total += ctx.film_rgb(cam.width, cam.height)
        ctx.close()
    err = rel_err(total.reshape(-1, 3), whole.reshape(-1, 3))
    # shards are cut at a drain-window boundary (sharding.py): the exact kernel and the fast kernel both reproduce the whole
    # up to the fp32 rounding of adding two partial sums
    assert err.max() <= 1e-5, err.max()
    assert err.max() <= REL_TOL

"""vrlb_vrls_dice_noalloc (the entry point that allocates nothing: its scratch is a static __device__ array) against
vrlb_vrls_dice, through the C ABI, on two calls in a row.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import mitsuba_nlvrl_b200 as pkg

scenes, api, abi = pkg.scenes, pkg.api, pkg.abi

pytestmark = pytest.mark.gpu


def _dice_with_context(ctx, soa, k):
    ctx.upload_vrls(scenes.vrlset_desc(soa))
    n = ctx.dice_vrls(k)
    got, _ = ctx.download_vrls()
    assert got.shape[1] == n
    return got


def _dice_noalloc(lib, torch, soa, max_density, k, cap):
    n = soa.shape[1]
    d_in = torch.from_numpy(np.ascontiguousarray(soa, np.float32)).cuda()
    d_out = torch.full((10, cap), float("nan"), dtype=torch.float32, device="cuda")
    d_n = torch.full((2,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc = lib.vrlb_vrls_dice_noalloc(0, C.c_void_p(stream), C.c_void_p(d_in.data_ptr()), n, C.c_float(max_density), k,
                                    C.c_void_p(d_out.data_ptr()), cap, C.c_void_p(d_n.data_ptr()))
    assert rc == abi.OK, lib.vrlb_last_error(None)
    torch.cuda.synchronize()
    total, status = (int(v) for v in d_n.cpu())
    return d_out.cpu().numpy(), total, status


def test_dice_noalloc_matches_dice_twice_in_a_row():
    import torch
    lib = api.load_library()
    sc = scenes.box_fog_scene(occluders=False)
    md = float(np.float32(sc.media[0].scale) * np.float32((np.float32(sc.media[0].sigma_t[0]) + np.float32(sc.media[0].sigma_t[1])
                                                            + np.float32(sc.media[0].sigma_t[2])) / np.float32(3.0)))
    ctx = api.Context(0)
    ctx.set_scene(sc)
    # second call smaller than the first and with another dice count: stale scratch contents must not leak into it
    for n, seed, k in ((6000, 31, 4), (1500, 32, 3)):
        soa = scenes.synthetic_vrls(n, seed=seed)
        want = _dice_with_context(ctx, soa, k)
        cap = want.shape[1] + 17
        got, total, status = _dice_noalloc(lib, torch, soa, md, k, cap)
        assert status == 0 and total == want.shape[1]
        assert np.array_equal(got[0:7, :total], want[0:7])                       # origin, direction, length: bit-exact
        assert np.allclose(got[7:10, :total], want[7:10], rtol=2e-6, atol=0)     # flux: the dice tests' tolerance
        assert np.isnan(got[:, total:]).all()                                    # nothing written past the diced set
    # capacity too small: reported in the status word, nothing written
    soa = scenes.synthetic_vrls(500, seed=33)
    got, total, status = _dice_noalloc(lib, torch, soa, md, 4, 100)
    assert status == 2 and total > 100 and np.isnan(got).all()
    # beyond the static scratch: a clear error, no launch
    big = np.zeros((10, abi.NOALLOC_MAX_VRLS + 1), np.float32)
    d_big = torch.from_numpy(big).cuda()
    d_n = torch.zeros(2, dtype=torch.int64, device="cuda")
    rc = lib.vrlb_vrls_dice_noalloc(0, None, C.c_void_p(d_big.data_ptr()), big.shape[1], C.c_float(md), 2,
                                    C.c_void_p(d_big.data_ptr()), big.shape[1], C.c_void_p(d_n.data_ptr()))
    assert rc == abi.ERR_INVALID and b"VRLB_NOALLOC_MAX_VRLS" in lib.vrlb_last_error(None)
    ctx.close()

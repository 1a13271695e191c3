"""N > 1 host logic on CPU (gloo, world size 2): VRL shards with global stream addressing sum to the single-rank
film, direct emitter radiance is counted once, the communicator id reaches every rank."""
import os
import socket

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp
import torch

import oracle
from mitsuba_nlvrl_b200 import scenes, sharding


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = scenes.box_fog_scene(occluders=True)
        cam = scenes.box_fog_camera(24, 14)
        opt = scenes.default_options()
        soa = scenes.synthetic_vrls(301, seed=8)           # odd size: ragged shards
        cid = sharding.broadcast_comm_id(dist, lambda: bytes(range(128)), rank)
        assert cid == bytes(range(128))
        shard, vd = sharding.shard_desc(soa, rank, world)
        lo, hi = sharding.shard_bounds(soa.shape[1], rank, world)
        assert vd.first_index == lo and vd.n == hi - lo and vd.total == 301
        osc = oracle.OracleScene(sc)
        segs, direct, aw = osc.build_segments(cam, opt, soa.shape[1])
        _, li64, ev = osc.gather(vd, segs, opt)             # this rank's shard, global addressing
        film = np.zeros((cam.height * cam.width, 3))
        np.add.at(film, segs["pixel"], li64 * segs["weight"])
        if rank == 0:
            film += direct.reshape(-1, 3)                  # emitter radiance only on rank 0 (segments.cu add_direct)
        t = torch.from_numpy(film)
        dist.all_reduce(t)                                 # the NCCL film sum, here over gloo
        if rank == 0:
            whole = sharding.shard_desc(soa, 0, 1)[1]
            _, ref64, _ = osc.gather(whole, segs, opt)
            ref = np.zeros_like(film)
            np.add.at(ref, segs["pixel"], ref64 * segs["weight"])
            ref += direct.reshape(-1, 3)
            np.save(os.path.join(out_dir, "err.npy"), np.abs(t.numpy() - ref).max() / np.abs(ref).max())
    finally:
        dist.destroy_process_group()


def test_two_rank_vrl_shards_sum_to_whole(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert float(np.load(tmp_path / "err.npy")) < 1e-12


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 100000):
        for w in (1, 2, 3, 8):
            b = [sharding.shard_bounds(n, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(w - 1))

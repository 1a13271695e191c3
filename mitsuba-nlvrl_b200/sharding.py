"""Host-side multi-GPU protocol: contiguous VRL shards with GLOBAL stream addressing and the film sum.
Used by bench.py under torchrun; the arithmetic is testable on CPU with the gloo backend (tests/test_multi_gpu_host.py)."""
import numpy as np

from . import scenes


def shard_bounds(n_total, rank, world):
    """Rank r owns VRLs [r*N/W, (r+1)*N/W): contiguous, so eval (s, j) keeps its PCG32 offset (s*N + j)*dpe."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside [0, {world})")
    # cuts on multiples of 512 VRLs where the set is large enough: the visibility kernel drains its trace buffer at those
    # GLOBAL boundaries (gather_fast.cu DRAIN_WINDOW), so the shards' partial films are the whole set's partial sums
    def cut(r):
        c = r * n_total // world
        return c if (r == 0 or r == world or n_total < 512 * world * 4) else (c + 256) // 512 * 512
    return cut(rank), cut(rank + 1)


def shard_desc(soa10, rank, world):
    """(shard array kept alive by the caller, vrlb_vrlset_desc with first_index/total/scale of the WHOLE set)."""
    n = soa10.shape[1]
    lo, hi = shard_bounds(n, rank, world)
    shard = np.ascontiguousarray(soa10[:, lo:hi])
    return shard, scenes.vrlset_desc(shard, scale=1.0 / n, first_index=lo, total=n)


def broadcast_comm_id(dist, make_id, rank, src=0):
    """Rank `src` creates the NCCL unique id (vrlb_comm_unique_id), everyone receives the 128 bytes."""
    box = [make_id() if rank == src else None]
    dist.broadcast_object_list(box, src=src)
    return box[0]

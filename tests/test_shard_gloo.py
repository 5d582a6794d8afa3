"""Multi-GPU host logic on CPU: frame sharding + temporal block boundaries + result gather, world_size 2 and 3 over
gloo.  The per-batch processor is a stateful stand-in with the same contract as
b2orb_stereo_frontend (history = last pair of the previous call), so "sharded == sequential" is checked without a GPU."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.conftest import pkg

NFRAMES, H, W = 23, 6, 8


def load(i0, i1):
    idx = np.arange(i0, i1, dtype=np.int64)
    L = ((idx[:, None, None] * 7 + np.arange(H)[None, :, None] * 3 + np.arange(W)[None, None, :]) % 251).astype(np.uint8)
    return L, (L + 1).astype(np.uint8)


class FakeFrontend:
    """n_match depends on the frame alone; n_circ on the frame AND the previous processed frame (temporal state)."""

    def __init__(self):
        self.prev = None
        self.calls = 0

    def reset(self):
        self.prev = None

    def __call__(self, L, R, temporal):
        self.calls += 1
        n = len(L)
        sig = L.reshape(n, -1).astype(np.int64).sum(1) + 2 * R.reshape(n, -1).astype(np.int64).sum(1)
        circ = np.zeros(n, np.int64)
        for k in range(n):
            circ[k] = 0 if self.prev is None or not temporal else (sig[k] * 31 + self.prev) % 1009
            self.prev = int(sig[k])
        kp = np.zeros((n, 4), pkg("frontend").KP_DTYPE)
        kp["x"] = sig[:, None].astype(np.float32)
        return {"n_match": sig.astype(np.int32), "n_circ": circ.astype(np.int32), "kp": kp}


def sequential(nframes=NFRAMES):
    shard = pkg("shard")
    return shard.run_block(FakeFrontend(), load, nframes, 0, 1, batch=4, temporal=True)


def worker(rank, world, port, q, nframes=NFRAMES):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    shard = pkg("shard")
    fe = FakeFrontend()
    local = shard.run_block(fe, load, nframes, rank, world, batch=4, temporal=True, reset=fe.reset)
    out = shard.gather_blocks(local, rank, world, dist)
    if rank == 0:
        q.put({k: v.tobytes() for k, v in out.items()})
    dist.barrier()
    dist.destroy_process_group()


def test_frame_blocks_partition_the_sequence():
    shard = pkg("shard")
    for n in (0, 1, 7, 10000):
        for world in (1, 2, 3, 4, 8):
            blocks = [shard.frame_block(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 1
    assert shard.local_range(100, 1, 4, True) == (24, 25, 50)
    assert shard.local_range(100, 0, 4, True) == (0, 0, 25)
    assert shard.local_range(100, 1, 4, False) == (25, 25, 50)


@pytest.mark.parametrize("world,nframes", [(2, NFRAMES), (3, NFRAMES), (3, 2)], ids=["world2", "world3", "world3_more_ranks_than_frames"])
def test_sharded_equals_sequential_over_gloo(world, nframes):
    """The last case leaves a rank without frames: it must still run every collective of the gather (it used to skip
    them, and the other ranks hung)."""
    ref = sequential(nframes)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000 + world + 7 * nframes
    procs = [ctx.Process(target=worker, args=(r, world, port, q, nframes)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert set(got) == set(ref)
    for k in ref:
        assert got[k] == ref[k].tobytes(), k

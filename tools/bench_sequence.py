#!/usr/bin/env python3
"""SURVEY.md §8(d) config 4 / BASELINE.json configs[3]: ONE long stereo sequence (default 10 000 EuRoC-shaped pairs)
cut into contiguous frame blocks over the ranks (pose-estimation_b200/shard.py), temporal matching on, no collective
on the data path; per-frame match counts are gathered to rank 0 at the end.  Strong scaling: the job is fixed, the time
is the max over ranks (device-synchronised wall clock around the whole block, host buffers through the C ABI).

  python tools/bench_sequence.py --pairs 10000                       # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
      tools/bench_sequence.py --pairs 10000                           # 2 GPUs

The sequence cycles through a pool of distinct synthetic pairs (frame i = pool[i % pool]); with --check the sharded
result is compared with an unsharded run of the first --check frames on rank 0.
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=10000)
    ap.add_argument("--workload", default="euroc")
    ap.add_argument("--batch", type=int, default=512, help="pairs per C-ABI call (pipelined in 64-pair chunks inside the call)")
    ap.add_argument("--pool", type=int, default=512, help="distinct synthetic pairs (>= batch); frame i = pool[i % pool]")
    ap.add_argument("--check", type=int, default=0, help="verify the first N frames against an unsharded run (rank 0)")
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    synth = importlib.import_module("pose-estimation_b200.synth")
    fe = importlib.import_module("pose-estimation_b200.frontend")
    shard = importlib.import_module("pose-estimation_b200.shard")
    w, h, nf, ini, mn = synth.SHAPES[args.workload]
    assert args.pool >= args.batch
    ls, rs = synth.sequence(args.workload, args.pool, seed=5)
    # the pool twice in pinned memory: any window of <= pool consecutive frames is one contiguous view (no host copies
    # inside the timed job — the reference reads its frames from files, that cost is not what is measured here)
    pl = fe.pinned_empty((2 * args.pool, h, w)); pr = fe.pinned_empty((2 * args.pool, h, w))
    pl[: args.pool] = ls; pl[args.pool:] = ls; pr[: args.pool] = rs; pr[args.pool:] = rs
    hd = fe.Handle(nf, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn, w, h, max_images=2 * args.batch, device=local)
    out = hd.alloc_stereo_out(args.batch, pinned=True, full=False)

    def load(i0, i1):
        s = i0 % args.pool
        return pl[s: s + (i1 - i0)], pr[s: s + (i1 - i0)]

    def process(L, R, temporal):
        n = len(L)
        hd.check(hd.L.b2orb_stereo_frontend(hd.h, L.ctypes.data, R.ctypes.data, n, w, h, w, w * h, 1 if temporal else 0, out[1]))
        return {"n_l": out[0]["n_l"][:n], "n_match": out[0]["n_match"][:n], "n_circ": out[0]["n_circ"][:n]}

    # warm-up (allocations, lanes), then the timed job
    shard.run_block(process, load, min(args.pairs, 2 * args.batch), 0, 1, args.batch, True, hd.reset_history)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    res = shard.run_block(process, load, args.pairs, rank, world, args.batch, True, hd.reset_history)
    torch.cuda.synchronize()
    t = torch.tensor([time.perf_counter() - t0], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    allres = shard.gather_blocks(res, rank, world, dist if world > 1 else None, device="cuda")
    if rank == 0:
        line = {"metric": "sequence_pairs_per_s", "workload": f"{args.workload} {w}x{h}, {nf} feats, stereo + temporal",
                "pairs": args.pairs, "n_gpus": world, "seconds": float(t.item()), "value": args.pairs / float(t.item()),
                "scaling": "strong", "sharding": "contiguous frame blocks, no data-path collective, boundary frame re-processed",
                "frames_gathered": int(len(allres["n_match"])), "stereo_matches": int(allres["n_match"].sum()),
                "circular_matches": int(allres["n_circ"].sum())}
        if args.check:
            ref = shard.run_block(process, load, args.check, 0, 1, args.batch, True, hd.reset_history)
            line["sharded_equals_sequential"] = bool(all(np.array_equal(ref[k], allres[k][: args.check]) for k in ref))
        print(json.dumps(line))
    hd.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

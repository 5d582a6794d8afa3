"""Frame sharding of a stereo sequence over the GPUs of one box (BASELINE.json configs[3]).

The reference processes frames strictly sequentially on one thread (src/euroc-state-estimation.cpp:53-94); one
stereo pair's extraction + stereo matching depends on nothing but its two images, so frames are the data-parallel
unit.  One process per GPU (torch.distributed), contiguous frame blocks, NO collective on the data path.  The only
cross-frame dependency on the path is the temporal matching of externalTrack (src/feature-tracker.cpp:231-328), which
needs frame t-1's kept descriptors: at a block boundary the rank re-processes frame start-1 locally (cheapest, no
communication) and drops its results.  At the end the per-frame result slabs are gathered to rank 0.
"""
import numpy as np


def frame_block(nframes, rank, world):
    """Contiguous block [start, stop) of rank `rank`: GPU g gets frames [g*F/G, (g+1)*F/G)."""
    return nframes * rank // world, nframes * (rank + 1) // world


def local_range(nframes, rank, world, temporal):
    """(first, start, stop): frames [first, stop) are processed, results of [start, stop) are kept."""
    start, stop = frame_block(nframes, rank, world)
    first = start - 1 if (temporal and start > 0 and stop > start) else start
    return first, start, stop


def run_block(process, load, nframes, rank, world, batch, temporal=True, reset=None):
    """Drive `process(L, R, temporal) -> dict of per-frame arrays` over this rank's block in batches of `batch`.

    load(i0, i1) -> (L, R) uint8 arrays of frames [i0, i1).  `process` is stateful when temporal (history =
    last pair of the previous call, like b2orb_stereo_frontend); `reset()` clears that history first.
    Returns {name: array[stop-start, ...]} in frame order."""
    first, start, stop = local_range(nframes, rank, world, temporal)
    if reset is not None:
        reset()
    chunks = []
    i = first
    while i < stop:
        # the warm-up frame goes alone so that batches of kept frames stay aligned
        j = start if i < start else min(i + batch, stop)
        L, R = load(i, j)
        out = process(L, R, temporal)
        if i >= start:
            chunks.append({k: np.array(v[: j - i]) for k, v in out.items()})
        i = j
    if not chunks:
        return {}
    return {k: np.concatenate([c[k] for c in chunks]) for k in chunks[0]}


def gather_blocks(local, rank, world, dist=None, device="cpu"):
    """Gather per-frame slabs to rank 0 in frame order (blocks are contiguous, so rank order = frame order).
    Works on any torch.distributed backend (gloo on CPU, nccl with device='cuda'); world == 1 needs no dist."""
    if world == 1:
        return local
    import torch
    # every rank must run the same collectives, also one whose block is empty (world > nframes): take the slab names,
    # dtypes and trailing shapes from the first rank that has any
    mine = {k: (np.asarray(v).dtype.descr if np.asarray(v).dtype.names else np.asarray(v).dtype.str,
                tuple(np.asarray(v).shape[1:])) for k, v in local.items()} if local else None
    schemas = [None] * world
    dist.all_gather_object(schemas, mine)
    schema = next((s for s in schemas if s), None)
    if schema is None:
        return {} if rank == 0 else None
    names = sorted(schema)
    if not local:
        local = {k: np.zeros((0,) + tuple(shp), np.dtype([tuple(f) for f in dt]) if isinstance(dt, list) else np.dtype(dt))
                 for k, (dt, shp) in schema.items()}
    n_local = len(local[names[0]])
    counts = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([n_local], dtype=torch.int64, device=device))
    counts = [int(c.item()) for c in counts]
    cap = max(counts)
    out = {}
    for k in names:
        a = np.ascontiguousarray(local[k])
        row = int(np.prod(a.shape[1:], dtype=np.int64)) * (a.dtype.itemsize if a.dtype.names else 1)
        raw = a.view(np.uint8).reshape(n_local, row) if a.dtype.names else a.reshape(n_local, row)
        padded = np.zeros((cap, raw.shape[1]), raw.dtype)
        padded[:n_local] = raw
        t = torch.from_numpy(padded).to(device)
        bufs = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
        dist.gather(t, bufs, dst=0)
        if rank == 0:
            parts = [b.cpu().numpy()[:c] for b, c in zip(bufs, counts)]
            cat = np.concatenate(parts)
            if a.dtype.names:
                cat = cat.view(a.dtype).reshape((-1,) + a.shape[1:])
            else:
                cat = cat.reshape((-1,) + a.shape[1:])
            out[k] = cat
    return out if rank == 0 else None

#!/usr/bin/env python3
"""Aggregate pinned host<->device copy rate of the box with 1, 2, 4, 8 GPUs copying at the same time (one thread and one
stream per GPU, 64 MiB copies), one direction at a time and both directions concurrently.  Context for the multi-GPU e2e numbers of bench.py, whose inputs cross this path.
Usage: python tools/h2d_aggregate.py"""
import json
import threading
import time

import torch


def run_both(n, seconds=0.4, size=64 << 20, d2h_ratio=1.0):
    """H2D and D2H at the same time on every GPU (two streams per GPU); d2h_ratio scales the D2H copy size relative to
    the H2D one (the e2e leg of bench.py moves ~0.93 MB in per pair and 0.01-0.29 MB out).  Returns (h2d GB/s, d2h GB/s)."""
    bufs = []
    dsz = max(int(size * d2h_ratio), 1 << 16)
    for g in range(n):
        with torch.cuda.device(g):
            bufs.append((torch.empty(size, dtype=torch.uint8).pin_memory(), torch.empty(size, dtype=torch.uint8, device=f"cuda:{g}"),
                         torch.empty(dsz, dtype=torch.uint8).pin_memory(), torch.empty(dsz, dtype=torch.uint8, device=f"cuda:{g}"),
                         torch.cuda.Stream(device=g), torch.cuda.Stream(device=g)))
    cin, cout = [0] * n, [0] * n
    start = threading.Barrier(2 * n + 1)
    stop = [False]

    def worker(g, out):
        hi, di, ho, do, s_in, s_out = bufs[g]
        st = s_out if out else s_in
        with torch.cuda.device(g), torch.cuda.stream(st):
            start.wait()
            while not stop[0]:
                for _ in range(4):
                    if out:
                        ho.copy_(do, non_blocking=True)
                    else:
                        di.copy_(hi, non_blocking=True)
                st.synchronize()
                (cout if out else cin)[g] += 4
    ts = [threading.Thread(target=worker, args=(g, o)) for g in range(n) for o in (False, True)]
    for t in ts: t.start()
    start.wait()
    t0 = time.perf_counter()
    time.sleep(seconds)
    stop[0] = True
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    return sum(cin) * size / dt / 1e9, sum(cout) * dsz / dt / 1e9


def run(n, direction, seconds=0.4, size=64 << 20):
    bufs = []
    for g in range(n):
        with torch.cuda.device(g):
            h = torch.empty(size, dtype=torch.uint8).pin_memory()
            d = torch.empty(size, dtype=torch.uint8, device=f"cuda:{g}")
            bufs.append((h, d, torch.cuda.Stream(device=g)))
    counts = [0] * n
    start = threading.Barrier(n + 1)
    stop = [False]

    def worker(g):
        h, d, st = bufs[g]
        with torch.cuda.device(g), torch.cuda.stream(st):
            start.wait()
            while not stop[0]:
                for _ in range(4):
                    (d if direction == "h2d" else h).copy_(h if direction == "h2d" else d, non_blocking=True)
                st.synchronize()
                counts[g] += 4
    ts = [threading.Thread(target=worker, args=(g,)) for g in range(n)]
    for t in ts: t.start()
    start.wait()
    t0 = time.perf_counter()
    time.sleep(seconds)
    stop[0] = True
    for t in ts: t.join()
    dt = time.perf_counter() - t0
    return sum(counts) * size / dt / 1e9


def main():
    ng = torch.cuda.device_count()
    for n in (1, 2, 4, 8):
        if n > ng:
            break
        bi = run_both(n)
        bs = run_both(n, d2h_ratio=0.32)     # the full result set of bench.py's e2e leg: 0.29 MB out per 0.93 MB in
        print(json.dumps({"gpus": n, "h2d_gbs": round(run(n, "h2d"), 1), "d2h_gbs": round(run(n, "d2h"), 1),
                          "both_h2d_gbs": round(bi[0], 1), "both_d2h_gbs": round(bi[1], 1),
                          "both_0.32_h2d_gbs": round(bs[0], 1), "both_0.32_d2h_gbs": round(bs[1], 1)}), flush=True)


if __name__ == "__main__":
    main()

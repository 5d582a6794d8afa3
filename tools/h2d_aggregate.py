#!/usr/bin/env python3
"""Aggregate pinned host<->device copy rate of the box with 1, 2, 4, 8 GPUs copying at the same time (one thread and one
stream per GPU, 64 MiB copies).  Context for the multi-GPU e2e numbers of bench.py, whose inputs cross this path.
Usage: python tools/h2d_aggregate.py"""
import json
import threading
import time

import torch


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
        print(json.dumps({"gpus": n, "h2d_gbs": round(run(n, "h2d"), 1), "d2h_gbs": round(run(n, "d2h"), 1)}), flush=True)


if __name__ == "__main__":
    main()

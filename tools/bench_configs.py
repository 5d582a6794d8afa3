#!/usr/bin/env python3
"""SURVEY.md §8(d) configurations beside the headline bench line: for every shape (EuRoC, KITTI, kitti.yml, cfsd, 4K)
  * single-frame latency through the host C ABI (one stereo pair per b2orb_stereo_frontend call, pinned host
    images in, host results out): p50 / p95 / max over N calls — config 3 asks whether the 5 ms period of the
    200 Hz cfsd stream is met;
  * throughput of the same call with a batch of pairs (host buffers, copies inside);
  * parity of the first pair against the CPU oracle (checker only, through tests/parity_configs.py).
One JSON object per shape on stdout.  Needs a GPU.  Usage: python tools/bench_configs.py [--calls 200] [--batch 64]
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
    ap.add_argument("--calls", type=int, default=200)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--shapes", default="euroc,kitti,kitti_yml,cfsd,4k")
    ap.add_argument("--no-parity", action="store_true")
    args = ap.parse_args()
    synth = importlib.import_module("pose-estimation_b200.synth")
    fe = importlib.import_module("pose-estimation_b200.frontend")
    for shape in args.shapes.split(","):
        w, h, nf, ini, mn = synth.SHAPES[shape]
        temporal = shape.startswith("kitti")
        batch = args.batch if shape != "4k" else max(args.batch // 8, 2)
        npool = max(batch, 8)
        ls, rs = synth.sequence(shape, npool, seed=11)
        pl = fe.pinned_empty(ls.shape); pl[...] = ls
        pr = fe.pinned_empty(rs.shape); pr[...] = rs
        hd = fe.Handle(nf, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn, w, h, max_images=2 * batch)
        out1 = hd.alloc_stereo_out(1, pinned=True)
        img = w * h
        # ---- latency: one pair per call ----
        lat = []
        for k in range(args.calls + 10):
            i = k % npool
            t0 = time.perf_counter()
            hd.check(hd.L.b2orb_stereo_frontend(hd.h, pl[i:].ctypes.data, pr[i:].ctypes.data, 1, w, h, w, img,
                                                1 if temporal else 0, out1[1]))
            lat.append((time.perf_counter() - t0) * 1e3)
        lat = np.sort(np.array(lat[10:]))
        dev_ms = hd.last_device_ms
        # ---- throughput: a batch per call ----
        outb = hd.alloc_stereo_out(batch, pinned=True)
        reps = max(2, min(20, 2000 // batch))
        for k in range(2):
            hd.check(hd.L.b2orb_stereo_frontend(hd.h, pl.ctypes.data, pr.ctypes.data, batch, w, h, w, img,
                                                1 if temporal else 0, outb[1]))
        t0 = time.perf_counter()
        for k in range(reps):
            hd.check(hd.L.b2orb_stereo_frontend(hd.h, pl.ctypes.data, pr.ctypes.data, batch, w, h, w, img,
                                                1 if temporal else 0, outb[1]))
        dt = time.perf_counter() - t0
        line = {"shape": shape, "width": w, "height": h, "nfeatures": nf, "temporal": temporal,
                "latency_ms": {"p50": round(float(lat[len(lat) // 2]), 4), "p95": round(float(lat[int(len(lat) * 0.95)]), 4),
                               "max": round(float(lat[-1]), 4), "calls": int(len(lat)), "device_ms_last_call": round(dev_ms, 4)},
                "batch_pairs_per_s": round(batch * reps / dt, 1), "batch": batch,
                "keypoints_left_pair0": int(outb[0]["n_l"][0]), "stereo_matches_pair0": int(outb[0]["n_match"][0])}
        if shape == "cfsd":
            line["deadline_5ms_met"] = bool(lat[int(len(lat) * 0.95)] < 5.0)
        if not args.no_parity:
            from tests.parity_configs import stereo_pair_bit_exact   # the oracle stays behind tests/
            line["bit_exact_vs_oracle_pair0"] = stereo_pair_bit_exact((nf, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn),
                                                                      ls[0], rs[0], outb[0])
        print(json.dumps(line), flush=True)
        hd.close()
    if "cfsd" in args.shapes.split(","):
        cfsd_live(args, synth, fe)


def cfsd_live(args, synth, fe):
    """Config 3 as the reference runs it (src/cfsd-state-estimation.cpp:117-127 + process()): one 2560x720 CV_8UC4
    frame per call -> fused resize + grey + split (f2) -> rectification (f1) -> stereo front end; host frame in, host
    results out.  p50 / p95 latency against the 5 ms period of the 200 Hz stream, CPU port of the ingest beside it."""
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    sw, sh = 2560, 720
    ls, rs = synth.sequence("cfsd", 8, seed=11)
    frames = fe.pinned_empty((8, sh, sw, 4))
    for i in range(8):
        frames[i] = synth.packed_frame(ls[i], rs[i], sw, sh, seed=i)
    hd = fe.Handle(nf, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn, w, h, max_images=2)
    maps = [synth.rectification_maps(w, h, e) for e in (0, 1)]
    hd.set_rectification(maps[0][0], maps[0][1], maps[1][0], maps[1][1])
    out1 = hd.alloc_stereo_out(1, pinned=True)
    lat = []
    for k in range(args.calls + 10):
        i = k % 8
        t0 = time.perf_counter()
        hd.check(hd.L.b2orb_stereo_frontend_packed(hd.h, frames[i:].ctypes.data, 1, sw, sh, sw * 4, sw * sh * 4, w, h, 0, out1[1]))
        lat.append((time.perf_counter() - t0) * 1e3)
    lat = np.sort(np.array(lat[10:]))
    line = {"shape": "cfsd_live", "frame": [sw, sh, 4], "eye": [w, h], "nfeatures": nf, "rectify": True,
            "latency_ms": {"p50": round(float(lat[len(lat) // 2]), 4), "p95": round(float(lat[int(len(lat) * 0.95)]), 4),
                           "max": round(float(lat[-1]), 4), "calls": int(len(lat)), "device_ms_last_call": round(hd.last_device_ms, 4)},
            "h2d_bytes_per_frame": sw * sh * 4, "deadline_5ms_met": bool(lat[int(len(lat) * 0.95)] < 5.0),
            "keypoints_left": int(out1[0]["n_l"][0])}
    if not args.no_parity:
        from tests.parity_configs import cfsd_live_reference
        kl, dl, ms = cfsd_live_reference((nf, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn), frames[0], w, h, maps)
        line["cpu_port_ingest_rectify_ms"] = round(ms, 2)
        hd.check(hd.L.b2orb_stereo_frontend_packed(hd.h, frames.ctypes.data, 1, sw, sh, sw * 4, sw * sh * 4, w, h, 0, out1[1]))
        n = int(out1[0]["n_l"][0])
        line["bit_exact_vs_oracle"] = bool(n == len(kl) and np.array_equal(out1[0]["desc_l"][0, :n], dl))
    print(json.dumps(line), flush=True)
    hd.close()


if __name__ == "__main__":
    main()

#!/usr/bin/env python3
"""Where does the single-pair latency go?  One KITTI/EuRoC/cfsd pair per b2orb_stereo_frontend call with (a) every
output array, (b) counts + kept-match lists only; prints p50 wall clock and the device time (upload + kernels).
Needs a GPU.  Usage: python tools/latency_probe.py [shape ...]"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    synth = importlib.import_module("pose-estimation_b200.synth")
    fe = importlib.import_module("pose-estimation_b200.frontend")
    for shape in (sys.argv[1:] or ["kitti", "euroc", "cfsd"]):
        w, h, nf, ini, mn = synth.SHAPES[shape]
        temporal = shape.startswith("kitti")
        ls, rs = synth.sequence(shape, 8, seed=11)
        pl = fe.pinned_empty(ls.shape); pl[...] = ls
        pr = fe.pinned_empty(rs.shape); pr[...] = rs
        hd = fe.Handle(nf, synth.SCALE_FACTOR, synth.N_LEVELS, ini, mn, w, h, max_images=2)
        for label, full in (("all outputs", True), ("counts + match lists", False)):
            out = hd.alloc_stereo_out(1, pinned=True, full=full)
            lat = []
            for k in range(310):
                i = k % 8
                t0 = time.perf_counter()
                hd.check(hd.L.b2orb_stereo_frontend(hd.h, pl[i:].ctypes.data, pr[i:].ctypes.data, 1, w, h, w, w * h,
                                                    1 if temporal else 0, out[1]))
                lat.append((time.perf_counter() - t0) * 1e3)
            lat = np.sort(np.array(lat[10:]))
            print(f"{shape:6s} {label:22s} p50 {lat[len(lat) // 2]:.4f} ms  p95 {lat[int(len(lat) * .95)]:.4f}  "
                  f"device (upload + kernels) {hd.last_device_ms:.4f} ms", flush=True)
        hd.profile_enable(True)
        for k in range(20):
            hd.check(hd.L.b2orb_stereo_frontend(hd.h, pl[k % 8:].ctypes.data, pr[k % 8:].ctypes.data, 1, w, h, w, w * h,
                                                1 if temporal else 0, out[1]))
        prof = hd.profile_read()
        hd.profile_enable(False)
        print("       per-kernel us per call (events between launches, no launch overlap): " +
              " ".join(f"{k}={v[0] * 1e3 / 20:.1f}({v[1] // 20})" for k, v in prof.items() if v[1]), flush=True)
        hd.close()


if __name__ == "__main__":
    main()

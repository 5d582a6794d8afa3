#!/usr/bin/env python3
"""Kernel list of ONE single-pair call (run under ncu --metrics gpu__time_duration.sum): 3 warm-up calls, then one.
Usage: ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/latency_kernels.py [shape]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
fe = importlib.import_module("pose-estimation_b200.frontend"); synth = importlib.import_module("pose-estimation_b200.synth")
shape = sys.argv[1] if len(sys.argv) > 1 else "kitti"
w, h, nf, ini, mn = synth.SHAPES[shape]
ls, rs = synth.sequence(shape, 4, seed=11)
pl = fe.pinned_empty(ls.shape); pl[...] = ls
pr = fe.pinned_empty(rs.shape); pr[...] = rs
hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
out = hd.alloc_stereo_out(1, pinned=True)
for i in range(4):
    hd.check(hd.L.b2orb_stereo_frontend(hd.h, pl[i:].ctypes.data, pr[i:].ctypes.data, 1, w, h, w, w * h, 1, out[1]))
print("device ms of the last call (not under ncu: meaningless here)", hd.last_device_ms)

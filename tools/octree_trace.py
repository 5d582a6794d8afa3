"""Phase timeline of the quadtree kernel for (image 0, level 0): build csrc with -DOT_TRACE (make NVEXTRA=-DOT_TRACE)."""
import ctypes as C, importlib, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import os as _os
fe = importlib.import_module("pose-estimation_b200.frontend"); synth = importlib.import_module("pose-estimation_b200.synth")
shape = sys.argv[1] if len(sys.argv) > 1 else "kitti"
w, h, nf, ini, mn = synth.SHAPES[shape]
ls, rs = synth.sequence(shape, 1, seed=0)
hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
for _ in range(3):
    hd.stereo_frontend(ls, rs)
t = np.zeros(64, np.int64)
hd.L.b2orb_debug_octree_trace(t.ctypes.data_as(C.c_void_p))
ghz = 1.965
names = ["start", "histogram", "pyramid+roots", "rounds", "best pass", "decode"]
print(shape, "M=%d lsz=%d use_pyr=%d D=%d" % (t[10], t[11], t[12], t[13]))
for i in range(1, 6):
    print("  %-14s %8.2f us" % (names[i], (t[i] - t[i - 1]) / ghz / 1e3))
r = [x for x in t[20:60] if x]
for i in range(1, len(r)):
    print("  round %d        %8.2f us" % (i, (r[i] - r[i - 1]) / ghz / 1e3))
if r: print("  last round     %8.2f us" % ((t[3] - r[-1]) / ghz / 1e3))

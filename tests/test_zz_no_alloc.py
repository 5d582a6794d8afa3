"""b2orb_distribute_static — the allocation-free quadtree entry point (no counterpart in the reference):
its workspace is a file-scope device array, the call allocates nothing.

Checked through the C ABI against the existing entry point: b2orb_extract's keypoints of one pyramid level must be
exactly what b2orb_distribute_static selects from that level's candidates (b2orb_debug_get_candidates) — same points,
same order, bit for bit (the tolerance every quadtree test here uses: none).  Two calls in a row on different inputs:
the second runs on the workspace the first one left behind.  Collected last on purpose (file name), so a failure here
cannot hide another test under `pytest -x`; runs alone in a fresh process as well as after the other tests."""
import ctypes as C

import numpy as np
import pytest

from tests.conftest import pkg

pytestmark = pytest.mark.gpu


def _bind(fe):
    L = fe.lib()
    L.b2orb_distribute_static.restype = C.c_int
    L.b2orb_distribute_static.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    return L


def _packed_candidates(fe, hd, level):
    n = C.c_int(0)
    hd.check(hd.L.b2orb_debug_get_candidates(hd.h, 0, level, None, 0, C.byref(n)))
    out = np.zeros(max(n.value, 1), np.uint32)
    hd.check(hd.L.b2orb_debug_get_candidates(hd.h, 0, level, out.ctypes.data_as(C.c_void_p), n.value, C.byref(n)))
    return out[:n.value].copy()


def test_static_quadtree_equals_extract_on_two_calls_in_a_row(synth):
    import torch
    fe = pkg("frontend")
    L = _bind(fe)
    w, h, nf, ini, mn = synth.SHAPES["kitti"]
    ls, rs = synth.sequence("kitti", 2, seed=5)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    scale = hd.tables()[0]
    cap = hd.keypoint_capacity
    d_sel = torch.zeros(cap, dtype=torch.int32, device="cuda")
    d_n = torch.zeros(1, dtype=torch.int32, device="cuda")
    checked = 0
    # call 1: left image, level 0 (count-pyramid path, ~18k candidates); call 2: right image of the next pair, level 3
    # (other geometry, fewer points) on the workspace call 1 left behind; then level 0 again
    for img, level in ((ls[0], 0), (rs[1], 3), (rs[1], 0)):
        kps, _ = hd.extract(img)                              # existing entry point
        want = kps[kps["octave"] == level]
        cand = _packed_candidates(fe, hd, level)
        assert len(cand) > 4 * len(want) > 0
        d_cand = torch.from_numpy(cand.view(np.int32)).cuda()
        d_sel.fill_(-1); d_n.fill_(-1)
        torch.cuda.synchronize()
        rc = L.b2orb_distribute_static(hd.h, level, d_cand.data_ptr(), len(cand), d_sel.data_ptr(), d_n.data_ptr())
        assert rc == 0, hd.L.b2orb_last_error(hd.h)
        hd.sync()
        n = int(d_n.cpu()[0])
        assert n == len(want), (level, n, len(want))
        sel = d_sel.cpu().numpy().view(np.uint32)[:n]
        x = (sel & 0xFFF).astype(np.float32) + np.float32(16)
        y = ((sel >> 12) & 0xFFF).astype(np.float32) + np.float32(16)
        if level:                                             # src/ORBextractor.cc:1095-1101: pt *= scale (fp32)
            x = x * scale[level]; y = y * scale[level]
        assert np.array_equal(x, want["x"]) and np.array_equal(y, want["y"]), level
        assert np.array_equal((sel >> 24).astype(np.float32), want["response"]), level
        assert (d_sel.cpu().numpy()[n:] == -1).all()          # nothing written beyond the count
        checked += 1
    assert checked == 3
    hd.close()


def test_static_quadtree_rejects_sizes_beyond_its_workspace(synth):
    import torch
    fe = pkg("frontend")
    L = _bind(fe)
    big = fe.Handle(20000, 1.2, 8, 20, 7, 640, 480, max_images=1)      # level-0 quota 4342 > 4096
    cand = torch.tensor([5 | 7 << 12 | 30 << 24, 40 | 9 << 12 | 50 << 24, 41 | 60 << 12 | 20 << 24], dtype=torch.int32, device="cuda")
    sel = torch.zeros(big.keypoint_capacity, dtype=torch.int32, device="cuda")
    n = torch.zeros(1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = L.b2orb_distribute_static(big.h, 0, cand.data_ptr(), 3, sel.data_ptr(), n.data_ptr())
    assert rc == fe.ERR_CAPACITY and b"B2ORB_STATIC_MAX_FEATURES" in big.L.b2orb_last_error(big.h)
    assert L.b2orb_distribute_static(big.h, 5, cand.data_ptr(), 3, sel.data_ptr(), n.data_ptr()) == 0   # smaller level: fine
    big.sync()
    # all three points fall into the same child of the root, so the list size does not change and DistributeOctTree stops
    # (src/ORBextractor.cc:669): one node, its best response wins
    assert int(n.cpu()[0]) == 1 and int(sel.cpu()[0]) == (40 | 9 << 12 | 50 << 24)
    rc = L.b2orb_distribute_static(big.h, 5, cand.data_ptr(), 300000, sel.data_ptr(), n.data_ptr())
    assert rc == fe.ERR_CAPACITY and b"B2ORB_STATIC_MAX_CANDIDATES" in big.L.b2orb_last_error(big.h)
    assert L.b2orb_distribute_static(big.h, 9, cand.data_ptr(), 3, sel.data_ptr(), n.data_ptr()) == fe.ERR_INVALID
    big.close()

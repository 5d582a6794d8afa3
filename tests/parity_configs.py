"""Oracle checks used by tools/bench_configs.py for its bit_exact columns.  They live under tests/ because only tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs may touch oracle/ (the oracle is the checker, never the
product); the tool passes in what the CUDA path produced and gets a verdict back."""
import time

import numpy as np

from oracle import oracle as orc


def stereo_pair_bit_exact(cfg, image_l, image_r, out, pair=0):
    """out: arrays of Handle.stereo_frontend; True when keypoints, descriptors and BF matches of `pair` equal the oracle's."""
    ex = orc.OracleExtractor(*cfg)
    kl, dl = ex(image_l); kr, dr = ex(image_r)
    nl, nr = int(out["n_l"][pair]), int(out["n_r"][pair])
    ok = nl == len(kl) and nr == len(kr) and np.array_equal(out["desc_l"][pair, :nl], dl) and \
        np.array_equal(out["desc_r"][pair, :nr], dr) and \
        all(np.array_equal(out["kp_l"][pair, :nl][f], kl[f]) for f in ("x", "y", "angle", "octave", "response"))
    ti, di = orc.bf_match(dl, dr)
    return bool(ok and np.array_equal(out["train_idx"][pair, :nl], ti) and np.array_equal(out["dist"][pair, :nl], di))


def cfsd_live_reference(cfg, frame, w, h, maps):
    """CPU port of ingest + rectification + extraction of the left eye: (keypoints, descriptors, ingest+rectify ms)."""
    t0 = time.perf_counter()
    gl, gr = orc.cfsd_ingest(frame, 2 * w, h)
    rl = orc.remap(gl, *maps[0]); orc.remap(gr, *maps[1])
    ms = (time.perf_counter() - t0) * 1e3
    kl, dl = orc.OracleExtractor(*cfg)(rl)
    return kl, dl, ms

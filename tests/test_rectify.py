"""Next row f1: rectification cv::remap in front of the pyramid, src/feature-tracker.cpp:75-76.

Known answers: tests/golden/rectify.json holds SHA-256 of REAL cv2.remap outputs (tools/gen_golden.py) on the
deterministic fixed-point maps of synth.rectification_maps.  CPU: the oracle restatement reproduces them (and matches
cv2 directly when cv2 is importable).  GPU: b2orb_remap reproduces them, and the fused path (raw frames in,
rectified on the device while staging) equals the un-fused one on pre-rectified frames, bit for bit."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests.conftest import pkg

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "rectify.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load(name, synth):
    g = GOLD[name]
    w, h, nf, ini, mn, seed, passes, frame = g["params"]
    cv = synth.canvas(w, h, seed, passes)
    left, right = synth.stereo_pair(cv, w, h, frame, seed)
    assert sha(left) == g["left_sha"] and sha(right) == g["right_sha"], "synthetic input generator drifted"
    maps = [synth.rectification_maps(w, h, e) for e in (0, 1)]
    for e in (0, 1):
        assert sha(maps[e][0]) + sha(maps[e][1]) == g[f"map_sha_{e}"], "map generator drifted"
    return g, (w, h, nf, ini, mn), (left, right), maps


@pytest.mark.parametrize("name", sorted(GOLD))
def test_oracle_remap_reproduces_cv2_known_answers(name, oracle, synth):
    g, _, imgs, maps = load(name, synth)
    for e in (0, 1):
        assert g[f"oob_pixels_{e}"] > 500                       # BORDER_CONSTANT taps are exercised
        r = oracle.remap(imgs[e], *maps[e])
        assert r[100, ::16].tolist() == g[f"rect_row100_{e}"]
        assert sha(r) == g[f"rect_sha_{e}"]


def test_oracle_remap_matches_cv2_on_real_rectify_maps(oracle, synth):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for (w, h) in ((752, 480), (333, 211)):
        K = np.array([[0.61 * w, 0, 0.49 * w], [0, 0.61 * w, 0.52 * h], [0, 0, 1.0]])
        D = np.array([-0.2834, 0.0739, 0.00019, 1.76e-05])
        R = cv2.Rodrigues(np.array([0.012, -0.02, 0.004]))[0]
        P = K.copy(); P[0, 2] += 7.5; P[1, 2] -= 3.25
        m1, m2 = cv2.initUndistortRectifyMap(K, D, R, P, (w, h), cv2.CV_16SC2)
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        assert np.array_equal(oracle.remap(img, m1, m2), cv2.remap(img, m1, m2, cv2.INTER_LINEAR))


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GOLD))
def test_cuda_remap_reproduces_cv2_known_answers(name, synth):
    fe = pkg("frontend")
    g, (w, h, nf, ini, mn), imgs, maps = load(name, synth)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    with pytest.raises(fe.B2Error):
        hd.remap(imgs[0], 0)                                    # no maps yet
    hd.set_rectification(maps[0][0], maps[0][1], maps[1][0], maps[1][1])
    for e in (0, 1):
        assert sha(hd.remap(imgs[e], e)) == g[f"rect_sha_{e}"]
    hd.close()


@pytest.mark.gpu
@pytest.mark.parametrize("shape,npairs,chunk", [("cfsd", 3, 0), ("kitti", 5, 2), ("odd", 2, 0)])
def test_fused_rectification_equals_prerectified_frames(shape, npairs, chunk, oracle, synth):
    """processImage = remap x2 + internalMatch: raw frames through the fused device path == oracle-rectified frames
    through the plain path (keypoints, descriptors, matches, temporal circular matches), also when the batch is
    pipelined in chunks over the lanes."""
    fe = pkg("frontend")
    if shape == "odd":
        w, h, nf, ini, mn = 333, 211, 300, 20, 7
        ls, rs = synth.sequence((w, h), npairs, seed=9)
    else:
        w, h, nf, ini, mn = synth.SHAPES[shape]
        ls, rs = synth.sequence(shape, npairs, seed=9)
    maps = [synth.rectification_maps(w, h, e) for e in (0, 1)]
    rl = np.stack([oracle.remap(im, *maps[0]) for im in ls]); rr = np.stack([oracle.remap(im, *maps[1]) for im in rs])
    plain = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs).debug_set("pipe_chunk", chunk)
    fused = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs).debug_set("pipe_chunk", chunk)
    fused.set_rectification(maps[0][0], maps[0][1], maps[1][0], maps[1][1])
    a = plain.stereo_frontend(rl, rr, temporal=True)
    b = fused.stereo_frontend(ls, rs, temporal=True)
    assert a["n_l"].min() > 100 and a["n_r"].min() > 100
    for k in ("n_l", "n_r", "n_match", "n_circ"):
        assert np.array_equal(a[k], b[k]), k
    for p in range(npairs):
        for key, cnt in (("kp_l", "n_l"), ("desc_l", "n_l"), ("kp_r", "n_r"), ("desc_r", "n_r"), ("train_idx", "n_l"),
                         ("dist", "n_l"), ("match_l", "n_match"), ("match_r", "n_match"), ("circ_cur", "n_circ"),
                         ("circ_hist", "n_circ")):
            n = int(a[cnt][p])
            assert np.array_equal(a[key][p, :n], b[key][p, :n]), f"pair {p}: {key}"
    # switching the maps off restores the plain behaviour
    fused.set_rectification()
    fused.reset_history(); plain.reset_history()
    c = fused.stereo_frontend(rl, rr, temporal=True)
    d = plain.stereo_frontend(rl, rr, temporal=True)
    assert np.array_equal(c["n_match"], d["n_match"]) and np.array_equal(c["desc_l"][0, :c["n_l"][0]], d["desc_l"][0, :d["n_l"][0]])
    plain.close(); fused.close()

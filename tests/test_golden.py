"""Known-answer tests against tests/golden/*.npz (made by tools/gen_golden.py from real OpenCV 4.13 primitives
wired like src/ORBextractor.cc:765-853,1043-1105 and src/feature-tracker.cpp:161-164,208-218 — NOT by the
oracle).  CPU: the oracle reproduces every vector.  GPU: the CUDA path (through the C ABI) does too."""
import glob
import hashlib
import os

import numpy as np
import pytest

from tests.conftest import pkg

GOLDEN = sorted(p for p in glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "*.npz"))
                if os.path.basename(p) != "triangulate.npz")   # that one belongs to tests/test_triangulate.py


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load(path, synth):
    g = np.load(path)
    w, h, nf, ini, mn, seed, passes, frame = [int(v) for v in g["params"]]
    cv = synth.canvas(w, h, seed, passes)
    left, right = synth.stereo_pair(cv, w, h, frame, seed)
    assert sha(left) == str(g["left_sha"]) and sha(right) == str(g["right_sha"]), "synthetic input generator drifted"
    return g, (w, h, nf, ini, mn), left, right


def check_extract(got, g, side):
    k, d = got
    gk, gd = g["kp_" + side], g["desc_" + side]
    assert len(k) == len(gk)
    for f in ("x", "y", "size", "angle", "response", "octave", "class_id"):
        assert np.array_equal(k[f], gk[f]), f"{side}: {f}"
    assert np.array_equal(d, gd)


def test_golden_files_exist():
    assert len(GOLDEN) >= 5


@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_oracle_reproduces_golden(path, oracle, synth):
    g, (w, h, nf, ini, mn), left, right = load(path, synth)
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    kl, dl = ex(left)
    for lvl in range(8):
        assert sha(ex.level(lvl)) == str(g["level_sha_l"][lvl])
        b = ex.level(lvl, blurred=True)
        assert (sha(b) if b is not None else "") == str(g["blur_sha_l"][lvl])
        assert len(ex.raw_corners(lvl)) == int(g["raw_count_l"][lvl])
    check_extract((kl, dl), g, "l")
    kr, dr = ex(right)
    check_extract((kr, dr), g, "r")
    ti, di = oracle.bf_match(dl, dr)
    assert np.array_equal(ti, g["train_idx"]) and np.array_equal(di, g["dist"])
    ml, mr = oracle.stereo_filter(ti, di, kl, kr, 2.0, 30.0, 0.1)
    assert np.array_equal(ml, g["match_l"]) and np.array_equal(mr, g["match_r"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", GOLDEN, ids=[os.path.basename(p)[:-4] for p in GOLDEN])
def test_cuda_reproduces_golden(path, synth):
    fe = pkg("frontend")
    g, (w, h, nf, ini, mn), left, right = load(path, synth)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    o = hd.stereo_frontend(left, right)
    nl, nr, m = int(o["n_l"][0]), int(o["n_r"][0]), int(o["n_match"][0])
    check_extract((o["kp_l"][0, :nl], o["desc_l"][0, :nl]), g, "l")
    check_extract((o["kp_r"][0, :nr], o["desc_r"][0, :nr]), g, "r")
    for lvl in range(8):
        assert sha(hd.debug_level(0, lvl)) == str(g["level_sha_l"][lvl])
        if str(g["blur_sha_l"][lvl]):
            assert sha(hd.debug_level(0, lvl, blurred=True)) == str(g["blur_sha_l"][lvl])
        assert len(hd.debug_candidates(0, lvl)) == int(g["raw_count_l"][lvl])
    assert np.array_equal(o["train_idx"][0, :nl], g["train_idx"]) and np.array_equal(o["dist"][0, :nl], g["dist"])
    assert m == len(g["match_l"])
    assert np.array_equal(o["match_l"][0, :m], g["match_l"]) and np.array_equal(o["match_r"][0, :m], g["match_r"])
    # the stand-alone entry points agree with the fused one
    k, d = hd.extract(left)
    check_extract((k, d), g, "l")
    ti, di = hd.match(g["desc_l"], g["desc_r"])
    assert np.array_equal(ti, g["train_idx"]) and np.array_equal(di, g["dist"])
    hd.close()


REF_IMAGES = ["/root/reference/benchmark/cvRectify/cfsd360pLeft.jpg", "/root/reference/benchmark/cvRectify/kitti0.png"]


@pytest.mark.parametrize("path", REF_IMAGES)
def test_oracle_vs_cv2_hybrid_on_reference_images(path, oracle):
    """Real camera frames shipped with the reference (input fixtures only): oracle == cv2 pipeline.
    Skipped where /root/reference is not mounted (the GPU box)."""
    cv2 = pytest.importorskip("cv2")
    if not os.path.exists(path):
        pytest.skip("reference tree not mounted")
    from tests import hybrid_cv2
    img = cv2.imread(path, cv2.IMREAD_GRAYSCALE)
    ini = 12 if "kitti" in path else 20
    ex = oracle.OracleExtractor(1000, 1.2, 8, ini, 7)
    kps, desc = ex(img)
    _, _, _, hk, hd = hybrid_cv2.extract(cv2, img, 1000, 1.2, 8, ini, 7)
    hk = np.array([k[:6] for k in hk], np.float32)
    assert len(kps) == len(hk) and len(kps) > 500
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(kps[f], hk[:, i]), f
    assert np.array_equal(kps["octave"], hk[:, 5].astype(np.int32)) and np.array_equal(desc, hd)

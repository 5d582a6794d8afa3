"""Pin every OpenCV / libm primitive of the CPU oracle against the cv2 4.13 wheel of this image.

The reference has no tests or golden vectors (test/vio-test.cpp is empty); the arithmetic of its hot
path lives in OpenCV (un-vendored) and libm, so the oracle's restatements are checked here against
the only OpenCV available.  Each test cites the reference call site it pins.
"""
import ctypes
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

SHAPES = [(752, 480), (1241, 376), (672, 376), (1392, 512)]


def _texture(w, h, seed, synth, passes=3):
    return synth.canvas(w, h, seed, passes)[:h, :w].copy()


@pytest.mark.parametrize("wh", SHAPES)
def test_resize_chain_matches_cv2(wh, oracle, synth):
    """cv::resize INTER_LINEAR chain of ComputePyramid (src/ORBextractor.cc:1107-1132)."""
    w, h = wh
    img = _texture(w, h, 1, synth)
    ex = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    _, inv, _, _ = ex.tables()
    a = b = img
    for lvl in range(1, 8):
        dw = int(np.rint(np.float32(w) * inv[lvl])); dh = int(np.rint(np.float32(h) * inv[lvl]))
        a = oracle.resize(a, dw, dh)
        b = cv2.resize(b, (dw, dh), interpolation=cv2.INTER_LINEAR)
        assert np.array_equal(a, b), f"level {lvl}"


def test_resize_odd_sizes(oracle):
    """Downscale / identity / one-axis cases.  (Two-axis UPscaling differs from cv2 4.13 by +1 on a few
    pixels; the path never upscales — scaleFactor > 1 is required, src/ORBextractor.cc:1111-1120.)"""
    rng = np.random.default_rng(3)
    for (sw, sh, dw, dh) in [(37, 23, 31, 19), (64, 48, 64, 48), (50, 40, 17, 13), (9, 7, 3, 2), (100, 3, 83, 3),
                             (50, 40, 61, 40), (333, 211, 278, 176)]:
        img = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
        assert np.array_equal(oracle.resize(img, dw, dh), cv2.resize(img, (dw, dh), interpolation=cv2.INTER_LINEAR))


@pytest.mark.parametrize("wh", [(97, 113), (752, 480), (210, 134), (8, 9)])
def test_gaussian_blur_matches_cv2(wh, oracle):
    """GaussianBlur(7x7, sigma 2, REFLECT_101) of src/ORBextractor.cc:1086 (cv2 4.x taps)."""
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (wh[1], wh[0]), dtype=np.uint8)
    ref = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
    assert np.array_equal(oracle.blur(img), ref)
    assert not np.array_equal(oracle.blur(img, oracle.TAPS_CV32), ref)


@pytest.mark.parametrize("thr", [7, 12, 20])
def test_fast_roi_matches_cv2(thr, oracle, synth):
    """cv::FAST(roi, thr, nms=true) on sub-rectangles (src/ORBextractor.cc:809-815): same points,
    responses and order."""
    img = _texture(320, 240, 2, synth)
    det = cv2.FastFeatureDetector_create(thr, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    rois = [(0, 0, 320, 240), (16, 16, 52, 54), (100, 50, 137, 88), (283, 200, 304, 224), (5, 5, 11, 30), (7, 9, 14, 16)]
    for (x0, y0, x1, y1) in rois:
        kps = det.detect(np.ascontiguousarray(img[y0:y1, x0:x1]))
        ref = np.array([[k.pt[0], k.pt[1], k.response] for k in kps], np.float32).reshape(-1, 3)
        got = oracle.fast(img, (x0, y0, x1, y1), thr)
        assert got.shape == ref.shape and np.array_equal(got, ref), (x0, y0, x1, y1)


def test_fast_noise_matches_cv2(oracle):
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (90, 120), dtype=np.uint8)
    for thr in (7, 20, 60):
        det = cv2.FastFeatureDetector_create(thr, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        kps = det.detect(img)
        ref = np.array([[k.pt[0], k.pt[1], k.response] for k in kps], np.float32).reshape(-1, 3)
        assert np.array_equal(oracle.fast(img, (0, 0, 120, 90), thr), ref)


def test_fast_atan2_matches_cv2(oracle):
    """cv::fastAtan2 of IC_Angle (src/ORBextractor.cc:103): bit-exact incl. axes and (0,0)."""
    rng = np.random.default_rng(7)
    ys = rng.integers(-200000, 200000, 20000).astype(np.float32)
    xs = rng.integers(-200000, 200000, 20000).astype(np.float32)
    pts = list(zip(ys, xs)) + [(0, 0), (0, 5), (5, 0), (0, -5), (-5, 0), (3, 3), (-3, 3), (3, -3), (-3, -3)]
    for y, x in pts:
        a = np.float32(oracle.fast_atan2(y, x)); b = np.float32(cv2.fastAtan2(float(y), float(x)))
        assert a.tobytes() == b.tobytes(), (y, x, a, b)


def test_sincosf_restatement_matches_libm_sampled(oracle):
    """glibc cosf/sinf used at src/ORBextractor.cc:113: sampled check here (stride through every
    float in [0, 2*pi]); tests/test_oracle_slow.py and tools/check_sincosf.c run it exhaustively."""
    L = oracle.lib()
    hi = int(np.float32(6.2832).view(np.uint32))
    assert L.orc_sincos_check(0, 2_000_000) == 0
    assert L.orc_sincos_check(hi - 4_000_000, hi) == 0
    libm = ctypes.CDLL("libm.so.6")
    libm.cosf.restype = ctypes.c_float; libm.cosf.argtypes = [ctypes.c_float]
    libm.sinf.restype = ctypes.c_float; libm.sinf.argtypes = [ctypes.c_float]
    for u in range(1, hi, 104729):
        a = float(np.uint32(u).view(np.float32))
        assert np.float32(L.orc_cosf(a)).tobytes() == np.float32(libm.cosf(a)).tobytes()
        assert np.float32(L.orc_sinf(a)).tobytes() == np.float32(libm.sinf(a)).tobytes()


def test_bf_match_matches_cv2_with_ties(oracle):
    """BFMatcher(NORM_HAMMING).match (src/feature-tracker.cpp:161-163): first minimum wins."""
    rng = np.random.default_rng(9)
    protos = rng.integers(0, 256, (6, 32), dtype=np.uint8)
    q = protos[rng.integers(0, 6, 500)].copy(); t = protos[rng.integers(0, 6, 700)].copy()
    flip = rng.integers(0, 500, 100); q[flip, rng.integers(0, 32, 100)] ^= 1
    ms = cv2.BFMatcher(cv2.NORM_HAMMING).match(q, t)
    ti, di = oracle.bf_match(q, t)
    assert [m.queryIdx for m in ms] == list(range(500))
    assert np.array_equal(ti, np.array([m.trainIdx for m in ms]))
    assert np.array_equal(di, np.array([int(m.distance) for m in ms]))

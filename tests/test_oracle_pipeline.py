"""Pin the C++ oracle END TO END against the independent cv2-primitive pipeline (tests/hybrid_cv2.py):
pyramid, raw per-cell FAST corners incl. order, blurred levels, quadtree selection incl. order, angles,
descriptors, final coordinates.  Reference: src/ORBextractor.cc:1043-1105."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
from tests import hybrid_cv2


CASES = [
    # (w, h, nfeatures, ini, min, seed, passes)
    (752, 480, 1000, 20, 7, 0, 3),
    (672, 376, 1000, 20, 7, 1, 3),
    (400, 300, 500, 20, 7, 2, 8),    # low texture -> minThFAST fallback cells
    (1241, 376, 2000, 12, 7, 3, 3),
]


@pytest.mark.parametrize("case", CASES)
def test_oracle_equals_cv2_hybrid(case, oracle, synth):
    w, h, nf, ini, mn, seed, passes = case
    img = synth.canvas(w, h, seed, passes)[:h, :w].copy()
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    kps, desc = ex(img)
    pyr, raw, blur, hk, hd = hybrid_cv2.extract(cv2, img, nf, 1.2, 8, ini, mn)
    sf, inv, quota, umax = hybrid_cv2.tables(nf, 1.2, 8)
    s, i, q, u = ex.tables()
    assert np.array_equal(s, np.array(sf, np.float32)) and np.array_equal(q, np.array(quota))
    assert list(u) == umax
    for lvl in range(8):
        assert np.array_equal(ex.level(lvl), pyr[lvl]), f"pyramid {lvl}"
        r = ex.raw_corners(lvl)
        assert np.array_equal(r, np.array(raw[lvl], np.float32).reshape(-1, 3)), f"raw corners {lvl}"
        if blur[lvl] is not None:
            assert np.array_equal(ex.level(lvl, blurred=True), blur[lvl]), f"blur {lvl}"
    assert len(kps) == len(hk)
    hk = np.array([k[:5] for k in hk], np.float32)
    assert np.array_equal(kps["x"], hk[:, 0]) and np.array_equal(kps["y"], hk[:, 1])
    assert np.array_equal(kps["size"], hk[:, 2])
    assert np.array_equal(kps["angle"], hk[:, 3])
    assert np.array_equal(kps["response"], hk[:, 4])
    assert np.array_equal(desc, hd)
    assert nf <= len(kps) <= nf + 3 * 8


def test_quota_and_scale_tables(oracle):
    """ctor tables (src/ORBextractor.cc:410-470) for the configs of SURVEY App. B."""
    ex = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    s, i, q, u = ex.tables()
    assert list(q) == [217, 181, 151, 126, 105, 87, 73, 60]
    assert list(u) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert [int(31 * float(x)) for x in s] == [31, 37, 44, 53, 64, 77, 92, 111]
    assert list(oracle.OracleExtractor(800, 1.2, 8, 20, 7).tables()[2]) == [174, 145, 121, 101, 84, 70, 58, 47]
    assert list(oracle.OracleExtractor(2000, 1.2, 8, 12, 7).tables()[2]) == [434, 362, 302, 251, 209, 175, 145, 122]
    assert list(oracle.OracleExtractor(8000, 1.2, 8, 20, 7).tables()[2]) == [1737, 1448, 1207, 1005, 838, 698, 582, 485]


def test_octtree_python_vs_cpp_random(oracle):
    """DistributeOctTree (src/ORBextractor.cc:539-763): C++ oracle == plain-Python list simulation on
    random point sets incl. clustered ones and N larger than the point count."""
    rng = np.random.default_rng(4)
    for trial in range(12):
        w, h = int(rng.integers(80, 700)), int(rng.integers(60, 400))
        n = int(rng.integers(1, 3000))
        if trial % 3 == 0:
            xs = np.clip(rng.normal(w / 2, w / 10, n), 3, w - 4).astype(int); ys = np.clip(rng.normal(h / 2, h / 12, n), 3, h - 4).astype(int)
        else:
            xs = rng.integers(3, w - 3, n); ys = rng.integers(3, h - 3, n)
        pts = np.unique(np.stack([ys, xs], 1), axis=0)  # unique, row-major like raw FAST output
        resp = rng.integers(7, 60, len(pts))
        xyr = np.stack([pts[:, 1], pts[:, 0], resp], 1).astype(np.float32)
        n_want = int(rng.integers(1, 400))
        a = oracle.octtree(xyr, 16, 16 + w, 16, 16 + h, n_want)
        b = hybrid_cv2.octtree([tuple(map(float, r)) for r in xyr], 16, 16 + w, 16, 16 + h, n_want)
        assert np.array_equal(a, np.array(b, np.float32).reshape(-1, 3)), trial

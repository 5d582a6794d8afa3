"""Next row f4: cv::triangulatePoints + the depth gates of featurePoolUpdate / structFromMotion
(src/feature-tracker.cpp:365-379, :487-509).

Known answers: tests/golden/triangulate.npz holds REAL cv2.triangulatePoints outputs (tools/gen_golden.py) for the
EuRoC / KITTI / cfsd rectified projection pairs (keypoint-like float coordinates, zero and negative disparities
included) and for a general un-rectified camera pair.  Floating point (fp64), but every operation of OpenCV's 4x4 Jacobi
SVD is restated in order without FMA, so the bar here is BIT-EXACT: oracle == cv2 (CPU), CUDA == oracle == golden (GPU).
The depth gates are compared on every point (NaN / inf from W -> 0 included)."""
import os

import numpy as np
import pytest

from tests.conftest import pkg

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triangulate.npz"))
CASES = sorted({k.rsplit("_", 1)[0] for k in GOLD.files if k.endswith("_P1")})


def bits(a):
    return np.ascontiguousarray(a, np.float64).view(np.uint64)


def case(name):
    return [GOLD[f"{name}_{k}"] for k in ("P1", "P2", "pl", "pr", "points4d")]


@pytest.fixture(scope="module")
def fe():
    return pkg("frontend")


# ---------------------------------------------------------------------------------------------- CPU: oracle pinned
@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_cv2_known_answers(name, oracle):
    P1, P2, pl, pr, ref = case(name)
    got = oracle.triangulate(P1, P2, pl, pr)
    assert got.shape == ref.shape == (4, len(pl))
    assert np.array_equal(bits(got), bits(ref))
    if name != "general":
        assert (pl[:, 0] == pr[:, 0]).sum() > 0          # zero disparity is exercised


def test_oracle_matches_cv2_directly(oracle):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(11)
    P1, P2 = case("kitti")[:2]
    n = 5000
    pl = np.stack([rng.uniform(0, 1241, n), rng.uniform(0, 376, n)], 1).astype(np.float32).astype(np.float64)
    pr = pl - np.stack([rng.uniform(-2, 80, n), rng.normal(0, 0.05, n)], 1).astype(np.float32)
    ref = cv2.triangulatePoints(P1, P2, pl.T.copy(), pr.T.copy())
    assert np.array_equal(bits(oracle.triangulate(P1, P2, pl, pr)), bits(ref))
    # a general pair, un-normalised pixel magnitudes
    P1g, P2g = case("general")[:2]
    ref = cv2.triangulatePoints(P1g, P2g, pl.T.copy(), pr.T.copy())
    assert np.array_equal(bits(oracle.triangulate(P1g, P2g, pl, pr)), bits(ref))


def test_oracle_depth_gate_follows_the_reference_comparisons(oracle):
    # columns: in range, beyond maxDepth, closer than 0.1, negative depth, W = 0 (inf), 0/0 (nan)
    p4 = np.array([[1, 2, 0.01, 1, 1, 0], [1, 2, 0.01, 1, 1, 0], [5, 50, 0.05, -3, 1, 0], [1, 2, 1, 1, 0, 0]], np.float64)
    pc, fl = oracle.depth_gate(p4, 20.0)
    assert fl.tolist() == [3, 0, 1, 1, 0, 3]
    # NaN depth: `depth > maxDepth` and `depth < 0.1` are both false, so the reference keeps the point in both paths
    assert fl[5] == 3 and np.isnan(pc[5, 2]) and np.isinf(pc[4, 2])
    assert np.array_equal(pc[0], [1, 1, 5]) and np.array_equal(pc[1], [1, 1, 25])


# ---------------------------------------------------------------------------------------------- GPU: CUDA parity
@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_triangulate_reproduces_cv2_known_answers(name, fe, oracle):
    P1, P2, pl, pr, ref = case(name)
    hd = fe.Handle(500, 1.2, 8, 20, 7, 128, 128, max_images=2)
    got = hd.triangulate(P1, P2, pl, pr)
    assert np.array_equal(bits(got), bits(ref))
    assert np.array_equal(bits(got), bits(oracle.triangulate(P1, P2, pl, pr)))
    hd.close()


@pytest.mark.gpu
def test_cuda_triangulate_large_and_edge_sizes(fe, oracle):
    rng = np.random.default_rng(5)
    P1, P2 = case("euroc")[:2]
    P1g, P2g = case("general")[:2]
    hd = fe.Handle(500, 1.2, 8, 20, 7, 128, 128, max_images=2)
    assert hd.triangulate(P1, P2, np.zeros((0, 2)), np.zeros((0, 2))).shape == (4, 0)      # empty match list
    for n in (1, 127, 129, 4097, 200_000):                                                 # scratch growth, ragged tail
        pl = np.stack([rng.uniform(0, 752, n), rng.uniform(0, 480, n)], 1).astype(np.float32).astype(np.float64)
        pr = pl - np.stack([rng.uniform(-1, 70, n), rng.normal(0, 0.05, n)], 1)
        for (a, b) in ((P1, P2), (P1g, P2g)):
            assert np.array_equal(bits(hd.triangulate(a, b, pl, pr)), bits(oracle.triangulate(a, b, pl, pr)))
    # identical pixels in both views with identical projections: rank-deficient system, still the same bits
    pl = np.stack([rng.uniform(0, 752, 64), rng.uniform(0, 480, 64)], 1)
    assert np.array_equal(bits(hd.triangulate(P1, P1, pl, pl)), bits(oracle.triangulate(P1, P1, pl, pl)))
    with pytest.raises(fe.B2Error):
        hd.check(hd.L.b2orb_triangulate(hd.h, None, None, None, None, 4, None))
    hd.close()


def _expected_fused(o, p, P1, P2, max_depth, oracle):
    m = int(o["n_match"][p])
    il, ir = o["match_l"][p, :m], o["match_r"][p, :m]
    kl, kr = o["kp_l"][p], o["kp_r"][p]
    pl = np.stack([kl["x"][il], kl["y"][il]], 1).astype(np.float64)     # Point2f -> Point2d, :214-215
    pr = np.stack([kr["x"][ir], kr["y"][ir]], 1).astype(np.float64)
    p4 = oracle.triangulate(P1, P2, pl, pr)
    pc, fl = oracle.depth_gate(p4, max_depth)
    return m, p4, pc, fl


@pytest.mark.gpu
@pytest.mark.parametrize("shape,npairs,chunk", [("euroc", 2, 0), ("cfsd", 7, 3)])
def test_fused_triangulation_behind_the_stereo_filter(shape, npairs, chunk, fe, synth, oracle):
    """b2orb_set_projection: the kept matches of every pair are triangulated in the same device pass; also through the
    pipelined host path (chunks over lanes) and with the caller's cap larger than the handle's."""
    w, h, nf, ini, mn = synth.SHAPES[shape]
    P1, P2 = case(shape)[:2]
    P2 = P2.copy(); P2[0, 3] = -P2[0, 3]     # synth.stereo_pair displaces the right eye's content towards +x
    max_depth = 4.0
    ls, rs = synth.sequence(shape, npairs, seed=21)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs).debug_set("pipe_chunk", chunk)
    plain = hd.stereo_frontend(ls, rs, temporal=True)
    assert "points4d" not in plain
    hd.reset_history()
    hd.set_projection(P1, P2, max_depth)
    o = hd.stereo_frontend(ls, rs, temporal=True)
    for k in ("n_match", "match_l", "match_r", "n_circ"):                # everything else is untouched
        assert np.array_equal(o[k], plain[k])
    seen = set()
    for p in range(npairs):
        m, p4, pc, fl = _expected_fused(o, p, P1, P2, max_depth, oracle)
        assert m > 20
        assert np.array_equal(bits(o["points4d"][p][:, :m]), bits(p4))
        assert np.array_equal(bits(o["point_cam"][p][:m]), bits(pc))
        assert np.array_equal(o["depth_flags"][p][:m], fl)
        seen |= set(fl.tolist())
    assert {0, 3} <= seen                                                # both sides of the maxDepth gate occur
    # larger caller capacity: row-wise copies
    cap = hd.keypoint_capacity + 13
    arrays = {"n_match": np.zeros(npairs, np.int32), "match_l": np.zeros((npairs, cap), np.int32),
              "match_r": np.zeros((npairs, cap), np.int32), "kp_l": np.zeros((npairs, cap), fe.KP_DTYPE),
              "kp_r": np.zeros((npairs, cap), fe.KP_DTYPE), "points4d": np.zeros((npairs, 4, cap)),
              "point_cam": np.zeros((npairs, cap, 3)), "depth_flags": np.zeros((npairs, cap), np.uint8)}
    s = fe.B2StereoOut()
    s.cap = cap
    for k, v in arrays.items():
        setattr(s, k, v.ctypes.data)
    hd.stereo_frontend(ls, rs, temporal=False, out=(arrays, s))
    for p in range(npairs):
        m, p4, pc, fl = _expected_fused(arrays, p, P1, P2, max_depth, oracle)
        assert np.array_equal(bits(arrays["points4d"][p][:, :m]), bits(p4))
        assert np.array_equal(bits(arrays["point_cam"][p][:m]), bits(pc)) and np.array_equal(arrays["depth_flags"][p][:m], fl)
    hd.set_projection()                                                   # off again
    assert "points4d" not in hd.stereo_frontend(ls[:1], rs[:1])
    hd.close()


@pytest.mark.gpu
def test_feature_tracker_mirror_triangulates(fe, synth, oracle):
    w, h, nf, ini, mn = synth.SHAPES["euroc"]
    P1, P2 = case("euroc")[:2]
    ls, rs = synth.sequence("euroc", 1, seed=4)
    ft = fe.FeatureTracker(w, h, nf, 1.2, 8, ini, mn)
    ft.setProjection(P1, P2, 20.0)
    ft.internalMatch(ls[0], rs[0])
    p4, pc, fl = ft.triangulate()
    ref = oracle.triangulate(P1, P2, ft._curPixelsL, ft._curPixelsR)
    assert p4.shape[1] == len(ft._curPixelsL) > 20 and np.array_equal(bits(p4), bits(ref))
    rpc, rfl = oracle.depth_gate(ref, 20.0)
    assert np.array_equal(bits(pc), bits(rpc)) and np.array_equal(fl, rfl)

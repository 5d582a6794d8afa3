"""GPU parity against the REFERENCE'S OWN extractor: the CUDA path (through the C ABI) vs tests/golden/ref/*.npz, the
committed outputs of /root/reference/src/ORBextractor.cc compiled unmodified (oracle/ref/, tools/gen_ref_golden.py) —
and, when oracle/_ref travelled with the snapshot, vs live runs of that library.  Synthetic shapes of every
BASELINE.json config that fits (incl. config/kitti.yml's 1392x512 and euroc.yml's 800 features), six real camera
frames shipped with the reference, and the OpenCV-3.2 Gaussian taps the reference's Dockerfile names.
Bar: bit-exact keypoints (all seven cv::KeyPoint fields, order included) and descriptors."""
import glob
import os

import numpy as np
import pytest

from tests.conftest import ROOT, pkg
from tests.test_ref_parity import REAL_CASES, SYNTH_CASES, real_case, same, synth_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    return pkg("frontend")


@pytest.mark.parametrize("name", SYNTH_CASES)
def test_cuda_equals_reference_outputs_synthetic(name, fe, synth):
    g, (nf, ini, mn), il, ir = synth_case(synth, name)
    h, w = il.shape
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    for eye, img in (("l", il), ("r", ir)):
        k, d = hd.extract(img)
        assert same(k, d, g["kp_" + eye], g["desc_" + eye]), (name, eye)
    hd.close()
    h32 = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2, blur_taps=[18, 34, 49, 55, 49, 34, 18])
    k, d = h32.extract(il)
    assert same(k, d, g["kp_l_cv32"], g["desc_l_cv32"]), (name, "OpenCV 3.2 taps")
    h32.close()


@pytest.mark.parametrize("name", REAL_CASES)
def test_cuda_equals_reference_outputs_real_frames(name, fe):
    g, (nf, ini, mn), img = real_case(name)
    h, w = img.shape
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    k, d = hd.extract(img)
    assert same(k, d, g["kp"], g["desc"]), name
    hd.close()


def test_cuda_real_stereo_pairs_match_oracle_end_to_end(fe, oracle):
    """internalMatch on the real stereo pairs (cfsd360p L/R, cfsd-imageset 110 L/R): extraction, BF match, filters."""
    for ln, rn in (("cfsd360p_left", "cfsd360p_right"), ("cfsd_set_110_left", "cfsd_set_110_right")):
        gl, (nf, ini, mn), il = real_case(ln)
        gr, _, ir = real_case(rn)
        h, w = il.shape
        hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
        o = hd.stereo_frontend(il, ir)
        nl, nr, m = int(o["n_l"][0]), int(o["n_r"][0]), int(o["n_match"][0])
        assert same(o["kp_l"][0, :nl], o["desc_l"][0, :nl], gl["kp"], gl["desc"])
        assert same(o["kp_r"][0, :nr], o["desc_r"][0, :nr], gr["kp"], gr["desc"])
        ti, di = oracle.bf_match(gl["desc"], gr["desc"])
        assert np.array_equal(o["train_idx"][0, :nl], ti) and np.array_equal(o["dist"][0, :nl], di)
        ml, mr = oracle.stereo_filter(ti, di, gl["kp"], gr["kp"], 2.0, 30.0, 0.1)
        assert m == len(ml) and np.array_equal(o["match_l"][0, :m], ml) and np.array_equal(o["match_r"][0, :m], mr)
        hd.close()


def test_cuda_equals_live_reference_library(fe, oracle, synth):
    """When oracle/_ref/libref_orb.so is present (built in the container, shipped with the snapshot): fresh inputs the
    committed vectors do not cover, straight against the reference's code."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not in this snapshot")
    for (w, h, nf, ini, mn, seed, passes) in ((640, 400, 900, 20, 7, 41, 3), (1241, 376, 2000, 12, 7, 42, 3),
                                              (411, 277, 350, 20, 7, 43, 8), (1392, 512, 2000, 12, 7, 44, 5)):
        ls, rs = synth.sequence((w, h), 2, seed=seed, passes=passes)
        rf = oracle.RefExtractor(nf, 1.2, 8, ini, mn)
        hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
        for img in (ls[1], rs[1]):
            kr, dr = rf(img)
            k, d = hd.extract(img)
            assert same(k, d, kr, dr), (w, h, nf)
        hd.close()


def test_blur_taps_saturate_like_opencv32(fe, oracle):
    """OpenCV 3.2's rounded taps sum to 257, so (acc + 32768) >> 16 reaches 257 on a white region and must saturate
    (FixedPtCastEx + saturate_cast<uchar>).  Oracle and CUDA blurred levels on an image with 255-plateaus."""
    rng = np.random.default_rng(8)
    img = rng.integers(0, 256, (240, 320), dtype=np.uint8)
    img[20:120, 30:200] = 255
    img[130:200, 100:300] = np.where(rng.random((70, 200)) < 0.5, 255, 254).astype(np.uint8)
    taps = [18, 34, 49, 55, 49, 34, 18]
    ref = oracle.blur(img, taps)
    assert ref.max() == 255 and (ref[30:110, 40:190] == 255).all()
    ex = oracle.OracleExtractor(400, 1.2, 8, 20, 7, taps=taps)
    ko, do = ex(img)
    hd = fe.Handle(400, 1.2, 8, 20, 7, 320, 240, max_images=2, blur_taps=taps)
    k, d = hd.extract(img)
    assert len(ko) > 100 and same(k, d, ko, do)
    assert np.array_equal(hd.debug_level(0, 0, blurred=True), ref)
    for lvl in range(1, 8):
        rb = ex.level(lvl, blurred=True)
        if rb is not None:
            assert np.array_equal(hd.debug_level(0, lvl, blurred=True), rb), lvl
    hd.close()


def test_sfm_shaped_full_set_matching_between_consecutive_frames(fe, oracle, synth):
    """structFromMotion (src/feature-tracker.cpp:414-545): left+right extraction, stereo match, then
    matcher.match(descriptors of frame t, reference-keyframe descriptors) on FULL extracted sets (:447, :494), both
    directions, on two consecutive extracted frames — not synthetic prototypes."""
    w, h, nf, ini, mn = synth.SHAPES["euroc"]
    ls, rs = synth.sequence("euroc", 3, seed=61)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    k0, d0 = hd.extract(ls[0])                      # _refKeypointsL/_refDescriptorsL (:419-426)
    ko, do = ex(ls[0])
    assert same(k0, d0, ko, do)
    for t in (1, 2):
        kl, dl = hd.extract(ls[t]); kr, dr = hd.extract(rs[t])
        ti, di = hd.match(dl, dr)                   # :447
        oi, od = oracle.bf_match(dl, dr)
        assert np.array_equal(ti, oi) and np.array_equal(di, od)
        ml, mr = oracle.stereo_filter(oi, od, kl, kr, 2.0, 30.0, 0.1)
        inl = dl[ml]                                # descriptors of the kept matches (stand-in for the RANSAC inliers)
        for q, t_ in ((inl, d0), (d0, inl), (dl, d0), (d0, dl)):   # :494 and the reverse direction
            a = hd.match(q, t_); b = oracle.bf_match(q, t_)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    hd.close()


def _check_pair(o, p, kl, dl, kr, dr, oracle):
    nl, nr, m = int(o["n_l"][p]), int(o["n_r"][p]), int(o["n_match"][p])
    assert same(o["kp_l"][p, :nl], o["desc_l"][p, :nl], kl, dl), f"pair {p} left"
    assert same(o["kp_r"][p, :nr], o["desc_r"][p, :nr], kr, dr), f"pair {p} right"
    ti, di = oracle.bf_match(dl, dr)
    assert np.array_equal(o["train_idx"][p, :nl], ti) and np.array_equal(o["dist"][p, :nl], di)
    ml, mr = oracle.stereo_filter(ti, di, kl, kr, 2.0, 30.0, 0.1)
    assert m == len(ml) and np.array_equal(o["match_l"][p, :m], ml) and np.array_equal(o["match_r"][p, :m], mr)
    return dl[ml], dr[mr]


def test_bench_configuration_every_pair_checked(fe, oracle, synth):
    """bench.py's own configuration: KITTI 1241x376, 2000 features, stereo + temporal, 128 pairs per step, two handles on
    two streams, device-resident entry point + b2orb_stereo_download — EVERY pair of both handles against the oracle
    (extraction, stereo match, filter, temporal circular lists)."""
    import torch
    w, h, nf, ini, mn = synth.SHAPES["kitti"]
    n = 128
    ls, rs = synth.sequence("kitti", n, seed=0)
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    want = [ex(ls[p]) + ex(rs[p]) for p in range(n)]
    dl_ = torch.from_numpy(ls).cuda(); dr_ = torch.from_numpy(rs).cuda()
    hds = [fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * n) for _ in range(2)]
    outs = [hd.alloc_stereo_out(n) for hd in hds]
    torch.cuda.synchronize()
    for rep in range(2):                      # the second step starts from the first one's history, like the bench loop
        for hd in hds:
            hd.stereo_frontend_device(dl_.data_ptr(), dr_.data_ptr(), n, w, h, w, w * h, temporal=True)
        for hd, out in zip(hds, outs):
            hd.stereo_download(n, out)
    for hd, out in zip(hds, outs):
        o = out[0]
        prev = None
        for p in range(n):
            kl, dl, kr, dr = want[p]
            cur = _check_pair(o, p, kl, dl, kr, dr, oracle)
            if prev is not None:
                tl, d1 = oracle.bf_match(cur[0], prev[0]); tr, d2 = oracle.bf_match(cur[1], prev[1])
                ql, hl = oracle.threshold_filter(tl, d1, len(prev[0]), 2.0, 30.0)
                qr, hr = oracle.threshold_filter(tr, d2, len(prev[1]), 2.0, 30.0)
                mp = dict(zip(ql.tolist(), hl.tolist()))
                circ = [(q, t) for q, t in zip(qr.tolist(), hr.tolist()) if mp.get(q) == t]
                nc = int(o["n_circ"][p])
                assert nc == len(circ), f"pair {p}"
                assert list(zip(o["circ_cur"][p, :nc].tolist(), o["circ_hist"][p, :nc].tolist())) == circ
            prev = cur
        assert o["n_circ"][0] > 0             # pair 0 of step 2 matched against pair 127 of step 1
        hd.close()


def test_sequence_sharded_over_two_handles_equals_sequential_full_slabs(fe, synth):
    """Config 4 in miniature: one EuRoC-shaped sequence cut into two contiguous frame blocks, each on its own handle
    (the per-GPU unit of pose-estimation_b200/shard.py), the second block re-processing frame start-1 for the temporal
    chain.  Every output array of every frame must equal the sequential run's — full slabs, not counts."""
    w, h, nf, ini, mn = synth.SHAPES["euroc"]
    n = 24
    ls, rs = synth.sequence("euroc", n, seed=3)
    seq = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * n)
    a = seq.stereo_frontend(ls, rs, temporal=True)
    cut = 11
    h0 = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * n)
    h1 = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * n)
    b0 = h0.stereo_frontend(ls[:cut], rs[:cut], temporal=True)
    b1 = h1.stereo_frontend(ls[cut - 1:], rs[cut - 1:], temporal=True)     # frame cut-1 recomputed locally
    keys = (("kp_l", "n_l"), ("desc_l", "n_l"), ("kp_r", "n_r"), ("desc_r", "n_r"), ("train_idx", "n_l"), ("dist", "n_l"),
            ("match_l", "n_match"), ("match_r", "n_match"), ("circ_cur", "n_circ"), ("circ_hist", "n_circ"))
    for p in range(n):
        b, q = (b0, p) if p < cut else (b1, p - cut + 1)
        for key, cnt in keys:
            m = int(a[cnt][p])
            assert m == int(b[cnt][q]), (p, cnt)
            assert a[key][p, :m].tobytes() == b[key][q, :m].tobytes(), (p, key)
    assert a["n_circ"][1:].min() > 5
    for hd in (seq, h0, h1):
        hd.close()


def test_unchanged_caller_path_async_extract_match_and_stereo_filter(fe, oracle, synth):
    """The reference's own call sequence with nothing fused (src/feature-tracker.cpp:153-164, 208-218): two extractor
    instances run concurrently (b2orb_extract_async on two handles from one thread = the two std::threads), then
    BFMatcher.match, then the host filter loops as b2orb_stereo_filter.  Must equal the fused b2orb_stereo_frontend and
    the oracle."""
    w, h, nf, ini, mn = synth.SHAPES["euroc"]
    ls, rs = synth.sequence("euroc", 2, seed=71)
    left = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    right = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    fused = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    for t in range(2):
        left.extract_async(ls[t]); right.extract_async(rs[t])
        kl, dl = left.extract_wait(); kr, dr = right.extract_wait()
        ko, do = ex(ls[t])
        assert same(kl, dl, ko, do)
        ti, di = left.match(dl, dr)
        f = left.stereo_filter(ti, di, kl, dl, kr, dr)
        o = fused.stereo_frontend(ls[t], rs[t], out=fused.alloc_stereo_out(1, kept=True))
        m = int(o["n_match"][0])
        assert m == len(f["match_l"]) > 20
        for k in ("match_l", "match_r", "pix_l", "pix_r", "cur_desc_l", "cur_desc_r"):
            assert np.array_equal(f[k], o[k][0, :m]), k
        ml, mr = oracle.stereo_filter(ti, di, kl, kr, 2.0, 30.0, 0.1)
        assert np.array_equal(f["match_l"], ml) and np.array_equal(f["match_r"], mr)
        assert np.array_equal(f["pix_l"], np.stack([kl["x"][ml], kl["y"][ml]], 1).astype(np.float64))
    # empty image / wait without async / empty sets
    left.extract_async(np.zeros((0, 0), np.uint8))
    assert len(left.extract_wait()[0]) == 0
    with pytest.raises(fe.B2Error):
        left.extract_wait()
    assert len(left.stereo_filter(np.zeros(0, np.int32), np.zeros(0, np.int32), kl[:0], dl[:0], kr, dr)["match_l"]) == 0
    for hd in (left, right, fused):
        hd.close()

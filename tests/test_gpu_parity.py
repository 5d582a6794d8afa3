"""GPU parity: the CUDA path (through the C ABI, include/b2orb.h) vs the CPU oracle, bit-exact.

Stage by stage (pyramid, blur, raw FAST corners, quadtree selection, angles, descriptors, keypoints)
and end to end (stereo match + filters, temporal circular matching) on seeded synthetic inputs at the
shapes of BASELINE.json's configs.  Bar: bit-exact for every integer/byte/index output AND for the float
fields (angle, x, y) — the north star allows 1e-4 rad on angles, we assert exact equality.
"""
import numpy as np
import pytest

from tests.conftest import pkg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def fe():
    return pkg("frontend")


def assert_extract_equal(got, ref, tag=""):
    gk, gd = got
    rk, rd = ref
    assert len(gk) == len(rk), f"{tag}: keypoint count {len(gk)} vs {len(rk)}"
    for f in ("octave", "x", "y", "response", "size", "angle", "class_id"):
        assert np.array_equal(gk[f], rk[f]), f"{tag}: field {f} differs at {np.flatnonzero(gk[f] != rk[f])[:8]}"
    assert np.array_equal(gd, rd), f"{tag}: descriptors differ on rows {np.flatnonzero((gd != rd).any(1))[:8]}"


CASES = [
    # name, w, h, nfeatures, ini, min, seed, passes
    ("euroc", 752, 480, 1000, 20, 7, 0, 3),
    ("euroc_yml800", 752, 480, 800, 20, 7, 1, 3),
    ("cfsd", 672, 376, 1000, 20, 7, 2, 3),
    ("kitti", 1241, 376, 2000, 12, 7, 3, 3),
    ("lowtex", 400, 300, 500, 20, 7, 4, 8),
    ("odd", 333, 211, 300, 20, 7, 5, 3),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_extract_stages_match_oracle(case, fe, oracle, synth):
    _, w, h, nf, ini, mn, seed, passes = case
    img = synth.canvas(w, h, seed, passes)[:h, :w].copy()
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    ref = ex(img)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    got = hd.extract(img)
    for lvl in range(8):
        assert np.array_equal(hd.debug_level(0, lvl), ex.level(lvl)), f"pyramid level {lvl}"
        rb = ex.level(lvl, blurred=True)
        if rb is not None:
            assert np.array_equal(hd.debug_level(0, lvl, blurred=True), rb), f"blurred level {lvl}"
        c = hd.debug_candidates(0, lvl)
        r = ex.raw_corners(lvl).astype(np.int32)
        cs = c[np.lexsort((c[:, 0], c[:, 1]))]
        rs = r[np.lexsort((r[:, 0], r[:, 1]))]
        assert cs.shape == rs.shape and np.array_equal(cs, rs), f"raw FAST corners level {lvl}"
    assert_extract_equal(got, ref, case[0])
    s, i, s2, i2, q = hd.tables()
    os_, oi, oq, _ = ex.tables()
    assert np.array_equal(s, os_) and np.array_equal(i, oi) and np.array_equal(q, oq)
    hd.close()


@pytest.mark.parametrize("mode", ["", "points"])
def test_quadtree_paths_agree_with_oracle(mode, fe, oracle, synth):
    """The quadtree has two sources of child counts (octree.cu): the count pyramid and the point passes
    (debug_set("octree_points", 1) forces the latter).  Both must reproduce DistributeOctTree; the second image packs
    all texture into a few small patches, so nodes below the pyramid's depth have to be split and the kernel redoes
    those levels with point passes on its own."""
    w, h, nf, ini, mn = 1241, 376, 2000, 12, 7
    dense = synth.canvas(w, h, 17, 3)[:h, :w].copy()
    clustered = np.full((h, w), 96, np.uint8)
    for (y, x) in ((30, 60), (150, 420), (250, 800), (60, 1100)):
        clustered[y:y + 90, x:x + 90] = dense[y:y + 90, x:x + 90]
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2).debug_set("octree_points", mode == "points")
    for tag, img in (("dense", dense), ("clustered", clustered)):
        ref = ex(img)
        assert len(ref[0]) > 500
        before = hd.debug_octree_fallbacks()
        assert_extract_equal(hd.extract(img), ref, f"{tag} [{mode or 'pyramid'}]")
        redone = hd.debug_octree_fallbacks() - before
        assert (redone > 0) == (tag == "clustered" and mode == ""), (tag, mode, redone)
    hd.close()


@pytest.mark.parametrize("opt", [("fast_band", 0), ("fast_band", 1), ("fast_band", 2), ("fast_band", 3), ("blur_band", 0), ("blur_band", 1),
                                 ("blur_band", 16), ("blur_band", 17),   # TMA-staged blur: only in the variants build
                                 ("fork_max", 0), ("pdl", 0),
                                 ("pdl", 1), ("export_max", 0),
                                 ("describe_tma", 0), ("describe_tma", 1)],   # descriptor windows by cp.async / by TMA boxes
                                ids=lambda o: f"{o[0]}={o[1]}")
def test_launch_time_choices_agree_with_oracle(opt, fe, oracle, synth):
    """Every launch-time choice the library makes by batch size (FAST band length, side-stream blur, programmatic
    launch chain, direct export, how the descriptor windows are staged) must give the oracle's bits: pinned one at a time with b2orb_debug_set on a
    low-texture image (both FAST thresholds in play)."""
    w, h = 400, 300
    ls, rs = synth.sequence((w, h), 2, seed=21, passes=8)
    hd = fe.Handle(500, 1.2, 8, 20, 7, w, h, max_images=4).debug_set(*opt)
    ex = oracle.OracleExtractor(500, 1.2, 8, 20, 7)
    out = hd.stereo_frontend(ls, rs, temporal=True)
    for p in range(2):
        for img, kk, dk, nk in ((ls[p], "kp_l", "desc_l", "n_l"), (rs[p], "kp_r", "desc_r", "n_r")):
            k, d = ex(img)
            n = int(out[nk][p])
            assert n == len(k) and n > 100
            assert out[kk][p, :n].tobytes() == k.tobytes()
            assert np.array_equal(out[dk][p, :n], d)
    hd.close()


@pytest.mark.parametrize("nfeat", [8000, 12000])
def test_stress_4k_pair_matches_oracle(nfeat, fe, oracle, synth):
    """BASELINE config 5: 3840x2160 stereo, 8000 features per eye, 8k x 8k brute-force Hamming per pair — the largest
    shape the reference configs name.  12000 features push the level-0 quadtree workspace past the shared-memory budget
    (global-memory workspace path of octree.cu)."""
    w, h, _, ini, mn = synth.SHAPES["4k"]
    ls, rs = synth.sequence("4k", 1, seed=2)
    hd = fe.Handle(nfeat, 1.2, 8, ini, mn, w, h, max_images=2)
    o = hd.stereo_frontend(ls, rs)
    ex = oracle.OracleExtractor(nfeat, 1.2, 8, ini, mn)
    kl, dl = ex(ls[0]); kr, dr = ex(rs[0])
    nl, nr, m = int(o["n_l"][0]), int(o["n_r"][0]), int(o["n_match"][0])
    assert nl >= nfeat and nr >= nfeat
    assert_extract_equal((o["kp_l"][0, :nl], o["desc_l"][0, :nl]), (kl, dl), "4k left")
    assert_extract_equal((o["kp_r"][0, :nr], o["desc_r"][0, :nr]), (kr, dr), "4k right")
    ti, di = oracle.bf_match(dl, dr)
    assert np.array_equal(o["train_idx"][0, :nl], ti) and np.array_equal(o["dist"][0, :nl], di)
    ml, mr = oracle.stereo_filter(ti, di, kl, kr, 2.0, 30.0, 0.1)
    assert m == len(ml) > 100 and np.array_equal(o["match_l"][0, :m], ml) and np.array_equal(o["match_r"][0, :m], mr)
    hd.close()


def test_extract_batch_and_resize_of_workspace(fe, oracle, synth):
    ls, rs = synth.sequence("cfsd", 3, seed=7)
    hd = fe.Handle(1000, 1.2, 8, 20, 7, 752, 480, max_images=6)
    res = hd.extract_batch(np.concatenate([ls, rs]))
    ex = oracle.OracleExtractor(1000, 1.2, 8, 20, 7)
    for k, img in enumerate(list(ls) + list(rs)):
        assert_extract_equal(res[k], ex(img), f"batch image {k}")
    # another size on the same handle (workspace is rebuilt), given as a strided view
    img = synth.canvas(500, 400, 9)[:400, :500]
    assert_extract_equal(hd.extract(img), ex(np.ascontiguousarray(img)), "strided 500x400")
    hd.close()


def test_empty_and_degenerate_inputs(fe, oracle):
    hd = fe.Handle(500, 1.2, 8, 20, 7, 256, 256, max_images=2)
    k, d = hd.extract(np.zeros((0, 0), np.uint8))
    assert len(k) == 0 and d.shape == (0, 32)
    flat = np.full((200, 256), 128, np.uint8)            # no corners anywhere
    k, d = hd.extract(flat)
    assert len(k) == 0
    ex = oracle.OracleExtractor(500, 1.2, 8, 20, 7)
    rng = np.random.default_rng(3)
    tiny = rng.integers(0, 256, (70, 90), dtype=np.uint8)  # upper levels have no FAST cells
    assert_extract_equal(hd.extract(tiny), ex(tiny), "tiny 90x70")
    hd.close()


def test_orbextractor_mirror_matches_oracle(fe, oracle, synth):
    img = synth.canvas(640, 360, 11)[:360, :640].copy()
    ext = fe.ORBextractor(700, 1.2, 8, 20, 7)
    kps, desc = ext(img, None)
    assert_extract_equal((kps, desc), oracle.OracleExtractor(700, 1.2, 8, 20, 7)(img), "ORBextractor mirror")
    assert ext.GetLevels() == 8 and abs(ext.GetScaleFactor() - 1.2) < 1e-6
    assert np.array_equal(ext.GetScaleFactors(), oracle.OracleExtractor(700, 1.2, 8, 20, 7).tables()[0])


@pytest.mark.parametrize("nq,nt", [(1, 1), (500, 700), (1000, 1000), (37, 2051), (2000, 129)])
def test_bf_match_matches_oracle(nq, nt, fe, oracle):
    rng = np.random.default_rng(nq * 7 + nt)
    protos = rng.integers(0, 256, (9, 32), dtype=np.uint8)
    q = protos[rng.integers(0, 9, nq)].copy(); t = protos[rng.integers(0, 9, nt)].copy()
    q[rng.integers(0, nq, nq // 2 + 1), rng.integers(0, 32, nq // 2 + 1)] ^= 0x10
    t[rng.integers(0, nt, nt // 2 + 1), rng.integers(0, 32, nt // 2 + 1)] ^= 0x01
    hd = fe.Handle(500, 1.2, 8, 20, 7, 128, 128, max_images=2)
    ti, di = hd.match(q, t)
    oi, od = oracle.bf_match(q, t)
    assert np.array_equal(ti, oi) and np.array_equal(di, od)
    ti, di = hd.match(q, np.zeros((0, 32), np.uint8))
    assert (ti == -1).all()
    hd.close()


@pytest.mark.parametrize("nq,nt", [(5, 7), (300, 9), (257, 1025), (33, 3000)])
def test_bf_match_extreme_rows(nq, nt, fe, oracle):
    """Saturated descriptors (all zeros / all ones: distances 0 and 256), random rows, ragged tile sizes."""
    rng = np.random.default_rng(nq + 31 * nt)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8); t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    q[0] = 0; q[nq // 2] = 0xFF; q[-1] = 0xFF
    t[nt // 3] = 0; t[-1] = 0xFF; t[0] = 0xFF
    hd = fe.Handle(500, 1.2, 8, 20, 7, 128, 128, max_images=2)
    for qq, tt in ((q, t), (q, t[-1:]), (q[:1], t[:1]), (np.zeros((3, 32), np.uint8), np.full((2, 32), 0xFF, np.uint8))):
        ti, di = hd.match(qq, tt)
        oi, od = oracle.bf_match(qq, tt)
        assert np.array_equal(ti, oi) and np.array_equal(di, od)
    assert hd.match(np.zeros((3, 32), np.uint8), np.full((2, 32), 0xFF, np.uint8))[1].tolist() == [256, 256, 256]
    hd.close()


@pytest.mark.parametrize("shape,npairs", [("euroc", 3), ("kitti", 2), ("cfsd", 4)])
def test_stereo_frontend_matches_oracle(shape, npairs, fe, oracle, synth):
    """internalMatch (extract L/R + BF match + min-dist/row filter) and the temporal circular matching."""
    w, h, nf, ini, mn = synth.SHAPES[shape]
    ls, rs = synth.sequence(shape, npairs, seed=21)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs)
    o = hd.stereo_frontend(ls, rs, temporal=True)
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    prev = None
    for p in range(npairs):
        kl, dl = ex(ls[p]); kr, dr = ex(rs[p])
        nl, nr = int(o["n_l"][p]), int(o["n_r"][p])
        assert_extract_equal((o["kp_l"][p, :nl], o["desc_l"][p, :nl]), (kl, dl), f"pair {p} left")
        assert_extract_equal((o["kp_r"][p, :nr], o["desc_r"][p, :nr]), (kr, dr), f"pair {p} right")
        ti, di = oracle.bf_match(dl, dr)
        assert np.array_equal(o["train_idx"][p, :nl], ti) and np.array_equal(o["dist"][p, :nl], di)
        ml, mr = oracle.stereo_filter(ti, di, kl, kr, 2.0, 30.0, 0.1)
        m = int(o["n_match"][p])
        assert m == len(ml) and m > 20, f"pair {p}: {m} matches"
        assert np.array_equal(o["match_l"][p, :m], ml) and np.array_equal(o["match_r"][p, :m], mr)
        cur_l, cur_r = dl[ml], dr[mr]
        if prev is None:
            assert o["n_circ"][p] == 0
        else:
            tl, d1 = oracle.bf_match(cur_l, prev[0]); tr, d2 = oracle.bf_match(cur_r, prev[1])
            ql, hl = oracle.threshold_filter(tl, d1, len(prev[0]), 2.0, 30.0)
            qr, hr = oracle.threshold_filter(tr, d2, len(prev[1]), 2.0, 30.0)
            mp = dict(zip(ql.tolist(), hl.tolist()))
            circ = [(q, t) for q, t in zip(qr.tolist(), hr.tolist()) if mp.get(q) == t]
            nc = int(o["n_circ"][p])
            assert nc == len(circ) and nc > 5
            assert list(zip(o["circ_cur"][p, :nc].tolist(), o["circ_hist"][p, :nc].tolist())) == circ
        prev = (cur_l, cur_r)
    # history carries over to the next call: pair 0 of a second call matches against the last pair
    o2 = hd.stereo_frontend(ls[:1], rs[:1], temporal=True)
    assert o2["n_circ"][0] > 0
    hd.reset_history()
    assert hd.stereo_frontend(ls[:1], rs[:1], temporal=True)["n_circ"][0] == 0
    hd.close()


def _stereo_equal(a, b, npairs):
    for k in ("n_l", "n_r", "n_match", "n_circ"):
        assert np.array_equal(a[k][:npairs], b[k][:npairs]), k
    for p in range(npairs):
        for key, cnt in (("kp_l", "n_l"), ("desc_l", "n_l"), ("kp_r", "n_r"), ("desc_r", "n_r"), ("train_idx", "n_l"),
                         ("dist", "n_l"), ("match_l", "n_match"), ("match_r", "n_match"), ("circ_cur", "n_circ"),
                         ("circ_hist", "n_circ")):
            n = int(a[cnt][p])
            assert np.array_equal(a[key][p, :n], b[key][p, :n]), f"pair {p}: {key}"


@pytest.mark.parametrize("chunk,lanes", [(2, 3), (3, 2)])
def test_pipelined_host_path_equals_single_call(chunk, lanes, fe, synth):
    """A long host batch is cut into chunks that rotate over internal lanes (streams): same results as one call,
    including the temporal chain across chunk, lane and call boundaries."""
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    ls, rs = synth.sequence("cfsd", 7, seed=33)
    ref = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=14).debug_set("pipe_chunk", 0)
    pip = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=14).debug_set("pipe_chunk", chunk).debug_set("pipe_lanes", lanes)
    a, b = ref.stereo_frontend(ls, rs, temporal=True), pip.stereo_frontend(ls, rs, temporal=True)
    assert a["n_circ"][1:].min() > 5
    _stereo_equal(a, b, 7)
    # second call: history comes from the lane that ran the last chunk; then a short (un-pipelined) call on the root
    _stereo_equal(ref.stereo_frontend(ls[:5], rs[:5], temporal=True), pip.stereo_frontend(ls[:5], rs[:5], temporal=True), 5)
    a1, b1 = ref.stereo_frontend(ls[6:], rs[6:], temporal=True), pip.stereo_frontend(ls[6:], rs[6:], temporal=True)
    assert a1["n_circ"][0] > 0
    _stereo_equal(a1, b1, 1)
    _stereo_equal(ref.stereo_frontend(ls, rs, temporal=True), pip.stereo_frontend(ls, rs, temporal=True), 7)
    pip.reset_history(); ref.reset_history()
    b2 = pip.stereo_frontend(ls, rs, temporal=True)
    assert b2["n_circ"][0] == 0
    _stereo_equal(ref.stereo_frontend(ls, rs, temporal=True), b2, 7)
    assert pip.launch_count > ref.launch_count
    ref.close(); pip.close()


@pytest.mark.parametrize("npairs", [1, 3])
def test_direct_export_to_pinned_outputs_equals_copy_path(npairs, fe, synth):
    """Latency path: with pinned output arrays one kernel stores the valid prefixes straight into host memory instead
    of ~14 device-to-host copies.  Same results as the copy path (pageable arrays), temporal chain and triangulated
    points included; only entries [0, n) are defined on either path."""
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    ls, rs = synth.sequence("cfsd", 2 * npairs, seed=5)
    fx = 0.58 * w
    P1 = [[fx, 0, 0.49 * w, 0], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
    P2 = [[fx, 0, 0.49 * w, 0.537 * fx], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
    a_hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs)
    b_hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs)
    for hd in (a_hd, b_hd):
        hd.set_projection(P1, P2, 20.0)
    pinned = b_hd.alloc_stereo_out(npairs, pinned=True)
    for arr in pinned[0].values():
        arr.view(np.uint8)[...] = 0xAB
    for rep in range(2):                                   # second call: temporal history + cached export plan
        i = rep * npairs
        a = a_hd.stereo_frontend(ls[i:i + npairs], rs[i:i + npairs], temporal=True)
        b = b_hd.stereo_frontend(ls[i:i + npairs], rs[i:i + npairs], temporal=True, out=pinned)
        _stereo_equal(a, b, npairs)
        for p in range(npairs):
            n = int(a["n_match"][p])
            assert n > 10
            assert np.array_equal(a["points4d"][p, :, :n].view(np.uint64), b["points4d"][p, :, :n].view(np.uint64))
            assert np.array_equal(a["point_cam"][p, :n].view(np.uint64), b["point_cam"][p, :n].view(np.uint64))
            assert np.array_equal(a["depth_flags"][p, :n], b["depth_flags"][p, :n])
            nl = int(a["n_l"][p])
            assert np.all(b["desc_l"][p, nl:] == 0xAB)     # nothing beyond the valid prefix is touched
    assert a["n_circ"].min() > 5
    a_hd.close(); b_hd.close()


def test_direct_export_with_capacity_not_a_multiple_of_four(fe, synth):
    """750 features -> capacity 782: the 1-byte depth-flag rows of pairs 1.. start at addresses that are not word
    aligned; the export kernel must fall back to byte copies there instead of faulting."""
    w, h, nf, ini, mn = 672, 376, 750, 20, 7
    npairs = 3
    ls, rs = synth.sequence("cfsd", npairs, seed=15)
    fx = 0.58 * w
    P1 = [[fx, 0, 0.49 * w, 0], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
    P2 = [[fx, 0, 0.49 * w, 0.537 * fx], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
    a_hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs)
    b_hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * npairs)
    assert a_hd.keypoint_capacity % 4 != 0
    for hd in (a_hd, b_hd):
        hd.set_projection(P1, P2, 20.0)
    pinned = b_hd.alloc_stereo_out(npairs, pinned=True)
    a = a_hd.stereo_frontend(ls, rs, temporal=True)
    b = b_hd.stereo_frontend(ls, rs, temporal=True, out=pinned)
    _stereo_equal(a, b, npairs)
    for p in range(npairs):
        n = int(a["n_match"][p])
        assert n > 10 and np.array_equal(a["depth_flags"][p, :n], b["depth_flags"][p, :n])
        assert np.array_equal(a["points4d"][p, :, :n].view(np.uint64), b["points4d"][p, :, :n].view(np.uint64))
    a_hd.close(); b_hd.close()


def test_bench_sized_batch_matches_oracle_and_single_calls(fe, oracle, synth):
    """The throughput configuration itself: 136 pairs in ONE call (default pipeline: chunks of 64 + 64 + 8 over the
    internal lanes, 64-row FAST bands, no programmatic launch chain, copy-path outputs).  Every 9th pair is checked
    against the oracle, and the whole batch against the same pairs sent one call at a time (16-row bands, PDL,
    side-stream blur): batching must not change a bit."""
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    n = 136
    ls, rs = synth.sequence("cfsd", n, seed=77)
    big = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2 * n)
    one = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    a = big.stereo_frontend(ls, rs, temporal=True)
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    for p in range(0, n, 9):
        kl, dl = ex(ls[p]); kr, dr = ex(rs[p])
        nl, nr = int(a["n_l"][p]), int(a["n_r"][p])
        assert nl == len(kl) and nr == len(kr)
        assert a["kp_l"][p, :nl].tobytes() == kl.tobytes() and a["kp_r"][p, :nr].tobytes() == kr.tobytes()
        assert np.array_equal(a["desc_l"][p, :nl], dl) and np.array_equal(a["desc_r"][p, :nr], dr)
        ti, di = oracle.bf_match(dl, dr)
        assert np.array_equal(a["train_idx"][p, :nl], ti) and np.array_equal(a["dist"][p, :nl], di)
    for p in range(n):
        b = one.stereo_frontend(ls[p], rs[p], temporal=True)
        for key, cnt in (("kp_l", "n_l"), ("desc_l", "n_l"), ("kp_r", "n_r"), ("desc_r", "n_r"), ("train_idx", "n_l"),
                         ("dist", "n_l"), ("match_l", "n_match"), ("match_r", "n_match"), ("circ_cur", "n_circ"),
                         ("circ_hist", "n_circ")):
            m = int(a[cnt][p])
            assert m == int(b[cnt][0]), f"pair {p}: {cnt}"
            assert np.array_equal(a[key][p, :m], b[key][0, :m]), f"pair {p}: {key}"
    assert a["n_circ"][1:].min() > 5
    big.close(); one.close()


def test_temporal_track_entry_point(fe, oracle):
    rng = np.random.default_rng(5)
    hist_l = rng.integers(0, 256, (300, 32), dtype=np.uint8); hist_r = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    pick = rng.permutation(300)[:200]
    cur_l, cur_r = hist_l[pick].copy(), hist_r[pick].copy()
    cur_l[::3, 0] ^= 0x3; cur_r[::5, 1] ^= 0x7; cur_r[::7] = rng.integers(0, 256, (len(cur_r[::7]), 32), dtype=np.uint8)
    hd = fe.Handle(500, 1.2, 8, 20, 7, 128, 128, max_images=2)
    cc, ch, nl, nr = hd.temporal_track(cur_l, cur_r, hist_l, hist_r)
    tl, d1 = oracle.bf_match(cur_l, hist_l); tr, d2 = oracle.bf_match(cur_r, hist_r)
    ql, hl = oracle.threshold_filter(tl, d1, 300, 2.0, 30.0); qr, hr = oracle.threshold_filter(tr, d2, 300, 2.0, 30.0)
    mp = dict(zip(ql.tolist(), hl.tolist()))
    circ = [(q, t) for q, t in zip(qr.tolist(), hr.tolist()) if mp.get(q) == t]
    assert (nl, nr) == (len(ql), len(qr))
    assert list(zip(cc.tolist(), ch.tolist())) == circ and len(circ) > 50
    assert hd.temporal_track(cur_l, cur_r, hist_l[:0], hist_r[:0])[0].size == 0
    hd.close()


def test_error_codes(fe):
    with pytest.raises(fe.B2Error):
        fe.Handle(1000, 0.9, 8, 20, 7, 100, 100)          # scaleFactor must be > 1
    hd = fe.Handle(500, 1.2, 8, 20, 7, 128, 128, max_images=2)
    with pytest.raises(fe.B2Error) as e:
        hd.extract(np.zeros((300, 300), np.uint8))        # larger than configured
    assert e.value.code == fe.ERR_INVALID
    with pytest.raises(fe.B2Error) as e:
        hd.extract_batch(np.zeros((3, 100, 100), np.uint8))  # more than max_images
    assert e.value.code == fe.ERR_CAPACITY
    hd.close()


@pytest.mark.parametrize("shape", ["kitti", (401, 300), (640, 480)], ids=["kitti_1241", "odd_401", "even_640"])
def test_pinned_host_images_read_by_the_ingest_kernel(shape, fe, synth):
    """One-pair calls whose images sit in PINNED host memory are read by a kernel straight over the bus (no DMA copy + repack);
    the result must be the one of the DMA path (pageable images, or b2orb_debug_set(h, "ingest_host", 0)): odd and even
    widths, image bases that are not 16-byte aligned (image 1 of a pinned stack of odd-sized images), rows with a pitch
    larger than the width, and b2orb_extract on a single pinned image."""
    import ctypes as C
    if isinstance(shape, str):
        w, h, nf, ini, mn = synth.SHAPES[shape]
    else:
        (w, h), nf, ini, mn = shape, 600, 20, 7
    ls, rs = synth.sequence(shape, 2, seed=33)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    ref = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2).debug_set("ingest_host", 0)
    pl = fe.pinned_empty(ls.shape); pl[...] = ls
    pr = fe.pinned_empty(rs.shape); pr[...] = rs
    keys = ("n_l", "n_r", "n_match", "kp_l", "kp_r", "desc_l", "desc_r", "match_l", "match_r")
    for i in range(2):                                    # i = 1: base address = i * w * h, odd for odd shapes
        want = ref.stereo_frontend(ls[i], rs[i])
        l0 = hd.launch_count
        got = hd.stereo_frontend(pl[i], pr[i])
        if i:                                             # (call 0 also launches the one-off problem-table kernel)
            # one ingest launch instead of two re-pitching launches (rows that already have the staging pitch: plain DMA)
            assert hd.launch_count - l0 == ref_launches(ref, ls[i], rs[i]) - (1 if w % 32 else 0)
        for k in keys:
            assert np.array_equal(got[k], want[k]), (i, k)
    # pitched pinned rows (step > width), through the C ABI directly
    step = w + 37
    pp = fe.pinned_empty((2, h, step)); pp[...] = 0xAB
    pp[0, :, :w] = ls[0]; pp[1, :, :w] = rs[0]
    out = hd.alloc_stereo_out(1)
    hd.check(hd.L.b2orb_stereo_frontend(hd.h, pp[0].ctypes.data, pp[1].ctypes.data, 1, w, h, step, step * h, 0, C.byref(out[1])))
    want = ref.stereo_frontend(ls[0], rs[0])
    for k in keys:
        assert np.array_equal(out[0][k], want[k]), ("pitched", k)
    # single image
    k1, d1 = hd.extract(pl[1]); k0, d0 = ref.extract(ls[1])
    assert k1.tobytes() == k0.tobytes() and np.array_equal(d1, d0)
    hd.close(); ref.close()


def ref_launches(hd, l, r):
    l0 = hd.launch_count
    hd.stereo_frontend(l, r)
    return hd.launch_count - l0

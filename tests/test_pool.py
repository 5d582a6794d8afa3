"""Next row f3 — the device-resident feature pool (src/feature-tracker.cpp:231-328 matching parts with useRANSAC=false,
:330-412 featurePoolUpdate bookkeeping) against the oracle's restatement on the same std::unordered_map.

CPU: properties of the oracle pool itself (ageing arithmetic, erase above maxFeatureAge, container-order rebuild).
GPU: a 30-frame sequence with a keyframe every 3rd frame and maxFeatureAge = 8: feature ids, ages, pool order, pool
descriptor rows, circular matches, _matchedFeatureIDs and _curFeatureMask equal the oracle's on every frame, for the
stand-alone track (b2orb_pool_track), the fused one (b2orb_stereo_frontend(temporal=2)) and with the depth gate."""
import numpy as np
import pytest

from tests.conftest import pkg


def test_oracle_pool_ageing_and_order(oracle):
    rng = np.random.default_rng(1)
    base_l = rng.integers(0, 256, (300, 32), dtype=np.uint8); base_r = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    pool = oracle.OracleFeaturePool(max_feature_age=4)
    # frame 0: empty pool -> everything inserted with age 0, ids 0..n-1 in order, no matching (:237-240)
    t = pool.track(base_l[:100], base_r[:100])
    assert len(t["circ_cur"]) == 0 and t["mask"].all() and t["n_hist"] == 0
    assert pool.update(False) == (100, 0)                       # `_features.empty() ||` keyframe
    ids, ages, hl, hr = pool.state()
    assert ids.tolist() == list(range(100)) and not ages.any() and np.array_equal(hl, base_l[:100])
    # frame 1 (not a keyframe): identical descriptors match circularly, pool unchanged
    t = pool.track(base_l[:100], base_r[:100])
    assert t["circ_cur"].tolist() == list(range(100)) and not t["mask"].any()
    assert [int(ids[j]) for j in t["circ_hist"]] == t["matched_ids"].tolist() == list(range(100))
    assert pool.update(False) == (0, 0) and np.array_equal(pool.state()[0], ids)
    # keyframe with half matched: matched age +1, unmatched +2, new ones appended after the rebuilt rows
    cur_l = np.concatenate([base_l[:50], base_l[200:230]]); cur_r = np.concatenate([base_r[:50], base_r[200:230]])
    t = pool.track(cur_l, cur_r)
    assert t["mask"].tolist() == [0] * 50 + [1] * 30
    ins, er = pool.update(True)
    assert (ins, er) == (30, 0)
    ids2, ages2, hl2, _ = pool.state()
    age = dict(zip(ids2.tolist(), ages2.tolist()))
    assert all(age[i] == 1 for i in range(50)) and all(age[i] == 2 for i in range(50, 100)) and all(age[i] == 0 for i in range(100, 130))
    assert ids2[-30:].tolist() == list(range(100, 130)) and np.array_equal(hl2[-30:], base_l[200:230])
    assert sorted(ids2[:100].tolist()) == list(range(100))      # rebuilt part: container iteration order, same members
    # two more keyframes without matches: 2 -> 4 -> 6 > maxFeatureAge: erased
    far_l = rng.integers(0, 256, (5, 32), dtype=np.uint8)
    pool.track(far_l, far_l); pool.update(True, accept=np.zeros(5, np.uint8))
    pool.track(far_l, far_l); ins, er = pool.update(True, accept=np.zeros(5, np.uint8))
    # ages before: 1 / 2 / 0 -> 3 / 4 / 2 -> 5 / 6 / 4: the first hundred are erased, the thirty newer ones stay
    assert ins == 0 and er == 100 and sorted(pool.state()[0].tolist()) == list(range(100, 130))


def _run_sequence(fe, oracle, synth, fused, with_depth):
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    nframes, kf_every, max_age = 30, 3, 8
    ls, rs = synth.sequence("cfsd", nframes, seed=40)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    if with_depth:
        fx = 0.58 * w
        P1 = [[fx, 0, 0.49 * w, 0], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
        P2 = [[fx, 0, 0.49 * w, 0.537 * fx], [0, fx, 0.49 * h, 0], [0, 0, 1, 0]]
        hd.set_projection(P1, P2, 15.0)
    hd.pool_reset(max_age)
    pool = oracle.OracleFeaturePool(2.0, 30.0, max_age)
    out = hd.alloc_stereo_out(1, kept=True)
    seen_circ = seen_erase = seen_gate = 0
    for t in range(nframes):
        o = hd.stereo_frontend(ls[t], rs[t], temporal=2 if fused else 0, out=out)
        m = int(o["n_match"][0])
        cur_l, cur_r = o["cur_desc_l"][0, :m], o["cur_desc_r"][0, :m]
        assert np.array_equal(cur_l, o["desc_l"][0][o["match_l"][0, :m]]) and np.array_equal(cur_r, o["desc_r"][0][o["match_r"][0, :m]])
        kl, kr = o["kp_l"][0][o["match_l"][0, :m]], o["kp_r"][0][o["match_r"][0, :m]]
        assert np.array_equal(o["pix_l"][0, :m], np.stack([kl["x"], kl["y"]], 1).astype(np.float64))
        assert np.array_equal(o["pix_r"][0, :m], np.stack([kr["x"], kr["y"]], 1).astype(np.float64))
        want = pool.track(cur_l, cur_r)
        got = hd.pool_last_track() if fused else hd.pool_track(0)
        for k in ("circ_cur", "circ_hist", "mask", "matched_ids"):
            assert np.array_equal(got[k], want[k]), (t, k)
        assert (got["n_left"], got["n_right"]) == (want["n_left"], want["n_right"]), t
        if fused:
            nc = int(o["n_circ"][0])
            assert nc == len(want["circ_cur"]) and np.array_equal(o["circ_cur"][0, :nc], want["circ_cur"])
            assert np.array_equal(o["cur_mask"][0, :m], want["mask"])
        accept = None
        if with_depth:
            accept = o["depth_flags"][0, :m] & 1
            seen_gate += int((accept == 0).sum())
        kf = t % kf_every == 0
        new_ids, erased = hd.pool_update(kf)                   # depth gate taken from the device flags when a projection is set
        ins_o, er_o = pool.update(kf, accept)
        assert (len(new_ids), erased) == (ins_o, er_o), t
        ids, ages, hl, hr = hd.pool_state()
        oi, oa, ol, orr = pool.state()
        assert np.array_equal(ids, oi) and np.array_equal(ages, oa), t
        assert np.array_equal(hl, ol) and np.array_equal(hr, orr), t
        if ins_o:
            assert np.array_equal(new_ids, oi[-ins_o:])
        seen_circ += len(want["circ_cur"]); seen_erase += er_o
    assert seen_circ > 300 and seen_erase > (0 if with_depth else 50) and hd.pool_size == len(oi) > 15
    assert not with_depth or seen_gate > 5
    hd.close()


@pytest.mark.gpu
@pytest.mark.parametrize("fused,with_depth", [(False, False), (True, False), (True, True)],
                         ids=["pool_track", "fused_temporal2", "fused_depth_gate"])
def test_device_pool_sequence_equals_oracle(fused, with_depth, oracle, synth):
    _run_sequence(pkg("frontend"), oracle, synth, fused, with_depth)


@pytest.mark.gpu
def test_pool_batch_against_the_same_pool_and_reset(oracle, synth):
    """temporal=2 on a batch: every pair of the call is matched against the SAME pool (no update in between), through
    the pipelined host path as well; b2orb_pool_reset empties it again."""
    fe = pkg("frontend")
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    ls, rs = synth.sequence("cfsd", 6, seed=41)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=12).debug_set("pipe_chunk", 2)
    hd.pool_reset(8)
    o = hd.stereo_frontend(ls[:1], rs[:1], temporal=2)
    assert o["n_circ"][0] == 0
    new_ids, _ = hd.pool_update(True)
    ids, ages, hl, hr = hd.pool_state()
    assert len(new_ids) == int(o["n_match"][0]) == len(ids) > 20
    out = hd.alloc_stereo_out(6, kept=True)
    o = hd.stereo_frontend(ls, rs, temporal=2, out=out)       # 3 chunks over the lanes
    for p in range(6):
        m = int(o["n_match"][p])
        cl, cr = o["desc_l"][p][o["match_l"][p, :m]], o["desc_r"][p][o["match_r"][p, :m]]
        tl, d1 = oracle.bf_match(cl, hl); tr, d2 = oracle.bf_match(cr, hr)
        ql, ml = oracle.threshold_filter(tl, d1, len(hl), 2.0, 30.0); qr, mr = oracle.threshold_filter(tr, d2, len(hr), 2.0, 30.0)
        mp = dict(zip(ql.tolist(), ml.tolist()))
        circ = [(q, t) for q, t in zip(qr.tolist(), mr.tolist()) if mp.get(q) == t]
        nc = int(o["n_circ"][p])
        assert list(zip(o["circ_cur"][p, :nc].tolist(), o["circ_hist"][p, :nc].tolist())) == circ and (nc > 0 or p > 2)
        mask = np.ones(m, np.uint8); mask[[q for q, _ in circ]] = 0
        assert np.array_equal(o["cur_mask"][p, :m], mask)
    assert int(o["n_circ"][0]) == int(o["n_match"][0])        # frame 0 against its own features: all circular
    # the small result set (what the reference keeps of a frame) into pinned arrays: written by the lanes themselves
    small = hd.alloc_stereo_out(6, pinned=True, full=False, kept=True)
    for a in small[0].values():
        a.view(np.uint8)[...] = 0xCD
    s = hd.stereo_frontend(ls, rs, temporal=2, out=small)
    assert "kp_l" not in s and np.array_equal(s["n_match"], o["n_match"]) and np.array_equal(s["n_circ"], o["n_circ"])
    for p in range(6):
        m, nc = int(o["n_match"][p]), int(o["n_circ"][p])
        for k, n in (("pix_l", m), ("pix_r", m), ("cur_desc_l", m), ("cur_desc_r", m), ("cur_mask", m), ("match_l", m),
                     ("match_r", m), ("circ_cur", nc), ("circ_hist", nc)):
            assert np.array_equal(s[k][p, :n], o[k][p, :n]), (p, k)
        assert (s["cur_desc_l"][p, m:] == 0xCD).all()         # only the valid prefix crosses the bus
    hd.pool_reset(8)
    assert hd.pool_size == 0 and hd.stereo_frontend(ls[:1], rs[:1], temporal=2)["n_circ"][0] == 0
    hd.close()

"""Next row f2: the cfsd live ingest of src/cfsd-state-estimation.cpp:117-127 — CV_8UC4 side-by-side frame
-> cv::resize(INTER_LINEAR) -> cvtColor(BGR2GRAY) -> left / right split.

Known answers: tests/golden/ingest.json = SHA-256 of REAL cv2 outputs (tools/gen_golden.py).  CPU: the oracle reproduces
them.  GPU: the fused kernel reproduces them through b2orb_cfsd_ingest, and b2orb_stereo_frontend_packed equals the plain
front end on the oracle's grey eyes (also with rectification on)."""
import hashlib
import json
import os

import numpy as np
import pytest

from tests.conftest import pkg

GOLD = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ingest.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def load(name, synth):
    g = GOLD[name]
    w, h, nf, ini, mn, seed, passes, frame = g["params"]
    cv = synth.canvas(w, h, seed, passes)
    left, right = synth.stereo_pair(cv, w, h, frame, seed)
    fr = synth.packed_frame(left, right, g["src"][0], g["src"][1], seed)
    assert sha(fr) == g["frame_sha"], "synthetic frame generator drifted"
    return g, fr, (w, h, nf, ini, mn)


@pytest.mark.parametrize("name", sorted(GOLD))
def test_oracle_ingest_reproduces_cv2_known_answers(name, oracle, synth):
    g, fr, (w, h, *_) = load(name, synth)
    gl, gr = oracle.cfsd_ingest(fr, 2 * w, h)
    assert gl[50, ::16].tolist() == g["gray_row50_l"]
    assert sha(gl) == g["gray_sha_l"] and sha(gr) == g["gray_sha_r"]


def test_oracle_ingest_matches_cv2_on_random_frames(oracle):
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(8)
    for (sw, sh, dw, dh) in ((640, 200, 334, 102), (300, 111, 300, 111), (513, 97, 200, 64)):
        fr = rng.integers(0, 256, (sh, sw, 4), dtype=np.uint8)
        g = cv2.cvtColor(cv2.resize(fr, (dw, dh)), cv2.COLOR_BGR2GRAY)
        gl, gr = oracle.cfsd_ingest(fr, dw, dh)
        assert np.array_equal(gl, g[:, :dw // 2]) and np.array_equal(gr, g[:, dw // 2:])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(GOLD))
def test_cuda_ingest_reproduces_cv2_known_answers(name, synth):
    fe = pkg("frontend")
    g, fr, (w, h, nf, ini, mn) = load(name, synth)
    hd = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=2)
    gl, gr = hd.cfsd_ingest(fr, w, h)
    assert sha(gl) == g["gray_sha_l"] and sha(gr) == g["gray_sha_r"]
    hd.close()


@pytest.mark.gpu
@pytest.mark.parametrize("rectify", [False, True])
def test_packed_frontend_equals_plain_frontend_on_oracle_eyes(rectify, oracle, synth):
    fe = pkg("frontend")
    w, h, nf, ini, mn = synth.SHAPES["cfsd"]
    ls, rs = synth.sequence("cfsd", 3, seed=4)
    frames = np.stack([synth.packed_frame(ls[i], rs[i], 2560, 720, seed=i) for i in range(3)])
    eyes = [oracle.cfsd_ingest(f, 2 * w, h) for f in frames]
    gl = np.stack([e[0] for e in eyes]); gr = np.stack([e[1] for e in eyes])
    plain = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=6)
    packed = fe.Handle(nf, 1.2, 8, ini, mn, w, h, max_images=6)
    if rectify:
        maps = [synth.rectification_maps(w, h, e) for e in (0, 1)]
        for x in (plain, packed):
            x.set_rectification(maps[0][0], maps[0][1], maps[1][0], maps[1][1])
        assert np.array_equal(packed.cfsd_ingest(frames[0], w, h)[1], oracle.remap(gr[0], *maps[1]))
    a = plain.stereo_frontend(gl, gr, temporal=True)
    b = packed.stereo_frontend_packed(frames, w, h, temporal=True)
    assert a["n_l"].min() > 100
    for k in ("n_l", "n_r", "n_match", "n_circ"):
        assert np.array_equal(a[k], b[k]), k
    for p in range(3):
        for key, cnt in (("kp_l", "n_l"), ("desc_l", "n_l"), ("kp_r", "n_r"), ("desc_r", "n_r"), ("train_idx", "n_l"),
                         ("dist", "n_l"), ("match_l", "n_match"), ("match_r", "n_match"), ("circ_cur", "n_circ")):
            n = int(a[cnt][p])
            assert np.array_equal(a[key][p, :n], b[key][p, :n]), f"pair {p}: {key}"
    with pytest.raises(fe.B2Error):
        packed.stereo_frontend_packed(np.zeros((4, 8, 8, 4), np.uint8), w, h)   # 2*4 > max_images
    plain.close(); packed.close()

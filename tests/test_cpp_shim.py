"""The C-ABI boundary and the C++ drop-in shim (include/b2orb_shim/ORBextractor.h).

CPU: libb2orb.so loads and exports every function include/b2orb.h declares; the shim compiles against a mock of the
OpenCV headers (this image has no OpenCV C++) and links.  GPU: the compiled demo drives two ORB_SLAM2::ORBextractor
instances from two threads + the BFMatcher replacement, exactly like src/feature-tracker.cpp:133-164, and its output
equals the oracle's."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

from tests.conftest import ROOT, pkg

HDR = os.path.join(ROOT, "include", "b2orb.h")
SHIM_DIR = os.path.join(ROOT, "include", "b2orb_shim")
MOCK = os.path.join(ROOT, "tests", "mock_opencv")
DEMO_SRC = os.path.join(ROOT, "tests", "cpp", "shim_demo.cpp")


def declared_functions():
    txt = re.sub(r"/\*.*?\*/", "", open(HDR).read(), flags=re.S)
    return sorted(set(re.findall(r"\b(b2orb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    fe = pkg("frontend")
    L = ctypes.CDLL(fe.LIB_PATH)
    names = declared_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b2orb.h but not exported"
    assert L.b2orb_abi_version() == 2
    # no CUDA device here (CPU suite): creation must fail loudly, never fall back
    import torch
    if not torch.cuda.is_available():
        with pytest.raises(fe.B2Error):
            fe.Handle(500, 1.2, 8, 20, 7, 128, 128)


def test_keypoint_struct_layout_matches_cv_keypoint():
    fe = pkg("frontend")
    assert fe.KP_DTYPE.itemsize == 28 and fe.KP_DTYPE.names == ("x", "y", "size", "angle", "response", "octave", "class_id")
    assert ctypes.sizeof(fe.B2Config) == 4 * 8 + 4 * 7 + 4 * 3


def build_demo(tmp_path):
    fe = pkg("frontend")
    exe = str(tmp_path / "shim_demo")
    libdir = os.path.dirname(fe.LIB_PATH)
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-Wall", "-I" + MOCK, "-I" + SHIM_DIR, DEMO_SRC, "-o", exe,
                           "-L" + libdir, "-lb2orb", "-Wl,-rpath," + libdir, "-pthread"])
    return exe


def test_shim_compiles_and_links_with_reference_signatures(tmp_path):
    build_demo(tmp_path)
    src = open(os.path.join(SHIM_DIR, "ORBextractor.h")).read()
    # the public surface of include/ORBextractor.h:45-85
    for needle in ("namespace ORB_SLAM2", "class ORBextractor", "int nfeatures_, float scaleFactor_, int nlevels_",
                   "cv::InputArray _image, cv::InputArray", "std::vector<cv::KeyPoint>& _keypoints", "cv::OutputArray _descriptors",
                   "GetLevels", "GetScaleFactor", "GetScaleFactors", "GetInverseScaleFactors", "GetScaleSigmaSquares",
                   "GetInverseScaleSigmaSquares", "mvImagePyramid"):
        assert needle in src, needle


@pytest.mark.gpu
def test_shim_demo_matches_oracle(tmp_path, oracle, synth):
    exe = build_demo(tmp_path)
    w, h, nf, ini, mn = 480, 320, 400, 20, 7
    ls, rs = synth.sequence((w, h), 1, seed=13)
    ls[0].tofile(tmp_path / "l.raw"); rs[0].tofile(tmp_path / "r.raw")
    out = tmp_path / "out.txt"
    subprocess.check_call([exe, str(w), str(h), str(nf), str(ini), str(mn), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(out)])
    lines = open(out).read().split("\n")
    nl, nr, nm, levels, sf7 = lines[0].split()
    ex = oracle.OracleExtractor(nf, 1.2, 8, ini, mn)
    kl, dl = ex(ls[0]); kr, dr = ex(rs[0])
    ti, di = oracle.bf_match(dl, dr)
    assert (int(nl), int(nr), int(nm), int(levels)) == (len(kl), len(kr), len(kl), 8)
    assert np.float32(sf7) == ex.tables()[0][7]
    rows = np.array([[float(v) for v in ln.split()] for ln in lines[1:1 + len(kl)]])
    for i, f in enumerate(("x", "y", "size", "angle", "response")):
        assert np.array_equal(rows[:, i].astype(np.float32), kl[f]), f
    assert np.array_equal(rows[:, 5].astype(np.int32), kl["octave"])
    assert np.array_equal(rows[:, 6:38].astype(np.uint8), dl)
    assert np.array_equal(rows[:, 38].astype(np.int32), ti) and np.array_equal(rows[:, 39].astype(np.int32), di)
    # b2orb::triangulatePoints on every match == the oracle (itself pinned to cv2.triangulatePoints)
    fx, cx, cy, bfx = [float(v) for v in next(ln for ln in lines if ln.startswith("P ")).split()[1:]]
    P1 = np.array([[fx, 0, cx, 0], [0, fx, cy, 0], [0, 0, 1, 0.0]]); P2 = P1.copy(); P2[0, 3] = bfx
    got = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("T ")]).T
    pl = np.stack([kl["x"], kl["y"]], 1).astype(np.float64); pr = np.stack([kr["x"][ti], kr["y"][ti]], 1).astype(np.float64)
    ref = oracle.triangulate(P1, P2, pl, pr)
    assert got.shape == ref.shape and np.array_equal(got.view(np.uint64), ref.view(np.uint64))

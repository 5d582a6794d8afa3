#!/usr/bin/env python3
"""Generate (1) tests/golden/real/*.npz — real camera frames shipped with the reference, decoded to raw u8 so the
GPU box needs neither the reference tree nor an image codec (INPUT fixtures), and (2) tests/golden/ref/*.npz —
known-answer OUTPUTS of the reference's own extractor: /root/reference/src/ORBextractor.cc compiled unmodified into
oracle/_ref/libref_orb.so (oracle/ref/Makefile; ascending-address arena = the documented tie rule) and run here on
the synthetic shapes and on the real frames.  Keypoints are stored whole (28-byte cv::KeyPoint records), descriptors
whole, so every consumer (oracle on CPU, CUDA path on the GPU box) can be compared bit for bit with what the
reference's code produced.

Run here (needs /root/reference and cv2 for decoding):  python tools/gen_ref_golden.py
"""
import hashlib
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as orc  # noqa: E402

synth = importlib.import_module("pose-estimation_b200.synth")
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")

# name: (path relative to /root/reference, nfeatures, iniThFAST, minThFAST)  — thresholds of the yml that matches the camera
REAL = {
    "cfsd360p_left": ("benchmark/cvRectify/cfsd360pLeft.jpg", 1000, 20, 7),     # config/cfsd.yml:101,118-119
    "cfsd360p_right": ("benchmark/cvRectify/cfsd360pRight.jpg", 1000, 20, 7),
    "kitti0": ("benchmark/cvRectify/kitti0.png", 2000, 12, 7),                  # config/kitti.yml:90,107-108
    "kitti2": ("benchmark/cvRectify/kitti2.png", 2000, 12, 7),
    "cfsd_set_110_left": ("tools/opencv-samples/cfsd-imageset/110_left.jpg", 1000, 20, 7),
    "cfsd_set_110_right": ("tools/opencv-samples/cfsd-imageset/110_right.jpg", 1000, 20, 7),
}
# name: (w, h, nfeatures, iniThFAST, minThFAST, seed, passes, frame) — same cases as tools/gen_golden.py + the yml shapes
SYNTH = {
    "small_320x240": (320, 240, 300, 20, 7, 0, 3, 0),
    "euroc_752x480": (752, 480, 1000, 20, 7, 0, 3, 0),
    "euroc_yml_800": (752, 480, 800, 20, 7, 1, 3, 2),        # config/euroc.yml:89 numberOfFeatures 800
    "kitti_1241x376": (1241, 376, 2000, 12, 7, 0, 3, 1),
    "kitti_yml_1392x512": (1392, 512, 2000, 12, 7, 3, 3, 0),  # config/kitti.yml:12-13
    "cfsd_672x376": (672, 376, 1000, 20, 7, 2, 3, 3),
    "lowtex_400x300": (400, 300, 500, 20, 7, 4, 8, 0),
}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def run_ref(img, nf, ini, mn, taps=None):
    ex = orc.RefExtractor(nf, 1.2, 8, ini, mn, taps=taps, arena=1)
    kp, desc = ex(img)
    levels = [sha(ex.level(l)) for l in range(8)]
    return kp, desc, levels


def main():
    assert orc.build_ref(), "oracle/_ref not built"
    import cv2
    os.makedirs(os.path.join(GOLD, "real"), exist_ok=True)
    os.makedirs(os.path.join(GOLD, "ref"), exist_ok=True)
    for name, (rel, nf, ini, mn) in REAL.items():
        img = cv2.imread(os.path.join(REF, rel), cv2.IMREAD_GRAYSCALE)
        assert img is not None and img.dtype == np.uint8, rel
        np.savez_compressed(os.path.join(GOLD, "real", name + ".npz"), image=img, source=rel)
        kp, desc, levels = run_ref(img, nf, ini, mn)
        np.savez_compressed(os.path.join(GOLD, "ref", "real_" + name + ".npz"), cfg=np.array([nf, ini, mn]), image_sha=sha(img),
                            kp=kp, desc=desc, level_sha=np.array(levels))
        print(name, img.shape, len(kp), "keypoints")
    for name, (w, h, nf, ini, mn, seed, passes, frame) in SYNTH.items():
        ls, rs = synth.sequence((w, h), frame + 1, seed=seed, passes=passes)
        out = {"case": np.array([w, h, nf, ini, mn, seed, passes, frame])}
        for eye, img in (("l", ls[frame]), ("r", rs[frame])):
            kp, desc, levels = run_ref(img, nf, ini, mn)
            out.update({"image_sha_" + eye: sha(img), "kp_" + eye: kp, "desc_" + eye: desc, "level_sha_" + eye: np.array(levels)})
        # OpenCV-3.2 Gaussian taps (the version the reference's Dockerfile names), left eye only
        kp32, desc32, _ = run_ref(ls[frame], nf, ini, mn, taps=orc.TAPS_CV32)
        out.update({"kp_l_cv32": kp32, "desc_l_cv32": desc32})
        np.savez_compressed(os.path.join(GOLD, "ref", name + ".npz"), **out)
        print(name, len(out["kp_l"]), len(out["kp_r"]), "keypoints; cv32 taps change",
              int((desc32 != out["desc_l"]).any(1).sum()), "descriptors")


if __name__ == "__main__":
    main()

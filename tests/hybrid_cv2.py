"""Second, independent restatement of ORBextractor::operator() built from REAL cv2 primitives
(resize / FastFeatureDetector on cell ROIs / GaussianBlur / fastAtan2) plus a plain-Python quadtree and
descriptor (libm cosf/sinf through ctypes).  Used only to pin the C++ oracle end to end; it follows
/root/reference/src/ORBextractor.cc:765-853 (cell loop), :539-763 (quadtree), :77-104, :108-147, :1043-1105.
"""
import ctypes
import math
import os

import numpy as np

_libm = ctypes.CDLL("libm.so.6")
_libm.cosf.restype = ctypes.c_float; _libm.cosf.argtypes = [ctypes.c_float]
_libm.sinf.restype = ctypes.c_float; _libm.sinf.argtypes = [ctypes.c_float]

F = np.float32


def load_pattern():
    here = os.path.dirname(os.path.abspath(__file__))
    txt = open(os.path.join(here, "..", "oracle", "orb_pattern.inc")).read()
    txt = txt[txt.index("*/") + 2:]
    v = [int(t) for t in txt.replace("\n", "").split(",") if t.strip()]
    assert len(v) == 1024
    return np.array(v, np.int32).reshape(512, 2)


def tables(nfeatures, scale_factor, nlevels):
    sf = [F(1.0)]
    d = float(F(scale_factor))
    for _ in range(1, nlevels):
        sf.append(F(float(sf[-1]) * d))
    inv = [F(1.0) / s for s in sf]
    factor = F(1.0 / d)
    nd = F(nfeatures * (1 - float(factor)) / (1 - float(F(math.pow(float(factor), nlevels)))))
    # the reference computes nfeatures*(1 - factor) in float (int*float) then divides in float
    nd = F(F(nfeatures) * (F(1) - factor)) / (F(1) - F(math.pow(float(factor), float(nlevels))))
    quota, tot = [], 0
    for _ in range(nlevels - 1):
        q = int(np.rint(nd)); quota.append(q); tot += q; nd = F(nd * factor)
    quota.append(max(nfeatures - tot, 0))
    umax = [0] * 16
    vmax = int(math.floor(15 * math.sqrt(2.0) / 2 + 1)); vmin = int(math.ceil(15 * math.sqrt(2.0) / 2))
    for v in range(vmax + 1):
        umax[v] = int(np.rint(math.sqrt(225.0 - v * v)))
    v0 = 0
    for v in range(15, vmin - 1, -1):
        while umax[v0] == umax[v0 + 1]:
            v0 += 1
        umax[v] = v0; v0 += 1
    return sf, inv, quota, umax


class _Node:
    __slots__ = ("ulx", "uly", "brx", "bry", "keys", "no_more", "seq")


def _divide(p):
    hx = int(math.ceil(float(F(p.brx - p.ulx) / F(2)))); hy = int(math.ceil(float(F(p.bry - p.uly) / F(2))))
    kids = []
    for (ax, ay, bx, by) in ((p.ulx, p.uly, p.ulx + hx, p.uly + hy), (p.ulx + hx, p.uly, p.brx, p.uly + hy),
                             (p.ulx, p.uly + hy, p.ulx + hx, p.bry), (p.ulx + hx, p.uly + hy, p.brx, p.bry)):
        n = _Node(); n.ulx, n.uly, n.brx, n.bry, n.keys, n.no_more, n.seq = ax, ay, bx, by, [], False, -1
        kids.append(n)
    mx, my = p.ulx + hx, p.uly + hy
    for k in p.keys:
        kids[(0 if k[0] < mx else 1) + (0 if k[1] < my else 2)].keys.append(k)
    for n in kids:
        if len(n.keys) == 1:
            n.no_more = True
    return kids


def octtree(keys, min_x, max_x, min_y, max_y, n_want):
    """keys: list of (x, y, response) -> kept keys in list order.  Tie rule: later-created first."""
    if not keys:
        return []
    n_ini = int(round(float(F(max_x - min_x) / F(max_y - min_y))))
    hx = F(max_x - min_x) / F(n_ini)
    seq = 0
    nodes = []  # python list used as the std::list (front = index 0)
    for i in range(n_ini):
        n = _Node(); n.ulx = int(hx * F(i)); n.brx = int(hx * F(i + 1)); n.uly = 0; n.bry = max_y - min_y
        n.keys = []; n.no_more = False; n.seq = seq; seq += 1
        nodes.append(n)
    roots = list(nodes)
    for k in keys:
        roots[int(F(k[0]) / hx)].keys.append(k)
    nodes = [n for n in nodes if n.keys]
    for n in nodes:
        if len(n.keys) == 1:
            n.no_more = True
    finish = False
    while not finish:
        prev = len(nodes)
        cands = []
        new_front, keep = [], []
        for n in nodes:  # children go to the front in creation order -> reversed at the end
            if n.no_more:
                keep.append(n); continue
            for c in _divide(n):
                if c.keys:
                    c.seq = seq; seq += 1
                    new_front.append(c)
                    if len(c.keys) > 1:
                        cands.append(c)
        nodes = new_front[::-1] + keep
        if len(nodes) >= n_want or len(nodes) == prev:
            finish = True
        elif len(nodes) + 3 * len(cands) > n_want:
            while not finish:
                prev = len(nodes)
                order = sorted(cands, key=lambda c: (len(c.keys), c.seq))
                cands = []
                for c in reversed(order):
                    front = []
                    for k in _divide(c):
                        if k.keys:
                            k.seq = seq; seq += 1
                            front.append(k)
                            if len(k.keys) > 1:
                                cands.append(k)
                    nodes.remove(c)
                    nodes = front[::-1] + nodes
                    if len(nodes) >= n_want:
                        break
                if len(nodes) >= n_want or len(nodes) == prev:
                    finish = True
    out = []
    for n in nodes:
        best = n.keys[0]
        for k in n.keys[1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return out


def extract(cv2, image, nfeatures, scale_factor, nlevels, ini_th, min_th):
    sf, inv, quota, umax = tables(nfeatures, scale_factor, nlevels)
    pattern = load_pattern()
    h0, w0 = image.shape
    pyr = [image]
    for lvl in range(1, nlevels):
        dw = int(np.rint(F(w0) * inv[lvl])); dh = int(np.rint(F(h0) * inv[lvl]))
        pyr.append(cv2.resize(pyr[-1], (dw, dh), interpolation=cv2.INTER_LINEAR))
    det_ini = cv2.FastFeatureDetector_create(ini_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    det_min = cv2.FastFeatureDetector_create(min_th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
    raw_all, kps_all, desc_all, blur_all = [], [], [], []
    for lvl in range(nlevels):
        im = pyr[lvl]
        min_b = 16; max_bx = im.shape[1] - 16; max_by = im.shape[0] - 16
        width = F(max_bx - min_b); height = F(max_by - min_b)
        n_cols = int(width / F(30)); n_rows = int(height / F(30))
        w_cell = int(math.ceil(float(width / F(n_cols)))); h_cell = int(math.ceil(float(height / F(n_rows))))
        raw = []
        for i in range(n_rows):
            ini_y = min_b + i * h_cell; max_y = ini_y + h_cell + 6
            if ini_y >= max_by - 3:
                continue
            max_y = min(max_y, max_by)
            for j in range(n_cols):
                ini_x = min_b + j * w_cell; max_x = ini_x + w_cell + 6
                if ini_x >= max_bx - 6:
                    continue
                max_x = min(max_x, max_bx)
                roi = np.ascontiguousarray(im[ini_y:max_y, ini_x:max_x])
                ks = det_ini.detect(roi)
                if not ks:
                    ks = det_min.detect(roi)
                for k in ks:
                    raw.append((k.pt[0] + j * w_cell, k.pt[1] + i * h_cell, k.response))
        raw_all.append(raw)
        sel = octtree(raw, min_b, max_bx, min_b, max_by, quota[lvl])
        kps = []
        for (x, y, r) in sel:
            x += min_b; y += min_b
            m01 = m10 = 0
            cx, cy = int(x), int(y)
            for u in range(-15, 16):
                m10 += u * int(im[cy, cx + u])
            for v in range(1, 16):
                d = umax[v]; vs = 0
                for u in range(-d, d + 1):
                    p, m = int(im[cy + v, cx + u]), int(im[cy - v, cx + u])
                    vs += p - m; m10 += u * (p + m)
                m01 += v * vs
            ang = cv2.fastAtan2(float(m01), float(m10))
            kps.append([F(x), F(y), F(int(F(31) * sf[lvl])), F(ang), F(r), lvl])
        blur = cv2.GaussianBlur(im, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101) if kps else None
        blur_all.append(blur)
        for kp in kps:
            ang = F(kp[3]) * F(math.pi / float(F(180.0)))
            a = F(_libm.cosf(float(ang))); b = F(_libm.sinf(float(ang)))
            cx, cy = int(kp[0]), int(kp[1])
            px = pattern[:, 0].astype(np.float32); py = pattern[:, 1].astype(np.float32)
            dy = np.rint((px * b).astype(np.float32) + (py * a).astype(np.float32)).astype(np.int32)
            dx = np.rint((px * a).astype(np.float32) - (py * b).astype(np.float32)).astype(np.int32)
            vals = blur[cy + dy, cx + dx].astype(np.int32)
            bits = (vals[0::2] < vals[1::2]).astype(np.uint8).reshape(32, 8)
            desc_all.append((bits << np.arange(8, dtype=np.uint8)).sum(axis=1).astype(np.uint8))
            if lvl:
                kp[0] = F(kp[0] * sf[lvl]); kp[1] = F(kp[1] * sf[lvl])
        kps_all.extend(kps)
    return pyr, raw_all, blur_all, kps_all, (np.array(desc_all, np.uint8).reshape(-1, 32))

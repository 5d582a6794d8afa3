"""ctypes binding of the CPU oracle (oracle/orb_oracle.cpp).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liborb_oracle.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

TAPS_CV4 = (18, 34, 48, 56, 48, 34, 18)   # OpenCV 4.x fixed-point 7x7 sigma=2 kernel
TAPS_CV32 = (18, 34, 49, 55, 49, 34, 18)  # what OpenCV 3.2 would round the float kernel to


def build(force=False):
    """Compile liborb_oracle.so with oracle/Makefile (g++ only)."""
    src = os.path.join(_HERE, "orb_oracle.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u8p, i32p, f32p, vp = C.POINTER(C.c_uint8), C.POINTER(C.c_int), C.POINTER(C.c_float), C.c_void_p
        L.orc_create.restype = vp
        L.orc_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orc_destroy.argtypes = [vp]
        L.orc_set_taps.argtypes = [vp, i32p]
        L.orc_set_tie_mode.argtypes = [vp, C.c_int]
        L.orc_get_ties.argtypes = [vp, i32p]
        L.orc_tables.argtypes = [vp, f32p, f32p, i32p, i32p]
        L.orc_extract.restype = C.c_int
        L.orc_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.orc_level_size.argtypes = [vp, C.c_int, i32p, i32p]
        L.orc_get_level.restype = C.c_int
        L.orc_get_level.argtypes = [vp, C.c_int, C.c_int, vp]
        L.orc_get_raw.restype = C.c_int
        L.orc_get_raw.argtypes = [vp, C.c_int, vp, C.c_int]
        L.orc_resize.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, C.c_int]
        L.orc_blur.argtypes = [vp, C.c_int, C.c_int, vp, i32p]
        L.orc_cfsd_ingest.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.orc_remap.argtypes = [vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, vp, C.c_int]
        L.orc_fast.restype = C.c_int
        L.orc_fast.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_fast_atan2.restype = C.c_float
        L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orc_cosf.restype = C.c_float
        L.orc_cosf.argtypes = [C.c_float]
        L.orc_sinf.restype = C.c_float
        L.orc_sinf.argtypes = [C.c_float]
        L.orc_sincos_check.restype = C.c_long
        L.orc_sincos_check.argtypes = [C.c_uint32, C.c_uint32]
        L.orc_octtree.restype = C.c_int
        L.orc_octtree.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int]
        L.orc_bf_match.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp]
        L.orc_stereo_filter.restype = C.c_int
        L.orc_stereo_filter.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, C.c_float, C.c_float, C.c_float, vp, vp]
        L.orc_threshold_filter.restype = C.c_int
        L.orc_threshold_filter.argtypes = [vp, vp, C.c_int, C.c_int, C.c_float, C.c_float, vp, vp]
        L.orc_triangulate.argtypes = [vp, vp, vp, vp, C.c_int, vp]
        L.orc_depth_gate.argtypes = [vp, C.c_int, C.c_double, vp, vp]
        L.orc_pool_create.restype = vp
        L.orc_pool_create.argtypes = [C.c_float, C.c_float, C.c_int]
        L.orc_pool_destroy.argtypes = [vp]
        L.orc_pool_track.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp, vp, vp]
        L.orc_pool_update.restype = C.c_int
        L.orc_pool_update.argtypes = [vp, C.c_int, vp, i32p]
        L.orc_pool_size.restype = C.c_int
        L.orc_pool_size.argtypes = [vp]
        L.orc_pool_get.argtypes = [vp, vp, vp, vp, vp]
        L.orc_stereo_pair.restype = C.c_int
        L.orc_stereo_pair.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float,
                                      C.c_int, vp, vp, vp, vp, C.c_int, i32p, i32p, vp, vp, vp, vp]
        L.orc_last_stage_seconds.argtypes = [vp]
        L.orc_run_pairs.restype = C.c_long
        L.orc_run_pairs.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int,
                                    C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(a):
    if not (isinstance(a, np.ndarray) and a.dtype == np.uint8 and a.ndim == 2 and a.strides[1] == 1 and a.strides[0] >= a.shape[1]):
        a = np.ascontiguousarray(a, dtype=np.uint8)   # row-strided views (cv::Mat ROIs) are passed through as they are
    assert a.ndim == 2
    return a


class OracleExtractor:
    """Mirror of ORB_SLAM2::ORBextractor (include/ORBextractor.h:45-111) on the CPU oracle."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast, taps=None, tie_mode=1):
        self.L = lib()
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.h = self.L.orc_create(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast)
        if taps is not None:
            t = (C.c_int * 7)(*taps)
            self.L.orc_set_taps(self.h, t)
        if tie_mode != 1:
            self.L.orc_set_tie_mode(self.h, tie_mode)

    def ties(self):
        """Per level of the last call: how many equal-count comparisons the quadtree's final partial pass consumed
        (0 = the reference's pointer tie-break at src/ORBextractor.cc:684 could not matter on that level)."""
        t = (C.c_int * self.nlevels)()
        self.L.orc_get_ties(self.h, t)
        return np.array(t[:], np.int32)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_destroy(self.h)
            self.h = None

    def tables(self):
        nl = self.nlevels
        s = (C.c_float * nl)(); i = (C.c_float * nl)(); q = (C.c_int * nl)(); u = (C.c_int * 16)()
        self.L.orc_tables(self.h, s, i, q, u)
        return (np.array(s[:], np.float32), np.array(i[:], np.float32), np.array(q[:], np.int32),
                np.array(u[:], np.int32))

    def __call__(self, image):
        image = _img(image)
        cap = self.nfeatures + 4 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self.L.orc_extract(self.h, _p(image), image.shape[1], image.shape[0], image.strides[0],
                               _p(kps), _p(desc), cap)
        assert n >= 0, "oracle capacity exceeded"
        return kps[:n].copy(), desc[:n].copy()

    def level(self, level, blurred=False):
        w, h = C.c_int(), C.c_int()
        self.L.orc_level_size(self.h, level, C.byref(w), C.byref(h))
        out = np.zeros((h.value, w.value), np.uint8)
        ok = self.L.orc_get_level(self.h, level, 1 if blurred else 0, _p(out))
        return out if ok else None

    def raw_corners(self, level):
        cap = 1 << 21
        out = np.zeros((cap, 3), np.float32)
        n = self.L.orc_get_raw(self.h, level, _p(out), cap)
        return out[:n].copy()


# ------------------------------------------------------------------------------------------------------------------
# oracle/_ref: the reference's own src/ORBextractor.cc, compiled unmodified against tests/mock_opencv (oracle/ref/)
# ------------------------------------------------------------------------------------------------------------------
REF_SRC = "/root/reference/src/ORBextractor.cc"
_REF_DIR = os.path.join(_HERE, "_ref")
_REF_LIBS = {1: "libref_orb.so", -1: "libref_orb_desc.so", 0: "libref_orb_sysmalloc.so"}
_ref_libs = {}


def build_ref():
    """(Re)build oracle/_ref/*.so when the reference tree is mounted (this container); on the GPU box the prebuilt
    files travel with the snapshot.  Returns True when the libraries exist afterwards."""
    lib()
    if os.path.exists(REF_SRC):
        subprocess.check_call(["make", "-C", os.path.join(_HERE, "ref"), "-s"])
    return all(os.path.exists(os.path.join(_REF_DIR, n)) for n in _REF_LIBS.values())


def ref_lib(arena=1):
    """arena: +1 ascending heap addresses (the documented tie rule), -1 descending, 0 the process malloc."""
    if arena not in _ref_libs:
        if not build_ref():
            return None
        L = C.CDLL(os.path.join(_REF_DIR, _REF_LIBS[arena]))
        vp, i32p, f32p = C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_float)
        L.ref_create.restype = vp
        L.ref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_destroy.argtypes = [vp]
        L.ref_set_taps.argtypes = [i32p]
        L.ref_tables.argtypes = [vp, f32p, f32p, f32p, f32p, i32p, i32p]
        L.ref_extract.restype = C.c_int
        L.ref_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int]
        L.ref_level_size.argtypes = [vp, C.c_int, i32p, i32p]
        L.ref_get_level.argtypes = [vp, C.c_int, vp]
        L.ref_get_level_bordered.argtypes = [vp, C.c_int, vp]
        L.ref_run_pairs.restype = C.c_long
        L.ref_run_pairs.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, C.c_int, C.c_int,
                                    C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_int]
        assert L.ref_arena_mode() == arena
        _ref_libs[arena] = L
    return _ref_libs[arena]


def ref_available():
    return ref_lib(1) is not None


class RefExtractor:
    """ORB_SLAM2::ORBextractor of the reference itself (oracle/_ref), same call surface as OracleExtractor."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast, taps=None, arena=1):
        self.L = ref_lib(arena)
        if self.L is None:
            raise RuntimeError("oracle/_ref is not built and /root/reference is not mounted")
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self.taps = tuple(taps) if taps is not None else TAPS_CV4
        self.h = self.L.ref_create(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_destroy(self.h)
            self.h = None

    def tables(self):
        """(scale, inv_scale, sigma2, inv_sigma2, quota, umax) — src/ORBextractor.cc:415-469."""
        nl = self.nlevels
        f = [(C.c_float * nl)() for _ in range(4)]
        q = (C.c_int * nl)(); u = (C.c_int * 16)()
        self.L.ref_tables(self.h, f[0], f[1], f[2], f[3], q, u)
        return tuple(np.array(a[:], np.float32) for a in f) + (np.array(q[:], np.int32), np.array(u[:], np.int32))

    def __call__(self, image):
        image = _img(image)
        cap = self.nfeatures + 4 * self.nlevels + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        self.L.ref_set_taps((C.c_int * 7)(*self.taps))
        n = self.L.ref_extract(self.h, _p(image), image.shape[1], image.shape[0], image.strides[0],
                               _p(kps), _p(desc), cap)
        assert n >= 0, "capacity exceeded"
        return kps[:n].copy(), desc[:n].copy()

    def level(self, level, bordered=False):
        w, h = C.c_int(), C.c_int()
        self.L.ref_level_size(self.h, level, C.byref(w), C.byref(h))
        if bordered:
            out = np.zeros((h.value + 38, w.value + 38), np.uint8)
            self.L.ref_get_level_bordered(self.h, level, _p(out))
        else:
            out = np.zeros((h.value, w.value), np.uint8)
            self.L.ref_get_level(self.h, level, _p(out))
        return out


def ref_run_pairs(cfg, imgs_l, imgs_r, ratio, min_match_dist, max_dy, workers, per_pair_threads, temporal=False):
    """run_pairs with the reference's own ORBextractor doing the extraction (oracle/_ref, process-malloc build: the
    arena builds exist for reproducible tie-breaks, not for speed)."""
    L = ref_lib(0)
    imgs_l = np.ascontiguousarray(imgs_l, np.uint8); imgs_r = np.ascontiguousarray(imgs_r, np.uint8)
    p, h, w = imgs_l.shape
    return L.ref_run_pairs(cfg[0], cfg[1], cfg[2], cfg[3], cfg[4], _p(imgs_l), _p(imgs_r), p, w, h, ratio, min_match_dist,
                           max_dy, workers, per_pair_threads, 1 if temporal else 0)


def resize(src, dw, dh):
    src = _img(src)
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_resize(_p(src), src.shape[1], src.shape[0], _p(dst), dw, dh)
    return dst


def cfsd_ingest(frame, dw, dh):
    """src/cfsd-state-estimation.cpp:117-127: (H, W, 4) u8 side-by-side frame -> resize to dw x dh, BGR2GRAY, split.
    Returns (grayL, grayR), each dh x dw/2."""
    frame = np.ascontiguousarray(frame, np.uint8)
    sh, sw, _ = frame.shape
    gl = np.zeros((dh, dw // 2), np.uint8); gr = np.zeros((dh, dw // 2), np.uint8)
    lib().orc_cfsd_ingest(frame.ctypes.data, sw, sh, sw * 4, dw, dh, _p(gl), _p(gr), dw // 2)
    return gl, gr


def remap(src, map1, map2):
    """cv::remap(src, map1 (H x W x 2 int16), map2 (H x W uint16), INTER_LINEAR), src/feature-tracker.cpp:75-76."""
    src = _img(src)
    map1 = np.ascontiguousarray(map1, np.int16); map2 = np.ascontiguousarray(map2, np.uint16)
    dh, dw = map2.shape
    dst = np.zeros((dh, dw), np.uint8)
    lib().orc_remap(_p(src), src.shape[1], src.shape[0], src.shape[1], map1.ctypes.data, map2.ctypes.data, dw, dh, _p(dst), dw)
    return dst


def blur(src, taps=None):
    src = _img(src)
    dst = np.zeros_like(src)
    t = (C.c_int * 7)(*taps) if taps is not None else None
    lib().orc_blur(_p(src), src.shape[1], src.shape[0], _p(dst), t)
    return dst


def fast(img, roi, thr):
    """FAST-9/16 + NMS on roi=(x0,y0,x1,y1) -> (n,3) float32 [x, y, response] relative to the ROI."""
    img = _img(img)
    cap = img.size // 2 + 16
    out = np.zeros((cap, 3), np.float32)
    n = lib().orc_fast(_p(img), img.shape[1], img.shape[0], roi[0], roi[1], roi[2], roi[3], thr, _p(out), cap)
    return out[:n].copy()


def fast_atan2(y, x):
    return lib().orc_fast_atan2(float(y), float(x))


def octtree(xyr, min_x, max_x, min_y, max_y, n_want):
    xyr = np.ascontiguousarray(xyr, np.float32)
    cap = n_want * 4 + 16
    out = np.zeros((cap, 3), np.float32)
    m = lib().orc_octtree(_p(xyr), len(xyr), min_x, max_x, min_y, max_y, n_want, _p(out), cap)
    return out[:m].copy()


def bf_match(q, t):
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
    t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    ti = np.full(len(q), -1, np.int32)
    di = np.full(len(q), 0x7fffffff, np.int32)
    if len(q):
        lib().orc_bf_match(_p(q), len(q), _p(t), len(t), _p(ti), _p(di))
    return ti, di


def stereo_filter(train_idx, dist, kp_l, kp_r, ratio, min_match_dist, max_dy):
    nq, nt = len(kp_l), len(kp_r)
    o_l = np.zeros(max(nq, 1), np.int32); o_r = np.zeros(max(nq, 1), np.int32)
    kp_l = np.ascontiguousarray(kp_l); kp_r = np.ascontiguousarray(kp_r)
    ti = np.ascontiguousarray(train_idx, np.int32); di = np.ascontiguousarray(dist, np.int32)
    m = lib().orc_stereo_filter(_p(ti), _p(di), nq, nt, _p(kp_l), _p(kp_r), ratio, min_match_dist, max_dy,
                                _p(o_l), _p(o_r))
    return o_l[:m].copy(), o_r[:m].copy()


def threshold_filter(train_idx, dist, nt, ratio, min_match_dist):
    nq = len(train_idx)
    o_q = np.zeros(max(nq, 1), np.int32); o_t = np.zeros(max(nq, 1), np.int32)
    ti = np.ascontiguousarray(train_idx, np.int32); di = np.ascontiguousarray(dist, np.int32)
    m = lib().orc_threshold_filter(_p(ti), _p(di), nq, nt, ratio, min_match_dist, _p(o_q), _p(o_t))
    return o_q[:m].copy(), o_t[:m].copy()


class OracleFeaturePool:
    """The feature pool of cfsd::FeatureTracker (src/feature-tracker.cpp:231-328 matching parts with useRANSAC=false,
    :330-412 pool bookkeeping), restated on the same std::unordered_map so the pool ORDER follows libstdc++'s."""

    def __init__(self, ratio=2.0, min_match_dist=30.0, max_feature_age=8):
        self.L = lib()
        self.h = self.L.orc_pool_create(ratio, min_match_dist, max_feature_age)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_pool_destroy(self.h)
            self.h = None

    def track(self, cur_l, cur_r):
        """externalTrack: returns dict(circ_cur, circ_hist, mask, matched_ids, n_left, n_right, n_hist)."""
        cl = np.ascontiguousarray(cur_l, np.uint8).reshape(-1, 32); cr = np.ascontiguousarray(cur_r, np.uint8).reshape(-1, 32)
        n = len(cl)
        cc = np.zeros(max(n, 1), np.int32); ch = np.zeros(max(n, 1), np.int32)
        mask = np.zeros(max(n, 1), np.uint8); ids = np.zeros(max(n, 1), np.uint64); cnt = np.zeros(4, np.int32)
        self.L.orc_pool_track(self.h, _p(cl), _p(cr), n, _p(cc), _p(ch), _p(mask), _p(ids), _p(cnt))
        k = int(cnt[0])
        return {"circ_cur": cc[:k].copy(), "circ_hist": ch[:k].copy(), "mask": mask[:n].copy(), "matched_ids": ids[:k].copy(),
                "n_left": int(cnt[1]), "n_right": int(cnt[2]), "n_hist": int(cnt[3])}

    def update(self, is_keyframe, accept=None):
        """featurePoolUpdate: returns (inserted, erased)."""
        er = C.c_int(0)
        acc = np.ascontiguousarray(accept, np.uint8) if accept is not None else None
        ins = self.L.orc_pool_update(self.h, 1 if is_keyframe else 0, _p(acc) if acc is not None else None, C.byref(er))
        return ins, er.value

    def state(self):
        """(_histFeatureIDs, ages, _histDescriptorsL, _histDescriptorsR) in pool order."""
        n = self.L.orc_pool_size(self.h)
        ids = np.zeros(max(n, 1), np.uint64); ages = np.zeros(max(n, 1), np.int32)
        hl = np.zeros((max(n, 1), 32), np.uint8); hr = np.zeros((max(n, 1), 32), np.uint8)
        self.L.orc_pool_get(self.h, _p(ids), _p(ages), _p(hl), _p(hr))
        return ids[:n].copy(), ages[:n].copy(), hl[:n].copy(), hr[:n].copy()


def triangulate(P1, P2, pix_l, pix_r):
    """cv::triangulatePoints(P1, P2, pixelsL, pixelsR, points4D) with vector<Point2d> inputs
    (src/feature-tracker.cpp:367, :489): pix_* are (n, 2) float64; returns points4D as a (4, n) float64 array."""
    P1 = np.ascontiguousarray(P1, np.float64).reshape(3, 4)
    P2 = np.ascontiguousarray(P2, np.float64).reshape(3, 4)
    pl = np.ascontiguousarray(pix_l, np.float64).reshape(-1, 2)
    pr = np.ascontiguousarray(pix_r, np.float64).reshape(-1, 2)
    assert len(pl) == len(pr)
    out = np.zeros((4, len(pl)), np.float64)
    if len(pl):
        lib().orc_triangulate(_p(P1), _p(P2), _p(pl), _p(pr), len(pl), _p(out))
    return out


def depth_gate(points4d, max_depth):
    """point_wrt_cam = (X/W, Y/W, Z/W) and the two depth gates (src/feature-tracker.cpp:371-378, :501-506):
    flags bit 0 = !(depth > maxDepth), bit 1 = !(depth < 0.1 || depth > maxDepth)."""
    p4 = np.ascontiguousarray(points4d, np.float64)
    n = p4.shape[1]
    pc = np.zeros((n, 3), np.float64)
    fl = np.zeros(n, np.uint8)
    if n:
        lib().orc_depth_gate(_p(p4), n, float(max_depth), _p(pc), _p(fl))
    return pc, fl


STAGES = ("pyramid", "fast_cells", "quadtree", "orientation", "blur", "describe", "match_stereo", "match_temporal")


def last_stage_seconds():
    """Per-stage CPU seconds of the last run_pairs, summed over workers (thread-seconds)."""
    out = np.zeros(8, np.float64)
    lib().orc_last_stage_seconds(_p(out))
    return dict(zip(STAGES, out.tolist()))


def run_pairs(cfg, imgs_l, imgs_r, ratio, min_match_dist, max_dy, workers, per_pair_threads, temporal=False):
    """CPU baseline driver: cfg = (nfeatures, scaleFactor, nlevels, iniTh, minTh); imgs_* = (P,H,W) u8."""
    imgs_l = np.ascontiguousarray(imgs_l, np.uint8); imgs_r = np.ascontiguousarray(imgs_r, np.uint8)
    p, h, w = imgs_l.shape
    return lib().orc_run_pairs(cfg[0], cfg[1], cfg[2], cfg[3], cfg[4], _p(imgs_l), _p(imgs_r), p, w, h,
                               ratio, min_match_dist, max_dy, workers, per_pair_threads, 1 if temporal else 0)

"""Host-side mirror of the reference's call surface over the C ABI (include/b2orb.h), via ctypes.

  * ORBextractor      <-> ORB_SLAM2::ORBextractor      (include/ORBextractor.h:45-111)
  * FeatureTracker    <-> cfsd::FeatureTracker's matching half (src/feature-tracker.cpp:133-328):
                          extractORB, internalMatch (useRANSAC=false), externalTrack matching parts
  * BFMatcher         <-> cv::BFMatcher(cv::NORM_HAMMING).match

There is NO CPU fallback: loading fails loudly when libb2orb.so is missing, and handle creation fails
when no CUDA device is present.  The reference's host language is C++ (the drop-in C++ shim is
include/b2orb_shim/ORBextractor.h); this Python mirror exists so tests and bench.py read like the
reference's call sites.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2ORB_LIB: harness-only override to load the experiment build (csrc `make variants` -> libb2orb_variants.so)
LIB_PATH = os.environ.get("B2ORB_LIB") or os.path.join(_HERE, "libb2orb.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])

OK, ERR_INVALID, ERR_CAPACITY, ERR_CUDA, ERR_NOMEM = 0, 1, 2, 3, 4
KERNEL_NAMES = ("resize", "fast", "octree", "blur", "describe", "match_stereo", "stereo_filter", "match_temporal",
                "circular", "fast_cells", "ingest", "triangulate")


class B2Config(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("ini_th_fast", C.c_int32), ("min_th_fast", C.c_int32), ("max_width", C.c_int32),
                ("max_height", C.c_int32), ("max_images", C.c_int32), ("blur_taps", C.c_int32 * 7),
                ("match_ratio", C.c_float), ("min_match_dist", C.c_float), ("max_vertical_pixel_dist", C.c_float)]


class B2StereoOut(C.Structure):
    _fields_ = [("cap", C.c_int32)] + [(n, C.c_void_p) for n in (
        "kp_l", "kp_r", "desc_l", "desc_r", "n_l", "n_r", "train_idx", "dist", "match_l", "match_r", "n_match",
        "circ_cur", "circ_hist", "n_circ", "points4d", "point_cam", "depth_flags", "pix_l", "pix_r", "cur_desc_l",
        "cur_desc_r", "cur_mask")]


class B2Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"b2orb error {code}: {msg}")
        self.code = code


_lib = None


def lib():
    """Load libb2orb.so (built by __graft_entry__.build() / csrc/Makefile).  Fails loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp, i32, f32, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
        L.b2orb_abi_version.restype = i32
        L.b2orb_create.restype = i32
        L.b2orb_create.argtypes = [C.POINTER(vp), i32, C.POINTER(B2Config)]
        L.b2orb_destroy.argtypes = [vp]
        L.b2orb_last_error.restype = C.c_char_p
        L.b2orb_last_error.argtypes = [vp]
        L.b2orb_keypoint_capacity.restype = i32
        L.b2orb_keypoint_capacity.argtypes = [vp]
        L.b2orb_get_tables.restype = i32
        L.b2orb_get_tables.argtypes = [vp, vp, vp, vp, vp, vp]
        L.b2orb_host_alloc.restype = vp
        L.b2orb_host_alloc.argtypes = [sz]
        L.b2orb_host_free.argtypes = [vp]
        L.b2orb_extract.restype = i32
        L.b2orb_extract.argtypes = [vp, vp, i32, i32, i32, vp, vp, i32, C.POINTER(i32)]
        L.b2orb_extract_batch.restype = i32
        L.b2orb_extract_batch.argtypes = [vp, vp, i32, i32, i32, i32, sz, vp, vp, i32, vp]
        L.b2orb_match.restype = i32
        L.b2orb_match.argtypes = [vp, vp, i32, vp, i32, vp, vp]
        L.b2orb_triangulate.restype = i32
        L.b2orb_triangulate.argtypes = [vp, vp, vp, vp, vp, i32, vp]
        L.b2orb_set_projection.restype = i32
        L.b2orb_set_projection.argtypes = [vp, vp, vp, C.c_double]
        L.b2orb_set_rectification.restype = i32
        L.b2orb_set_rectification.argtypes = [vp, i32, i32, vp, vp, vp, vp]
        L.b2orb_cfsd_ingest.restype = i32
        L.b2orb_cfsd_ingest.argtypes = [vp, vp, i32, i32, i32, i32, i32, vp, vp, i32]
        L.b2orb_stereo_frontend_packed.restype = i32
        L.b2orb_stereo_frontend_packed.argtypes = [vp, vp, i32, i32, i32, i32, sz, i32, i32, i32, C.POINTER(B2StereoOut)]
        L.b2orb_remap.restype = i32
        L.b2orb_remap.argtypes = [vp, i32, vp, i32, i32, i32, vp, i32]
        L.b2orb_stereo_frontend.restype = i32
        L.b2orb_stereo_frontend.argtypes = [vp, vp, vp, i32, i32, i32, i32, sz, i32, C.POINTER(B2StereoOut)]
        L.b2orb_stereo_frontend_device.restype = i32
        L.b2orb_stereo_frontend_device.argtypes = [vp, vp, vp, i32, i32, i32, i32, sz, i32]
        L.b2orb_stereo_download.restype = i32
        L.b2orb_stereo_download.argtypes = [vp, i32, C.POINTER(B2StereoOut)]
        L.b2orb_sync.restype = i32
        L.b2orb_sync.argtypes = [vp]
        L.b2orb_reset_history.argtypes = [vp]
        L.b2orb_temporal_track.restype = i32
        L.b2orb_temporal_track.argtypes = [vp, vp, vp, i32, vp, vp, i32, vp, vp, C.POINTER(i32), C.POINTER(i32),
                                           C.POINTER(i32)]
        L.b2orb_debug_level_size.restype = i32
        L.b2orb_debug_level_size.argtypes = [vp, i32, C.POINTER(i32), C.POINTER(i32)]
        L.b2orb_debug_get_level.restype = i32
        L.b2orb_debug_get_level.argtypes = [vp, i32, i32, i32, vp]
        L.b2orb_debug_get_candidates.restype = i32
        L.b2orb_debug_get_candidates.argtypes = [vp, i32, i32, vp, i32, C.POINTER(i32)]
        u8p = vp
        L.b2orb_pool_reset.restype = i32
        L.b2orb_pool_reset.argtypes = [vp, i32]
        L.b2orb_pool_size.restype = i32
        L.b2orb_pool_size.argtypes = [vp]
        for fn in (L.b2orb_pool_track, L.b2orb_pool_last_track):
            fn.restype = i32
        L.b2orb_pool_track.argtypes = [vp, i32, vp, vp, C.POINTER(i32), u8p, vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
        L.b2orb_pool_last_track.argtypes = [vp, vp, vp, C.POINTER(i32), u8p, vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
        L.b2orb_pool_update.restype = i32
        L.b2orb_pool_update.argtypes = [vp, i32, vp, vp, C.POINTER(i32), C.POINTER(i32)]
        L.b2orb_pool_get.restype = i32
        L.b2orb_pool_get.argtypes = [vp, vp, vp, vp, vp, i32, C.POINTER(i32)]
        L.b2orb_stereo_filter.restype = i32
        L.b2orb_stereo_filter.argtypes = [vp, vp, vp, i32, vp, vp, vp, vp, i32, vp, vp, C.POINTER(i32), vp, vp, vp, vp]
        L.b2orb_extract_async.restype = i32
        L.b2orb_extract_async.argtypes = [vp, vp, i32, i32, i32]
        L.b2orb_extract_wait.restype = i32
        L.b2orb_extract_wait.argtypes = [vp, vp, vp, i32, C.POINTER(i32)]
        L.b2orb_debug_set.restype = i32
        L.b2orb_debug_set.argtypes = [vp, C.c_char_p, i32]
        L.b2orb_debug_octree_fallbacks.restype = C.c_uint64
        L.b2orb_debug_octree_fallbacks.argtypes = [vp]
        L.b2orb_stream.restype = vp
        L.b2orb_stream.argtypes = [vp]
        L.b2orb_profile_enable.argtypes = [vp, i32]
        L.b2orb_profile_read.restype = i32
        L.b2orb_profile_read.argtypes = [vp, vp, vp]
        L.b2orb_launch_count.restype = C.c_uint64
        L.b2orb_launch_count.argtypes = [vp]
        L.b2orb_last_device_ms.restype = f32
        L.b2orb_last_device_ms.argtypes = [vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def pinned_empty(shape, dtype=np.uint8):
    """numpy array backed by cudaHostAlloc'd memory (b2orb_host_alloc); freed with the array."""
    L = lib()
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    ptr = L.b2orb_host_alloc(max(n, 1))
    if not ptr:
        raise MemoryError("b2orb_host_alloc failed")
    buf = (C.c_uint8 * max(n, 1)).from_address(ptr)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    buf._owner = _PinnedOwner(L, ptr)   # freed when the last array viewing `buf` is collected
    return arr


class _PinnedOwner:
    def __init__(self, L, p):
        self.L, self.p = L, p

    def __del__(self):
        try:
            self.L.b2orb_host_free(self.p)
        except Exception:
            pass


class Handle:
    """One b2orb_handle: one device, one stream, workspace for `max_images` images."""

    def __init__(self, nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast, max_width, max_height,
                 max_images=2, device=0, match_ratio=2.0, min_match_dist=30.0, max_vertical_pixel_dist=0.1,
                 blur_taps=None):
        self.L = lib()
        cfg = B2Config(nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast, max_width, max_height,
                       max_images, (C.c_int32 * 7)(*(blur_taps or [0] * 7)), match_ratio, min_match_dist,
                       max_vertical_pixel_dist)
        self.cfg = cfg
        self.h = C.c_void_p()
        rc = self.L.b2orb_create(C.byref(self.h), device, C.byref(cfg))
        if rc != OK:
            msg = self.L.b2orb_last_error(self.h).decode() if self.h else "b2orb_create failed (no CUDA device?)"
            if self.h:
                self.L.b2orb_destroy(self.h)
                self.h = None
            raise B2Error(rc, msg)
        self.nlevels = nlevels

    def close(self):
        if getattr(self, "h", None):
            self.L.b2orb_destroy(self.h)
            self.h = None

    __del__ = close

    def check(self, rc):
        if rc != OK:
            raise B2Error(rc, self.L.b2orb_last_error(self.h).decode())

    @property
    def keypoint_capacity(self):
        return self.L.b2orb_keypoint_capacity(self.h)

    @property
    def launch_count(self):
        return int(self.L.b2orb_launch_count(self.h))

    @property
    def last_device_ms(self):
        return float(self.L.b2orb_last_device_ms(self.h))

    @property
    def stream(self):
        return self.L.b2orb_stream(self.h)

    def profile_enable(self, on=True):
        self.L.b2orb_profile_enable(self.h, 1 if on else 0)

    def profile_read(self):
        ms = np.zeros(len(KERNEL_NAMES), np.float64); cnt = np.zeros(len(KERNEL_NAMES), np.uint64)
        self.check(self.L.b2orb_profile_read(self.h, _p(ms), _p(cnt)))
        return {k: (float(ms[i]), int(cnt[i])) for i, k in enumerate(KERNEL_NAMES)}

    def tables(self):
        nl = self.nlevels
        arrs = [np.zeros(nl, np.float32) for _ in range(4)] + [np.zeros(nl, np.int32)]
        self.check(self.L.b2orb_get_tables(self.h, *[_p(a) for a in arrs]))
        return arrs

    # ---- extraction
    def extract(self, image):
        image = np.ascontiguousarray(image, np.uint8) if image.strides[-1] != 1 else image
        hgt, w = image.shape if image.ndim == 2 else (0, 0)
        cap = self.keypoint_capacity
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        step = image.strides[0] if hgt else 0
        self.check(self.L.b2orb_extract(self.h, _p(image) if hgt else None, w, hgt, step, _p(kps), _p(desc), cap,
                                        C.byref(n)))
        return kps[:n.value].copy(), desc[:n.value].copy()

    def extract_async(self, image):
        """b2orb_extract_async: enqueue and return; extract_wait() collects.  Keeps `image` alive until then."""
        image = np.ascontiguousarray(image, np.uint8) if image.strides[-1] != 1 else image
        hgt, w = image.shape if image.ndim == 2 else (0, 0)
        self._async_img = image
        self.check(self.L.b2orb_extract_async(self.h, _p(image) if hgt else None, w, hgt, image.strides[0] if hgt else 0))

    def extract_wait(self):
        cap = self.keypoint_capacity
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        n = C.c_int(0)
        self.check(self.L.b2orb_extract_wait(self.h, _p(kps), _p(desc), cap, C.byref(n)))
        self._async_img = None
        return kps[:n.value].copy(), desc[:n.value].copy()

    def stereo_filter(self, train_idx, dist, kp_l, desc_l, kp_r, desc_r):
        """Host loops of internalMatch (src/feature-tracker.cpp:164, 208-218): returns dict(match_l, match_r, pix_l,
        pix_r, cur_desc_l, cur_desc_r)."""
        ti = np.ascontiguousarray(train_idx, np.int32); di = np.ascontiguousarray(dist, np.int32)
        kl = np.ascontiguousarray(kp_l); kr = np.ascontiguousarray(kp_r)
        dl = np.ascontiguousarray(desc_l, np.uint8).reshape(-1, 32); dr = np.ascontiguousarray(desc_r, np.uint8).reshape(-1, 32)
        nq, nt = len(kl), len(kr)
        ml = np.zeros(max(nq, 1), np.int32); mr = np.zeros(max(nq, 1), np.int32)
        pl = np.zeros((max(nq, 1), 2), np.float64); pr = np.zeros((max(nq, 1), 2), np.float64)
        cl = np.zeros((max(nq, 1), 32), np.uint8); cr = np.zeros((max(nq, 1), 32), np.uint8)
        m = C.c_int(0)
        self.check(self.L.b2orb_stereo_filter(self.h, _p(ti), _p(di), nq, _p(kl), _p(dl), _p(kr), _p(dr), nt, _p(ml), _p(mr),
                                              C.byref(m), _p(pl), _p(pr), _p(cl), _p(cr)))
        k = m.value
        return {"match_l": ml[:k].copy(), "match_r": mr[:k].copy(), "pix_l": pl[:k].copy(), "pix_r": pr[:k].copy(),
                "cur_desc_l": cl[:k].copy(), "cur_desc_r": cr[:k].copy()}

    def extract_batch(self, images):
        images = np.ascontiguousarray(images, np.uint8)
        cnt, hgt, w = images.shape
        cap = self.keypoint_capacity
        kps = np.zeros((cnt, cap), KP_DTYPE)
        desc = np.zeros((cnt, cap, 32), np.uint8)
        n = np.zeros(cnt, np.int32)
        self.check(self.L.b2orb_extract_batch(self.h, _p(images), cnt, w, hgt, w, w * hgt, _p(kps), _p(desc), cap, _p(n)))
        return [(kps[i, :n[i]].copy(), desc[i, :n[i]].copy()) for i in range(cnt)]

    # ---- matching
    def match(self, query, train):
        q = np.ascontiguousarray(query, np.uint8).reshape(-1, 32)
        t = np.ascontiguousarray(train, np.uint8).reshape(-1, 32)
        ti = np.full(len(q), -1, np.int32)
        di = np.full(len(q), 0x7fffffff, np.int32)
        self.check(self.L.b2orb_match(self.h, _p(q), len(q), _p(t), len(t), _p(ti), _p(di)))
        return ti, di

    def temporal_track(self, cur_l, cur_r, hist_l, hist_r):
        cl = np.ascontiguousarray(cur_l, np.uint8).reshape(-1, 32); cr = np.ascontiguousarray(cur_r, np.uint8).reshape(-1, 32)
        hl = np.ascontiguousarray(hist_l, np.uint8).reshape(-1, 32); hr = np.ascontiguousarray(hist_r, np.uint8).reshape(-1, 32)
        assert len(cl) == len(cr) and len(hl) == len(hr)
        cc = np.zeros(max(len(cl), 1), np.int32); ch = np.zeros(max(len(cl), 1), np.int32)
        n, nl, nr = C.c_int(0), C.c_int(0), C.c_int(0)
        self.check(self.L.b2orb_temporal_track(self.h, _p(cl), _p(cr), len(cl), _p(hl), _p(hr), len(hl), _p(cc), _p(ch),
                                               C.byref(n), C.byref(nl), C.byref(nr)))
        return cc[:n.value].copy(), ch[:n.value].copy(), nl.value, nr.value

    # ---- triangulation (next row f4)
    def triangulate(self, P1, P2, pix_l, pix_r):
        """cv::triangulatePoints(P1, P2, pixelsL, pixelsR, points4D) (src/feature-tracker.cpp:367, :489): pix_* are
        (n, 2) float64 (vector<Point2d>); returns points4D as a (4, n) float64 array."""
        P1 = np.ascontiguousarray(P1, np.float64).reshape(3, 4); P2 = np.ascontiguousarray(P2, np.float64).reshape(3, 4)
        pl = np.ascontiguousarray(pix_l, np.float64).reshape(-1, 2); pr = np.ascontiguousarray(pix_r, np.float64).reshape(-1, 2)
        assert len(pl) == len(pr)
        out = np.zeros((4, len(pl)), np.float64)
        self.check(self.L.b2orb_triangulate(self.h, _p(P1), _p(P2), _p(pl), _p(pr), len(pl), _p(out)))
        return out

    def set_projection(self, P1=None, P2=None, max_depth=20.0):
        """CameraModel::_P1/_P2 + maxDepth: afterwards the stereo front end triangulates the kept matches on the device
        (points4d / point_cam / depth_flags of the output).  No arguments: off."""
        if P1 is None:
            self.check(self.L.b2orb_set_projection(self.h, None, None, 0.0))
            self._proj = False
            return
        P1 = np.ascontiguousarray(P1, np.float64).reshape(3, 4); P2 = np.ascontiguousarray(P2, np.float64).reshape(3, 4)
        self.check(self.L.b2orb_set_projection(self.h, _p(P1), _p(P2), float(max_depth)))
        self._proj = True

    # ---- fused stereo front-end
    def alloc_stereo_out(self, npairs, pinned=False, full=True, kept=False):
        """full: every keypoint / descriptor / match of both eyes; kept: what the reference keeps of a frame
        (_curPixelsL/R, _curDescriptorsL/R, _curFeatureMask — src/feature-tracker.cpp:210-218); kept without full is
        the small result set (a few KB per pair)."""
        cap = self.keypoint_capacity
        mk = (lambda shape, dt: pinned_empty(shape, dt)) if pinned else (lambda shape, dt: np.zeros(shape, dt))
        a = {"n_l": mk((npairs,), np.int32), "n_r": mk((npairs,), np.int32), "n_match": mk((npairs,), np.int32),
             "n_circ": mk((npairs,), np.int32),
             "match_l": mk((npairs, cap), np.int32), "match_r": mk((npairs, cap), np.int32)}
        if kept:
            a.update({"pix_l": mk((npairs, cap, 2), np.float64), "pix_r": mk((npairs, cap, 2), np.float64),
                      "cur_desc_l": mk((npairs, cap, 32), np.uint8), "cur_desc_r": mk((npairs, cap, 32), np.uint8),
                      "cur_mask": mk((npairs, cap), np.uint8),
                      "circ_cur": mk((npairs, cap), np.int32), "circ_hist": mk((npairs, cap), np.int32)})
        if full:
            a.update({"kp_l": mk((npairs, cap), KP_DTYPE), "kp_r": mk((npairs, cap), KP_DTYPE),
                      "desc_l": mk((npairs, cap, 32), np.uint8), "desc_r": mk((npairs, cap, 32), np.uint8),
                      "train_idx": mk((npairs, cap), np.int32), "dist": mk((npairs, cap), np.int32),
                      "circ_cur": mk((npairs, cap), np.int32), "circ_hist": mk((npairs, cap), np.int32)})
        if getattr(self, "_proj", False):
            a.update({"points4d": mk((npairs, 4, cap), np.float64), "point_cam": mk((npairs, cap, 3), np.float64),
                      "depth_flags": mk((npairs, cap), np.uint8)})
        s = B2StereoOut()
        s.cap = cap
        for k, v in a.items():
            setattr(s, k, v.ctypes.data)
        return a, s

    def stereo_frontend(self, images_l, images_r, temporal=False, out=None):
        images_l = np.ascontiguousarray(images_l, np.uint8); images_r = np.ascontiguousarray(images_r, np.uint8)
        if images_l.ndim == 2:
            images_l, images_r = images_l[None], images_r[None]
        npairs, hgt, w = images_l.shape
        arrays, s = out if out is not None else self.alloc_stereo_out(npairs)
        self.check(self.L.b2orb_stereo_frontend(self.h, _p(images_l), _p(images_r), npairs, w, hgt, w, w * hgt,
                                                int(temporal), C.byref(s)))   # temporal: 0 / 1 (previous pair) / 2 (pool)
        return arrays

    def stereo_frontend_device(self, d_ptr_l, d_ptr_r, npairs, w, hgt, step, image_stride, temporal=False):
        self.check(self.L.b2orb_stereo_frontend_device(self.h, d_ptr_l, d_ptr_r, npairs, w, hgt, step, image_stride,
                                                       int(temporal)))

    def stereo_download(self, npairs, out):
        self.check(self.L.b2orb_stereo_download(self.h, npairs, C.byref(out[1])))
        return out[0]

    def set_rectification(self, map1_l=None, map2_l=None, map1_r=None, map2_r=None):
        """Maps of cv::initUndistortRectifyMap(..., CV_16SC2) per eye (include/cfsd/camera-model.hpp:76-77): map1 H x W x 2
        int16, map2 H x W uint16.  Afterwards stereo_frontend rectifies raw camera frames on the device
        (cv::remap, src/feature-tracker.cpp:75-76).  No arguments: off."""
        if map1_l is None:
            self.check(self.L.b2orb_set_rectification(self.h, 0, 0, None, None, None, None))
            self._maps = None
            return
        maps = [np.ascontiguousarray(map1_l, np.int16), np.ascontiguousarray(map2_l, np.uint16),
                np.ascontiguousarray(map1_r, np.int16), np.ascontiguousarray(map2_r, np.uint16)]
        hgt, w = maps[1].shape
        assert maps[0].shape == (hgt, w, 2) and maps[2].shape == (hgt, w, 2) and maps[3].shape == (hgt, w)
        self.check(self.L.b2orb_set_rectification(self.h, w, hgt, *[m.ctypes.data for m in maps]))
        self._maps = maps

    def remap(self, image, eye):
        """cv::remap of one image with the maps of set_rectification (eye 0 = left, 1 = right)."""
        image = np.ascontiguousarray(image, np.uint8)
        hgt, w = image.shape
        out = np.zeros((hgt, w), np.uint8)
        self.check(self.L.b2orb_remap(self.h, eye, _p(image), w, hgt, w, _p(out), w))
        return out

    def cfsd_ingest(self, frame, width, height):
        """src/cfsd-state-estimation.cpp:117-127: (H, W, 4) u8 side-by-side frame -> (grayL, grayR), height x width each."""
        frame = np.ascontiguousarray(frame, np.uint8)
        sh, sw, _ = frame.shape
        gl = np.zeros((height, width), np.uint8); gr = np.zeros((height, width), np.uint8)
        self.check(self.L.b2orb_cfsd_ingest(self.h, frame.ctypes.data, sw, sh, sw * 4, width, height, _p(gl), _p(gr), width))
        return gl, gr

    def stereo_frontend_packed(self, frames, width, height, temporal=False, out=None):
        """frames: (N, H, W, 4) u8 side-by-side colour frames; the ingest and the stereo front end run on the device."""
        frames = np.ascontiguousarray(frames, np.uint8)
        if frames.ndim == 3:
            frames = frames[None]
        n, sh, sw, _ = frames.shape
        arrays, s = out if out is not None else self.alloc_stereo_out(n)
        self.check(self.L.b2orb_stereo_frontend_packed(self.h, frames.ctypes.data, n, sw, sh, sw * 4, sw * sh * 4, width,
                                                       height, 1 if temporal else 0, C.byref(s)))
        return arrays

    # ---- device-resident feature pool (next row f3)
    def pool_reset(self, max_feature_age=8):
        self.check(self.L.b2orb_pool_reset(self.h, max_feature_age))

    @property
    def pool_size(self):
        return self.L.b2orb_pool_size(self.h)

    def _pool_lists(self, fn, *lead):
        cap = self.keypoint_capacity
        cc = np.zeros(cap, np.int32); ch = np.zeros(cap, np.int32); mask = np.zeros(cap, np.uint8); ids = np.zeros(cap, np.uint64)
        n, ncur, nl, nr = C.c_int(0), C.c_int(0), C.c_int(0), C.c_int(0)
        self.check(fn(self.h, *lead, _p(cc), _p(ch), C.byref(n), _p(mask), _p(ids), C.byref(ncur), C.byref(nl), C.byref(nr)))
        k = n.value
        return {"circ_cur": cc[:k].copy(), "circ_hist": ch[:k].copy(), "mask": mask[:ncur.value].copy(),
                "matched_ids": ids[:k].copy(), "n_left": nl.value, "n_right": nr.value}

    def pool_track(self, pair=0):
        """externalTrack(false) of the kept matches of `pair` of the last stereo_frontend call against the pool."""
        return self._pool_lists(self.L.b2orb_pool_track, pair)

    def pool_last_track(self):
        """Lists of the last track (after stereo_frontend(..., temporal=2): pair 0)."""
        return self._pool_lists(self.L.b2orb_pool_last_track)

    def pool_update(self, is_keyframe, accept=None):
        """featurePoolUpdate; returns (new feature ids, erased count)."""
        cap = self.keypoint_capacity
        ids = np.zeros(cap, np.uint64)
        ins, er = C.c_int(0), C.c_int(0)
        acc = np.ascontiguousarray(accept, np.uint8) if accept is not None else None
        self.check(self.L.b2orb_pool_update(self.h, 1 if is_keyframe else 0, _p(acc), _p(ids), C.byref(ins), C.byref(er)))
        return ids[:ins.value].copy(), er.value

    def pool_state(self):
        """(_histFeatureIDs, ages, _histDescriptorsL, _histDescriptorsR) in pool order (copies the rows back: tests)."""
        n = C.c_int(0)
        self.check(self.L.b2orb_pool_get(self.h, None, None, None, None, 0, C.byref(n)))
        m = max(n.value, 1)
        ids = np.zeros(m, np.uint64); ages = np.zeros(m, np.int32); hl = np.zeros((m, 32), np.uint8); hr = np.zeros((m, 32), np.uint8)
        self.check(self.L.b2orb_pool_get(self.h, _p(ids), _p(ages), _p(hl), _p(hr), m, C.byref(n)))
        return ids[:n.value].copy(), ages[:n.value].copy(), hl[:n.value].copy(), hr[:n.value].copy()

    def sync(self):
        self.check(self.L.b2orb_sync(self.h))

    def reset_history(self):
        self.L.b2orb_reset_history(self.h)

    # ---- debug
    def debug_set(self, key, value):
        """Pin a launch-time choice on this handle (include/b2orb.h: b2orb_debug_set)."""
        self.check(self.L.b2orb_debug_set(self.h, key.encode(), int(value)))
        return self

    def debug_level(self, slot, level, blurred=False):
        w, hh = C.c_int(), C.c_int()
        self.check(self.L.b2orb_debug_level_size(self.h, level, C.byref(w), C.byref(hh)))
        out = np.zeros((hh.value, w.value), np.uint8)
        self.check(self.L.b2orb_debug_get_level(self.h, slot, level, 1 if blurred else 0, _p(out)))
        return out

    def debug_octree_fallbacks(self):
        return int(self.L.b2orb_debug_octree_fallbacks(self.h))

    def debug_candidates(self, slot, level):
        n = C.c_int(0)
        self.check(self.L.b2orb_debug_get_candidates(self.h, slot, level, None, 0, C.byref(n)))
        out = np.zeros(max(n.value, 1), np.uint32)
        self.check(self.L.b2orb_debug_get_candidates(self.h, slot, level, _p(out), n.value, C.byref(n)))
        v = out[:n.value]
        return np.stack([v & 0xFFF, (v >> 12) & 0xFFF, v >> 24], 1).astype(np.int32)


class ORBextractor:
    """ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) — same ctor arguments and
    operator()(image, mask, keypoints, descriptors) semantics (mask ignored, as in the reference)."""

    def __init__(self, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST, max_width=4096, max_height=4096, device=0):
        self._args = (nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)
        self._device = device
        self._max = (max_width, max_height)
        self._h = None

    def _handle(self, w, hgt):
        if self._h is None:   # workspace is sized on first use, like the reference's lazily grown pyramid
            self._h = Handle(*self._args, max_width=min(self._max[0], max(w, 64)),
                             max_height=min(self._max[1], max(hgt, 64)), max_images=1, device=self._device)
        return self._h

    def __call__(self, image, mask=None):
        image = np.asarray(image)
        if image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        assert image.dtype == np.uint8 and image.ndim == 2, "image.type() == CV_8UC1"
        return self._handle(image.shape[1], image.shape[0]).extract(image)

    def GetLevels(self): return self._args[2]
    def GetScaleFactor(self): return float(np.float32(self._args[1]))
    def _tables(self):
        h = self._h or self._handle(64, 64)
        return h.tables()
    def GetScaleFactors(self): return self._tables()[0]
    def GetInverseScaleFactors(self): return self._tables()[1]
    def GetScaleSigmaSquares(self): return self._tables()[2]
    def GetInverseScaleSigmaSquares(self): return self._tables()[3]


class BFMatcher:
    """cv::BFMatcher(cv::NORM_HAMMING).match(query, train) -> (trainIdx, distance) arrays in query order."""

    def __init__(self, handle):
        self.h = handle

    def match(self, query, train):
        return self.h.match(query, train)


class FeatureTracker:
    """Matching half of cfsd::FeatureTracker (src/feature-tracker.cpp).  Members keep the reference's names:
    _curPixelsL/R, _curDescriptorsL/R (internalMatch, :208-218), _histDescriptorsL/R (pool stand-in)."""

    def __init__(self, width, height, numberOfFeatures=1000, scaleFactor=1.2, levelPyramid=8, iniThFAST=20,
                 minThFAST=7, matchRatio=2.0, minMatchDist=30.0, maxVerticalPixelDist=0.1, max_pairs=1, device=0):
        self.h = Handle(numberOfFeatures, scaleFactor, levelPyramid, iniThFAST, minThFAST, width, height,
                        max_images=2 * max_pairs, device=device, match_ratio=matchRatio, min_match_dist=minMatchDist,
                        max_vertical_pixel_dist=maxVerticalPixelDist)
        self._curPixelsL = self._curPixelsR = np.zeros((0, 2), np.float64)
        self._curDescriptorsL = self._curDescriptorsR = np.zeros((0, 32), np.uint8)
        self._histDescriptorsL = self._histDescriptorsR = np.zeros((0, 32), np.uint8)

    def extractORB(self, flag, img):
        return self.h.extract(img)

    def setRectification(self, rmap):
        """rmap[eye] = (map1, map2) as CameraModel::_rmap (include/cfsd/camera-model.hpp:76-77, 128)."""
        self.h.set_rectification(rmap[0][0], rmap[0][1], rmap[1][0], rmap[1][1])

    def processImage(self, grayLeft, grayRight):
        """remap both eyes + internalMatch in one device pass (src/feature-tracker.cpp:69-76 then :140-229)."""
        return self.internalMatch(grayLeft, grayRight)

    def internalMatch(self, imgLeft, imgRight):
        o = self.h.stereo_frontend(imgLeft, imgRight)
        m = int(o["n_match"][0])
        il, ir = o["match_l"][0, :m], o["match_r"][0, :m]
        kl, kr = o["kp_l"][0], o["kp_r"][0]
        self._curPixelsL = np.stack([kl["x"][il], kl["y"][il]], 1).astype(np.float64)
        self._curPixelsR = np.stack([kr["x"][ir], kr["y"][ir]], 1).astype(np.float64)
        self._curDescriptorsL = o["desc_l"][0][il].copy()
        self._curDescriptorsR = o["desc_r"][0][ir].copy()
        if "points4d" in o:
            self._points4D = o["points4d"][0][:, :m].copy()
            self._pointsCam = o["point_cam"][0][:m].copy()
            self._depthFlags = o["depth_flags"][0][:m].copy()
        return o

    def setProjection(self, P1, P2, maxDepth):
        """CameraModel::_P1, _P2 and the maxDepth key: internalMatch then also triangulates on the device."""
        self.h.set_projection(P1, P2, maxDepth)

    def triangulate(self):
        """cv::triangulatePoints(_P1, _P2, _curPixelsL, _curPixelsR, points4D) of featurePoolUpdate
        (src/feature-tracker.cpp:365-379): returns (points4D 4 x M, point_wrt_cam M x 3, depth flags M) of the kept
        matches of the last internalMatch — computed in that same device pass."""
        return self._points4D, self._pointsCam, self._depthFlags

    def externalTrack(self):
        """Matching parts with useRANSAC=false; returns (curIdx, histIdx) of circular matches."""
        cc, ch, nl, nr = self.h.temporal_track(self._curDescriptorsL, self._curDescriptorsR, self._histDescriptorsL,
                                               self._histDescriptorsR)
        return cc, ch

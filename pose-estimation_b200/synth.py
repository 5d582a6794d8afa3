"""Deterministic synthetic textured stereo sequences — inputs for tests and bench.py.

Library independent (numpy PCG64 only): uniform u8 noise on a canvas, `passes` rounds of a 3x3
integer box filter ((sum+4)//9), integer min-max stretch to 0..255.  Left(t) is a window of the
canvas that drifts 2 px/frame in x and 1 px every 2 frames in y; Right(t) is the same window with
each row band shifted left by an integer disparity in {4..40}, so true matches lie on the same row
(the reference's row filter |yL - yR| < 0.1 px, config/euroc.yml:112, only passes equal rows).
`passes=8` gives the low-texture variant that exercises the minThFAST fallback.
"""
import numpy as np

SHAPES = {
    # name: (width, height, nfeatures, iniThFAST, minThFAST)   reference config
    "euroc": (752, 480, 1000, 20, 7),    # config/euroc.yml:10-11,89-91,106-107 (yml has 800 feats)
    "kitti": (1241, 376, 2000, 12, 7),   # BASELINE.json shape; config/kitti.yml:90,107-108
    "kitti_yml": (1392, 512, 2000, 12, 7),  # config/kitti.yml:12-13
    "cfsd": (672, 376, 1000, 20, 7),     # config/cfsd.yml:21-22,101,118-119
    "4k": (3840, 2160, 8000, 20, 7),     # BASELINE.json stress config
}
SCALE_FACTOR = 1.2       # config/*.yml scaleFactor
N_LEVELS = 8             # config/*.yml levelPyramid
MATCH_RATIO = 2.0        # matchRatio
MIN_MATCH_DIST = 30.0    # minMatchDist
MAX_VERTICAL_PIXEL_DIST = 0.1  # maxVerticalPixelDist

_T = 64  # drift period / canvas slack


def _box3(a):
    p = np.pad(a.astype(np.int32), 1, mode="edge")
    s = np.zeros(a.shape, np.int32)
    for dy in range(3):
        for dx in range(3):
            s += p[dy:dy + a.shape[0], dx:dx + a.shape[1]]
    return (s + 4) // 9


def canvas(width, height, seed=0, passes=3):
    rng = np.random.Generator(np.random.PCG64(seed))
    c = rng.integers(0, 256, size=(height + _T, width + 64 + _T), dtype=np.uint8).astype(np.int32)
    for _ in range(passes):
        c = _box3(c)
    lo, hi = int(c.min()), int(c.max())
    c = (c - lo) * 255 // max(hi - lo, 1)
    return c.astype(np.uint8)


def stereo_pair(cv, width, height, t, seed=0):
    """Frame t of the sequence on canvas `cv` -> (left, right) u8 arrays of (height, width)."""
    y0 = (t // 2) % _T
    x0 = 48 + (2 * t) % _T
    left = cv[y0:y0 + height, x0:x0 + width]
    right = np.empty_like(left)
    band = 32
    rng = np.random.Generator(np.random.PCG64(seed * 7919 + 17))
    disp = rng.integers(4, 41, size=(height + _T) // band + 2)
    for r0 in range(0, height, band):
        d = int(disp[(y0 + r0) // band])
        r1 = min(r0 + band, height)
        right[r0:r1] = cv[y0 + r0:y0 + r1, x0 - d:x0 - d + width]
    return np.ascontiguousarray(left), right


def sequence(shape, nframes, seed=0, passes=3):
    """(L, R) arrays of shape (nframes, H, W) for a named shape or a (w, h) tuple."""
    w, h = SHAPES[shape][:2] if isinstance(shape, str) else shape
    cv = canvas(w, h, seed, passes)
    ls = np.empty((nframes, h, w), np.uint8)
    rs = np.empty((nframes, h, w), np.uint8)
    for t in range(nframes):
        ls[t], rs[t] = stereo_pair(cv, w, h, t, seed)
    return ls, rs


def rectification_maps(w, h, eye):
    """Deterministic fixed-point rectification maps in the format cv::initUndistortRectifyMap(..., CV_16SC2, ...) produces
    (include/cfsd/camera-model.hpp:76-77): map1 (h, w, 2) int16 = integer source (x, y), map2 (h, w) uint16 =
    fx | fy << 5 in 1/32 px.  A radial distortion plus a small per-eye rotation and shift; pure float64 + rint, so the
    same bits everywhere without OpenCV.  Corner pixels sample outside the image (BORDER_CONSTANT taps)."""
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float64)
    cx, cy, f = 0.5 * w + 3.25, 0.5 * h - 2.5, 0.55 * w
    th = (0.004, -0.003)[eye]
    xn, yn = (xs - cx) / f, (ys - cy) / f
    xr, yr = xn * np.cos(th) - yn * np.sin(th), xn * np.sin(th) + yn * np.cos(th)
    r2 = xr * xr + yr * yr
    d = 1.0 + (-0.17, -0.16)[eye] * r2 + 0.025 * r2 * r2
    X = cx + f * xr * d * 1.06 + (1.3, -0.7)[eye]
    Y = cy + f * yr * d * 1.06 + (-0.4, 0.6)[eye]
    ix, iy = np.rint(X * 32).astype(np.int64), np.rint(Y * 32).astype(np.int64)
    map1 = np.stack([ix >> 5, iy >> 5], -1).astype(np.int16)
    map2 = ((ix & 31) | ((iy & 31) << 5)).astype(np.uint16)
    return map1, map2


def packed_frame(left, right, src_w, src_h, seed=0):
    """A camera-style CV_8UC4 frame (both eyes side by side, src_w x src_h) whose grey content follows the given left /
    right images: nearest-neighbour upsample (pure integer index maps) with per-channel integer offsets so that the
    three colour channels differ, 4th channel = deterministic junk (cvtColor ignores it).  Library-independent."""
    h, w = left.shape
    both = np.concatenate([left, right], 1)
    yi = (np.arange(src_h) * h) // src_h
    xi = (np.arange(src_w) * (2 * w)) // src_w
    g = both[yi][:, xi].astype(np.int32)
    rng = np.random.Generator(np.random.PCG64(seed + 77))
    fr = np.empty((src_h, src_w, 4), np.uint8)
    fr[..., 0] = np.clip(g - 9, 0, 255)
    fr[..., 1] = np.clip(g + 4, 0, 255)
    fr[..., 2] = np.clip(g + (g >> 3) - 11, 0, 255)
    fr[..., 3] = rng.integers(0, 256, (src_h, src_w), dtype=np.uint8)
    return fr


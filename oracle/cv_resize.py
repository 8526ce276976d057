"""TEST INFRASTRUCTURE (oracle): cv::resize(src, dst, Size(w, h)) for 8-bit images with the default interpolation
(INTER_LINEAR), restated in numpy.  The reference calls it in ProcImage (yolo_core.cpp:108-109) before Mat2Image; the
algorithm itself lives in OpenCV (a system dependency the reference's CMake finds with find_package(OpenCV REQUIRED), no
version pinned; not part of /root/reference).  Restated from OpenCV's published imgproc/src/resize.cpp (4.x):

  * resize(): scale = 1 / ((double)dsize / ssize); INTER_LINEAR with an exact 2 x 2 decimation is replaced by INTER_AREA
    (`if (interpolation == INTER_LINEAR && is_area_fast && iscale_x == 2 && iscale_y == 2) interpolation = INTER_AREA`),
    whose fast path for 8-bit data is (a + b + c + d + 2) >> 2 (ResizeAreaFastVec / ResizeAreaFast_Invoker);
  * coordinate tables: fx = (float)((dx + 0.5) * scale - 0.5); sx = floor(fx); fx -= sx; sx < 0 -> (0, 0);
    sx >= ssize - 1 -> (ssize - 1, 0) — for x only: the y tables keep fy and clip the two row indices instead (found by
    comparing with the library: a clamped y table is off by one in the first / last rows of an enlargement);
    coefficients saturate_cast<short>((1 - fx) * 2048), saturate_cast<short>(fx * 2048) (cvRound: nearest, ties to even);
  * HResizeLinear<uchar, int, short, 2048>: row[dx] = S[sx] * a0 + S[sx + 1] * a1 (int, 11 fractional bits);
  * VResizeLinear<uchar, int, short, FixedPtCast<int, uchar, 22>>:
      dst = (((b0 * (row0 >> 4)) >> 16) + ((b1 * (row1 >> 4)) >> 16) + 2) >> 2.

PINNED against the real library: tests/golden/cv_resize.npz holds outputs of cv2.resize (opencv-python 4.13, the version in
the build container) for the cases of tests/golden/make_cv_resize_golden.py; tests/test_cv_resize_cpu.py checks this
restatement against every one of them bit for bit.  Only tests may import this module."""
import numpy as np

COEF_BITS = 11
COEF_SCALE = 1 << COEF_BITS


def _tables(ssize: int, dsize: int, clamp: bool = True):
    """source indices and the two 11-bit coefficients for every destination index.  clamp=True: the x tables (an index
    outside the row gets coefficient (2048, 0) on the border pixel); clamp=False: the y tables — resize() keeps the
    fractional weights there and only clips the two ROW indices (`clip(sy0 - ksize2 + 1 + k, 0, ssize.height)` in
    resizeGeneric_Invoker), so the border row is blended with itself through two truncating products"""
    inv_scale = float(dsize) / float(ssize)          # double
    scale = 1.0 / inv_scale                          # double
    d = np.arange(dsize, dtype=np.float64)
    f = ((d + 0.5) * scale - 0.5).astype(np.float32)
    s = np.floor(f).astype(np.int32)
    f = (f - s.astype(np.float32)).astype(np.float32)
    if clamp:
        lo = s < 0
        s[lo], f[lo] = 0, 0.0
        hi = s >= ssize - 1
        s[hi], f[hi] = ssize - 1, 0.0
    a0 = np.rint((np.float32(1.0) - f) * np.float32(COEF_SCALE)).astype(np.int32)   # rint: ties to even, like cvRound
    a1 = np.rint(f * np.float32(COEF_SCALE)).astype(np.int32)
    s1 = np.clip(s + 1, 0, ssize - 1)
    s = np.clip(s, 0, ssize - 1)
    return s, s1, a0, a1


def resize_u8(src: np.ndarray, dh: int, dw: int) -> np.ndarray:
    """src: uint8 [h][w][c] (or [h][w]); returns uint8 [dh][dw][c] == cv2.resize(src, (dw, dh))"""
    assert src.dtype == np.uint8
    squeeze = src.ndim == 2
    if squeeze:
        src = src[:, :, None]
    sh, sw, _ = src.shape
    if (dh, dw) == (sh, sw):
        out = src.copy()
    elif sh == 2 * dh and sw == 2 * dw:
        s = src.astype(np.int32)
        out = ((s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2).astype(np.uint8)
    else:
        x0, x1, a0, a1 = _tables(sw, dw)
        y0, y1, b0, b1 = _tables(sh, dh, clamp=False)
        s = src.astype(np.int32)
        rows = s[:, x0, :] * a0[None, :, None] + s[:, x1, :] * a1[None, :, None]      # [sh][dw][c], 11 fractional bits
        r0, r1 = rows[y0], rows[y1]                                                    # [dh][dw][c]
        out = ((((b0[:, None, None] * (r0 >> 4)) >> 16) + ((b1[:, None, None] * (r1 >> 4)) >> 16) + 2) >> 2)
        out = np.clip(out, 0, 255).astype(np.uint8)
    return out[:, :, 0] if squeeze else out

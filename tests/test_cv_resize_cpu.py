"""oracle/cv_resize.py (numpy restatement of cv::resize, 8-bit, INTER_LINEAR — what ProcImage calls, yolo_core.cpp:108-109)
against golden vectors produced by the real library (tests/golden/make_cv_resize_golden.py, cv2 4.13): bit for bit."""
import hashlib
import os

import numpy as np

from oracle import cv_resize as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cv_resize.npz")


def test_small_cases_bit_exact():
    g = np.load(GOLD)
    n = int(g["n_small"])
    assert n >= 30
    for i in range(n):
        src, dst = g[f"small{i}_src"], g[f"small{i}_dst"]
        got = R.resize_u8(src, dst.shape[0], dst.shape[1])
        assert np.array_equal(got, dst), f"case {i}: {src.shape} -> {dst.shape}"


def test_large_cases_by_digest():
    g = np.load(GOLD)
    for row in g["large"]:
        seed, sh, sw, dh, dw, digest = str(row).split()
        src = np.random.default_rng(int(seed)).integers(0, 256, (int(sh), int(sw), 3), dtype=np.uint8)
        got = R.resize_u8(src, int(dh), int(dw))
        assert hashlib.sha256(np.ascontiguousarray(got).tobytes()).hexdigest() == digest, f"{sh}x{sw} -> {dh}x{dw}"


def test_identity_and_area_path():
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, (12, 10, 3), dtype=np.uint8)
    assert np.array_equal(R.resize_u8(src, 12, 10), src)
    half = R.resize_u8(src, 6, 5)     # exact 2 x 2 decimation: INTER_AREA, (a + b + c + d + 2) >> 2
    s = src.astype(np.int32)
    assert np.array_equal(half, ((s[0::2, 0::2] + s[0::2, 1::2] + s[1::2, 0::2] + s[1::2, 1::2] + 2) >> 2).astype(np.uint8))

"""Golden vectors for oracle/cv_resize.py from the real library: cv2.resize (default INTER_LINEAR) of opencv-python in the
build container.  Run here (needs cv2): python tests/golden/make_cv_resize_golden.py -> tests/golden/cv_resize.npz.
Small cases are stored whole (source + result); large ones as the seed of the source and the SHA-256 of the result."""
import hashlib
import os

import cv2
import numpy as np

SMALL = [((17, 5), (33, 64)), ((2, 2), (7, 9)), ((1, 1), (4, 4)), ((40, 37), (20, 37)), ((64, 64), (32, 32)),
         ((64, 64), (32, 31)), ((5, 70), (3, 10)), ((17, 64), (33, 64)), ((33, 5), (33, 64)), ((3, 3), (48, 48)),
         ((1, 9), (5, 3)), ((9, 1), (3, 5)), ((48, 64), (61, 61)), ((61, 61), (61, 61)), ((30, 40), (64, 64))]
LARGE = [((480, 640), (608, 608)), ((1080, 1920), (608, 608)), ((720, 1280), (416, 416)), ((1216, 1216), (608, 608)),
         ((375, 500), (608, 608)), ((607, 609), (608, 608)), ((1200, 1600), (320, 320))]


def source(seed, sh, sw, c):
    return np.random.default_rng(seed).integers(0, 256, (sh, sw, c), dtype=np.uint8)


def main():
    out = {"cv2_version": np.array(cv2.__version__)}
    n = 0
    for i, ((sh, sw), (dh, dw)) in enumerate(SMALL):
        for c in (3, 1):
            src = source(1000 + i, sh, sw, c)
            dst = cv2.resize(src if c == 3 else src[:, :, 0], (dw, dh))
            out[f"small{n}_src"] = src
            out[f"small{n}_dst"] = dst.reshape(dh, dw, c)
            n += 1
    out["n_small"] = np.array(n)
    rows = []
    for i, ((sh, sw), (dh, dw)) in enumerate(LARGE):
        src = source(2000 + i, sh, sw, 3)
        dst = cv2.resize(src, (dw, dh))
        rows.append((2000 + i, sh, sw, dh, dw, hashlib.sha256(np.ascontiguousarray(dst).tobytes()).hexdigest()))
    out["large"] = np.array([f"{a} {b} {c} {d} {e} {h}" for a, b, c, d, e, h in rows])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cv_resize.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path), "bytes;", n, "small cases,", len(rows), "large cases; cv2", cv2.__version__)


if __name__ == "__main__":
    main()

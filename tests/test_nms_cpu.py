"""NmsSort parity on the CPU (no GPU needed: NMS is host code on both sides).

yc_nms_sort (csrc/network.cu) vs the UNMODIFIED reference NmsSort (box.cpp:393-419 through
oracle/_ref) on the same candidates: suppressed probabilities AND final array order identical,
bit for bit, including ties (both sides use libc qsort with the same comparator outcomes)."""
import numpy as np
import pytest

from oracle import pyoracle

pytestmark = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built")


def _cands(n, classes, seed, ties=False):
    rng = np.random.default_rng(seed)
    boxes = np.concatenate([rng.uniform(0.1, 0.9, (n, 2)), rng.uniform(0.05, 0.4, (n, 2))], axis=1).astype(np.float32)
    obj = rng.uniform(0.3, 1.0, n).astype(np.float32)
    probs = rng.uniform(0, 1, (n, classes)).astype(np.float32)
    probs[probs < 0.6] = 0  # sparse like thresholded detections
    if ties:
        probs = np.round(probs, 1)
        boxes[n // 2:] = boxes[:n - n // 2]  # duplicated boxes: IoU == 1
    return boxes, obj, probs


@pytest.mark.parametrize("n,classes,kind,ties", [(0, 3, 0, False), (1, 3, 0, False), (50, 4, 0, False), (300, 80, 0, False),
                                                  (200, 5, 1, False), (120, 3, 0, True), (120, 3, 1, True)])
def test_nms_sort_matches_reference(built, n, classes, kind, ties):
    from darknet_b200 import network as N
    boxes, obj, probs = _cands(n, classes, seed=n + classes, ties=ties)
    if n == 0:
        p, order = N.NmsSortArrays(np.zeros((0, 4), np.float32), np.zeros(0, np.float32), np.zeros((0, classes), np.float32), 0.45)
        assert p.shape == (0, classes)
        return
    rp, rorder = pyoracle.ref_nms(boxes, obj, probs, 0.45, kind, 0.6)
    gp, gorder = N.NmsSortArrays(boxes, obj, probs, 0.45, kind, 0.6)
    assert np.array_equal(gp, rp)
    assert np.array_equal(gorder, rorder)
    assert (rp == 0).sum() > (probs == 0).sum() or n < 10  # something was suppressed

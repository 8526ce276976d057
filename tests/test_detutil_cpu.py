"""The detection-identity checker itself (tests/detutil.py) against the reference's NmsSort: with all margins at zero
the interval NMS must decide every entry and agree with the compiled reference; with margins it may only move
entries into UNDECIDED, never flip a decided one."""
import numpy as np
import pytest

from oracle import pyoracle
from tests import detutil

needs_ref = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built")


def _random_dets(n, classes, seed):
    rng = np.random.default_rng(seed)
    boxes = np.concatenate([rng.uniform(0.2, 0.8, (n, 2)), rng.uniform(0.05, 0.4, (n, 2))], axis=1).astype(np.float32)
    obj = rng.uniform(0.3, 1.0, n).astype(np.float32)
    probs = (obj[:, None] * rng.uniform(0, 1, (n, classes))).astype(np.float32)
    probs[probs < 0.25] = 0
    return boxes, obj, probs


@needs_ref
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_interval_nms_equals_reference_at_zero_margin(seed):
    boxes, obj, probs = _random_dets(300, 6, seed)
    after, _ = pyoracle.ref_nms(boxes, obj, probs, 0.45)
    for c in range(probs.shape[1]):
        sel = np.nonzero(probs[:, c] > 0)[0]
        st = detutil.interval_nms(boxes[sel], probs[sel, c], np.ones(sel.size, bool), 0.45, 0.0, 0.0)
        assert not np.any(st == 2)
        assert np.array_equal(st == 1, after[sel, c] > 0)


@needs_ref
def test_interval_nms_margins_only_add_undecided():
    boxes, obj, probs = _random_dets(400, 3, 7)
    after, _ = pyoracle.ref_nms(boxes, obj, probs, 0.45)
    undecided = 0
    for c in range(probs.shape[1]):
        sel = np.nonzero(probs[:, c] > 0)[0]
        st = detutil.interval_nms(boxes[sel], probs[sel, c], probs[sel, c] > 0.28, 0.45, 0.03, 0.05)
        ref_alive = after[sel, c] > 0
        assert np.all(ref_alive[st == 1]) and not np.any(ref_alive[st == 0])
        undecided += int((st == 2).sum())
    assert undecided > 0


def test_compare_image_flags_index_mismatch():
    ids = np.array([[10, 0, 5], [10, 1, 7]], np.int32)
    rows = np.array([[.5, .5, .1, .1, .9], [.2, .2, .1, .1, .8]], np.float32)
    probs = np.array([[.7, 0], [0, .6]], np.float32)
    st = detutil.compare_image(ids, rows, probs, ids, rows, probs, probs, 0.25, 0.45)
    assert st["kept"] == 2 and st["undecided"] == 0
    other = ids.copy()
    other[1, 2] = 8  # same box, different cell: an index mismatch, not a proximity match
    with pytest.raises(AssertionError):
        detutil.compare_image(ids, rows, probs, other, rows, probs, probs, 0.25, 0.45)
    wrong_cls = probs[:, ::-1].copy()
    with pytest.raises(AssertionError):
        detutil.compare_image(ids, rows, probs, ids, rows, wrong_cls, wrong_cls, 0.25, 0.45)

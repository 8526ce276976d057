"""TEST INFRASTRUCTURE — detection-set identity between the reference and the B200 path, by INDEX.

north_star: "the final post-NMS detection set identical in class and index, with box coordinates within
tolerance".  A detection's index is (image, yolo layer, anchor, cell); its identity is (index, class).

The B200 path computes in bf16, so a score that sits within `margin` of the threshold may fall on either side, and
a greedy NMS decision that hinges on such a score (or on an IoU within `iou_margin` of nms_thresh, or on the order
of two scores closer than `margin`) may differ without either side being wrong.  The policy, stated exactly:

  candidates : every index whose objectness clears thresh by more than `margin` on one side must exist on the other;
  classes    : every (index, class) whose score clears thresh by more than `margin` on one side must be non-zero on the
               other;
  boxes      : centre within `tol_xy` (image units), size within `tol_wh` relative, for every common index;
  post-NMS   : an interval version of the reference's NmsSort (box.cpp:393-419) is run on the REFERENCE's candidates:
               each (index, class) ends as KEPT, SUPPRESSED or UNDECIDED, where UNDECIDED means the outcome depends on a
               comparison inside a margin.  Every KEPT entry must survive on the B200 side and every B200 survivor must
               be KEPT or UNDECIDED.  The share of UNDECIDED entries is reported and bounded, so the statement cannot
               become vacuous.

Nothing here is a box-proximity match: rows are tied together by index only.
"""
from __future__ import annotations

import numpy as np


def scan_head(dec, thresh, anchors, classes, gaussian):
    """Indices of one image's candidates in one YOLO head in the reference's scan order (GetYoloDetections,
    yolo_layer.cpp:794-832: anchor n outer, cell i inner, objectness > thresh).  dec: [C, H, W] decoded head.
    Returns (anchor[n], cell[n], objectness[n])."""
    C, H, W = dec.shape
    E = classes + (9 if gaussian else 5)
    obj = dec.reshape(anchors, E, H * W)[:, 8 if gaussian else 4, :]
    a, cell = np.nonzero(obj > np.float32(thresh))
    return a.astype(np.int32), cell.astype(np.int32), obj[a, cell]


def candidate_ids(heads, thresh):
    """heads: list of (layer_index, dec[C,H,W], anchors, classes, gaussian) in network order.
    Returns ids [n,3] = (layer, anchor, cell) and objectness [n], in the order GetNetworkBoxes emits them."""
    ids, objs = [], []
    for li, dec, anchors, classes, gaussian in heads:
        a, cell, o = scan_head(dec, thresh, anchors, classes, gaussian)
        ids.append(np.stack([np.full_like(a, li), a, cell], axis=1))
        objs.append(o)
    return np.concatenate(ids, axis=0), np.concatenate(objs)


def _key(ids):
    return (ids[:, 0].astype(np.int64) << 40) | (ids[:, 1].astype(np.int64) << 32) | ids[:, 2].astype(np.int64)


def box_iou(a, b):
    """Box::Iou (box.cpp:36-72) of every row of a [n,4] against every row of b [m,4] (x, y, w, h), fp32."""
    a = a.astype(np.float32)[:, None, :]
    b = b.astype(np.float32)[None, :, :]

    def overlap(x1, w1, x2, w2):
        left = np.maximum(x1 - w1 / 2, x2 - w2 / 2)
        right = np.minimum(x1 + w1 / 2, x2 + w2 / 2)
        return right - left
    w = overlap(a[..., 0], a[..., 2], b[..., 0], b[..., 2])
    h = overlap(a[..., 1], a[..., 3], b[..., 1], b[..., 3])
    inter = np.where((w < 0) | (h < 0), np.float32(0), w * h)
    union = a[..., 2] * a[..., 3] + b[..., 2] * b[..., 3] - inter
    eps = np.finfo(np.float32).eps
    with np.errstate(divide="ignore", invalid="ignore"):
        iou = np.where((np.abs(inter) < eps) | (np.abs(union) < eps), np.float32(0), inter / union)
    return iou.astype(np.float32)


def interval_nms(boxes, p, certain, nms_thresh, margin, iou_margin, boxes2=None):
    """Three-valued greedy NMS of ONE class.  boxes [n,4], p [n] scores (> 0), certain [n] bool (the candidate exists
    on both sides whatever the rounding); boxes2: the other side's boxes of the same candidates (an overlap then counts
    as certain only if BOTH sides' boxes say so, as possible if EITHER does).  Returns state [n]: 1 KEPT, 0 SUPPRESSED,
    2 UNDECIDED."""
    n = p.shape[0]
    order = np.argsort(-p, kind="stable")
    b, ps, cert = boxes[order], p[order], certain[order]
    iou = box_iou(b, b)
    iou_lo = iou_hi = iou
    if boxes2 is not None:
        iou2 = box_iou(boxes2[order], boxes2[order])
        iou_lo, iou_hi = np.minimum(iou, iou2), np.maximum(iou, iou2)
    hi = iou_lo > nms_thresh + iou_margin       # suppresses whatever the rounding
    lo = iou_hi > nms_thresh - iou_margin       # may suppress
    state = np.zeros(n, np.int8)
    for k in range(n):
        before = np.arange(n) < k
        sure_order = before & (ps - ps[k] > 2 * margin)   # both scores may move by up to `margin`
        if np.any(sure_order & hi[:, k] & (state == 1)):
            state[k] = 0
            continue
        # anything that may come first and may be alive and may overlap enough
        maybe_first = (before | (ps > ps[k] - 2 * margin)) & (np.arange(n) != k)
        alive = np.where(before, state != 0, True)
        if np.any(maybe_first & alive & lo[:, k]) or not cert[k]:
            state[k] = 2
        else:
            state[k] = 1
    out = np.zeros(n, np.int8)
    out[order] = state
    return out


def compare_image(ref_ids, ref_rows, ref_probs, our_ids, our_rows, our_probs_pre, our_probs_post, thresh, nms_thresh,
                  margin=0.03, iou_margin=0.05, tol_xy=2e-2, tol_wh=5e-2):
    """One image.  *_ids [n,3]; ref_rows [n,>=5] = x y w h objectness (pre-NMS, reference order); ref_probs [n,classes]
    pre-NMS; our_rows [m,>=5] likewise; our_probs_pre / our_probs_post [m,classes] before / after NmsSort.
    Returns a dict of counts; raises AssertionError with the offending indices on any violation."""
    rk, ok = _key(ref_ids), _key(our_ids)
    assert len(np.unique(rk)) == len(rk) and len(np.unique(ok)) == len(ok)
    r_of = {int(k): i for i, k in enumerate(rk)}
    o_of = {int(k): i for i, k in enumerate(ok)}
    thr = np.float32(thresh)
    # ---- candidates
    miss = [tuple(ref_ids[i]) for i, k in enumerate(rk) if ref_rows[i, 4] > thr + margin and int(k) not in o_of]
    extra = [tuple(our_ids[i]) for i, k in enumerate(ok) if our_rows[i, 4] > thr + margin and int(k) not in r_of]
    assert not miss and not extra, f"candidate indices differ outside the margin: missing {miss[:5]} extra {extra[:5]}"
    common = [(r_of[int(k)], i) for i, k in enumerate(ok) if int(k) in r_of]
    ri = np.array([c[0] for c in common], np.int64)
    oi = np.array([c[1] for c in common], np.int64)
    stats = dict(ref=len(rk), ours=len(ok), common=len(common), in_band=len(rk) + len(ok) - 2 * len(common))
    if len(common):
        # ---- classes
        rp, op = ref_probs[ri], our_probs_pre[oi]
        bad = np.argwhere(((rp > thr + margin) & (op <= 0)) | ((op > thr + margin) & (rp <= 0)))
        assert bad.shape[0] == 0, ("class sets differ outside the margin at (index, class): " +
                                   str([(tuple(ref_ids[ri[a]]), int(c)) for a, c in bad[:5]]))
        both = (rp > 0) & (op > 0)
        stats["max_dscore"] = float(np.abs(rp - op)[both].max()) if both.any() else 0.0
        stats["max_dobj"] = float(np.abs(ref_rows[ri, 4] - our_rows[oi, 4]).max())
        # ---- boxes
        dxy = np.abs(ref_rows[ri, 0:2] - our_rows[oi, 0:2]).max()
        dwh = (np.abs(ref_rows[ri, 2:4] - our_rows[oi, 2:4]) / np.maximum(np.abs(ref_rows[ri, 2:4]), 1e-6)).max()
        stats["max_dxy"], stats["max_dwh_rel"] = float(dxy), float(dwh)
        assert dxy <= tol_xy and dwh <= tol_wh, f"boxes differ: centre {dxy:.3e} (tol {tol_xy}) size {dwh:.3e} rel (tol {tol_wh})"
    # ---- post-NMS by interval NMS on the reference's candidates (union with ours for the in-band extras)
    if nms_thresh > 0:
        classes = ref_probs.shape[1]
        only_ours = np.array([i for i, k in enumerate(ok) if int(k) not in r_of], np.int64)
        boxes = np.concatenate([ref_rows[:, :4], our_rows[only_ours, :4]], axis=0)
        boxes_other = boxes.copy()   # the B200 side's box of every candidate it has
        for i, k in enumerate(rk):
            if int(k) in o_of:
                boxes_other[i] = our_rows[o_of[int(k)], :4]
        probs = np.concatenate([ref_probs, our_probs_pre[only_ours]], axis=0)
        keys = np.concatenate([rk, ok[only_ours]])
        exists_other = np.array([int(k) in o_of for k in rk] + [False] * len(only_ours))
        kept = undecided = suppressed = 0
        for c in range(classes):
            sel = np.nonzero(probs[:, c] > 0)[0]
            if sel.size == 0:
                continue
            p = probs[sel, c]
            cert = exists_other[sel] & (p > thr + margin)
            st = interval_nms(boxes[sel], p, cert, nms_thresh, margin, iou_margin, boxes_other[sel])
            kept += int((st == 1).sum()); undecided += int((st == 2).sum()); suppressed += int((st == 0).sum())
            for j, s in zip(sel, st):
                k = int(keys[j])
                alive = k in o_of and our_probs_post[o_of[k], c] > 0
                if s == 1:
                    assert alive, f"post-NMS: reference keeps (index {k >> 40, (k >> 32) & 255, k & 0xffffffff}, class {c}) and it is robust, ours dropped it"
                elif s == 0:
                    assert not alive, f"post-NMS: reference suppresses (index {k >> 40, (k >> 32) & 255, k & 0xffffffff}, class {c}) robustly, ours keeps it"
        # survivors of ours that the reference never had as a candidate of that class
        for i in range(our_probs_post.shape[0]):
            k = int(ok[i])
            for c in np.nonzero(our_probs_post[i] > 0)[0]:
                if k in r_of:
                    assert ref_probs[r_of[k], c] > 0 or our_probs_pre[i, c] <= thr + margin
        stats.update(kept=kept, undecided=undecided, suppressed=suppressed)
    return stats

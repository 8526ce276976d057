"""cfg generators for the five BASELINE.json configs.

The reference ships no cfg files (`.gitignore:25` excludes `cfg/`), so the architectures are
regenerated here from the upstream layer lists in the INI dialect the
reference parser reads (`src/parser.cpp:59-100`; conv keys `:179-242`, yolo `:312-415`,
Gaussian_yolo `:443-557`, maxpool `:640-659`, shortcut `:720-779`, route/upsample `:819-896`).
Self-check: the reference's own parser prints Total BFLOPS 6.910 / 65.879 / 128.459 for
yolov4-tiny-416 / yolov3-416 / yolov4-608 built from these strings (tests/test_cfgs.py).
"""
from __future__ import annotations


class _Cfg:
    def __init__(self, w, h, batch, subdivisions, extra_net=None):
        self.s = []
        self.n = 0  # number of layers emitted so far
        net = {
            "batch": batch, "subdivisions": subdivisions, "width": w, "height": h, "channels": 3,
            "momentum": 0.9, "decay": 0.0005, "learning_rate": 0.001, "burn_in": 0,
            "max_epoch": 300, "policy": "constant",
        }
        if extra_net:
            net.update(extra_net)
        self.s.append("[net]\n" + "".join(f"{k}={v}\n" for k, v in net.items()))

    def add(self, name, **kv):
        self.s.append(f"\n[{name}]\n" + "".join(f"{k}={v}\n" for k, v in kv.items()))
        self.n += 1
        return self.n - 1

    def conv(self, f, k, s, act="leaky", bn=1):
        kv = {}
        if bn:
            kv["batch_normalize"] = 1
        kv.update(filters=f, size=k, stride=s, pad=1, activation=act)
        return self.add("convolutional", **kv)

    def shortcut(self, frm, act="linear"):
        return self.add("shortcut", **{"from": frm, "activation": act})

    def route(self, *layers, **kv):
        return self.add("route", layers=",".join(str(x) for x in layers), **kv)

    def maxpool(self, size, stride):
        return self.add("maxpool", size=size, stride=stride)

    def upsample(self, stride=2):
        return self.add("upsample", stride=stride)

    def text(self):
        return "".join(self.s)


V3_ANCHORS = "10,13, 16,30, 33,23, 30,61, 62,45, 59,119, 116,90, 156,198, 373,326"
V4_ANCHORS = "12,16, 19,36, 40,28, 36,75, 76,55, 72,146, 142,110, 192,243, 459,401"
TINY_ANCHORS = "10,14, 23,27, 37,58, 81,82, 135,169, 344,319"


def yolov3(w=416, h=416, batch=1, subdivisions=1, gaussian=False, classes=80):
    """yolov3 / Gaussian-yolov3: 107 layers (75 conv, 23 shortcut, 4 route, 2 upsample, 3 heads)."""
    c = _Cfg(w, h, batch, subdivisions)
    head_f = 3 * (classes + (9 if gaussian else 5))

    def res(ch, n):
        for _ in range(n):
            c.conv(ch // 2, 1, 1)
            c.conv(ch, 3, 1)
            c.shortcut(-3)

    def head(ch, mask):
        for _ in range(3):
            c.conv(ch, 1, 1)
            c.conv(ch * 2, 3, 1)
        c.conv(head_f, 1, 1, act="linear", bn=0)
        if gaussian:
            c.add("Gaussian_yolo", mask=mask, anchors=V3_ANCHORS, classes=classes, num=9, jitter=.3,
                  ignore_thresh=.7, truth_thresh=1, iou_thresh=0.213, uc_normalizer=1.0,
                  cls_normalizer=1.0, iou_normalizer=0.5, iou_loss="giou", scale_x_y=1.0, random=0)
        else:
            c.add("yolo", mask=mask, anchors=V3_ANCHORS, classes=classes, num=9, jitter=.3,
                  ignore_thresh=.7, truth_thresh=1, random=0)

    c.conv(32, 3, 1)
    c.conv(64, 3, 2)
    res(64, 1)
    c.conv(128, 3, 2)
    res(128, 2)
    c.conv(256, 3, 2)
    res(256, 8)
    c.conv(512, 3, 2)
    res(512, 8)
    c.conv(1024, 3, 2)
    res(1024, 4)
    head(512, "6,7,8")
    c.route(-4)
    c.conv(256, 1, 1)
    c.upsample(2)
    c.route(-1, 61)
    head(256, "3,4,5")
    c.route(-4)
    c.conv(128, 1, 1)
    c.upsample(2)
    c.route(-1, 36)
    head(128, "0,1,2")
    assert c.n == 107, c.n
    return c.text()


def gaussian_yolov3(w=608, h=608, batch=1, subdivisions=1, classes=80):
    return yolov3(w, h, batch, subdivisions, gaussian=True, classes=classes)


def yolov4_tiny(w=416, h=416, batch=1, subdivisions=1, classes=80):
    """yolov4-tiny: 38 layers (21 conv, 11 route, 3 maxpool, 1 upsample, 2 yolo), all leaky."""
    c = _Cfg(w, h, batch, subdivisions)
    head_f = 3 * (classes + 5)
    ykw = dict(anchors=TINY_ANCHORS, classes=classes, num=6, jitter=.3, scale_x_y=1.05,
               cls_normalizer=1.0, iou_normalizer=0.07, iou_loss="ciou", ignore_thresh=.7,
               truth_thresh=1, random=0, nms_kind="greedynms", beta_nms=0.6)

    def csp(ch):
        c.conv(ch, 3, 1)
        c.route(-1, groups=2, group_id=1)
        c.conv(ch // 2, 3, 1)
        c.conv(ch // 2, 3, 1)
        c.route(-1, -2)
        c.conv(ch, 1, 1)
        c.route(-6, -1)
        c.maxpool(2, 2)

    c.conv(32, 3, 2)
    c.conv(64, 3, 2)
    csp(64)
    csp(128)
    csp(256)
    c.conv(512, 3, 1)
    c.conv(256, 1, 1)
    c.conv(512, 3, 1)
    c.conv(head_f, 1, 1, act="linear", bn=0)
    c.add("yolo", mask="3,4,5", **ykw)
    c.route(-4)
    c.conv(128, 1, 1)
    c.upsample(2)
    c.route(-1, 23)
    c.conv(256, 3, 1)
    c.conv(head_f, 1, 1, act="linear", bn=0)
    c.add("yolo", mask="1,2,3", **ykw)
    assert c.n == 38, c.n
    return c.text()


def yolov4(w=608, h=608, batch=1, subdivisions=1, classes=80):
    """yolov4: 162 layers (110 conv, 23 shortcut, 21 route, 3 maxpool, 2 upsample, 3 yolo)."""
    c = _Cfg(w, h, batch, subdivisions, extra_net={"mosaic": 1})
    head_f = 3 * (classes + 5)
    M = "mish"

    def yolo(mask, sxy):
        c.add("yolo", mask=mask, anchors=V4_ANCHORS, classes=classes, num=9, jitter=.3,
              ignore_thresh=.7, truth_thresh=1, scale_x_y=sxy, iou_thresh=0.213, cls_normalizer=1.0,
              iou_normalizer=0.07, iou_loss="ciou", nms_kind="greedynms", beta_nms=0.6, max_delta=5,
              random=0)

    c.conv(32, 3, 1, M)
    # stage 1
    c.conv(64, 3, 2, M)
    c.conv(64, 1, 1, M)
    c.route(-2)
    c.conv(64, 1, 1, M)
    c.conv(32, 1, 1, M)
    c.conv(64, 3, 1, M)
    c.shortcut(-3)
    c.conv(64, 1, 1, M)
    c.route(-1, -7)
    c.conv(64, 1, 1, M)

    def stage(ch, n):
        c.conv(ch, 3, 2, M)
        c.conv(ch // 2, 1, 1, M)
        c.route(-2)
        c.conv(ch // 2, 1, 1, M)
        for _ in range(n):
            c.conv(ch // 2, 1, 1, M)
            c.conv(ch // 2, 3, 1, M)
            c.shortcut(-3)
        c.conv(ch // 2, 1, 1, M)
        c.route(-1, -(4 + 3 * n))
        c.conv(ch, 1, 1, M)

    stage(128, 2)
    stage(256, 8)
    assert c.n - 1 == 54, c.n
    stage(512, 8)
    assert c.n - 1 == 85, c.n
    stage(1024, 4)
    c.conv(512, 1, 1)
    c.conv(1024, 3, 1)
    c.conv(512, 1, 1)
    # SPP
    c.maxpool(5, 1)
    c.route(-2)
    c.maxpool(9, 1)
    c.route(-4)
    c.maxpool(13, 1)
    c.route(-1, -3, -5, -6)
    c.conv(512, 1, 1)
    c.conv(1024, 3, 1)
    c.conv(512, 1, 1)
    c.conv(256, 1, 1)
    c.upsample(2)
    c.route(85)
    c.conv(256, 1, 1)
    c.route(-1, -3)
    for _ in range(2):
        c.conv(256, 1, 1)
        c.conv(512, 3, 1)
    c.conv(256, 1, 1)
    c.conv(128, 1, 1)
    c.upsample(2)
    c.route(54)
    c.conv(128, 1, 1)
    c.route(-1, -3)
    for _ in range(2):
        c.conv(128, 1, 1)
        c.conv(256, 3, 1)
    c.conv(128, 1, 1)
    c.conv(256, 3, 1)
    c.conv(head_f, 1, 1, act="linear", bn=0)
    yolo("0,1,2", 1.2)
    c.route(-4)
    c.conv(256, 3, 2)
    c.route(-1, -16)
    for _ in range(2):
        c.conv(256, 1, 1)
        c.conv(512, 3, 1)
    c.conv(256, 1, 1)
    c.conv(512, 3, 1)
    c.conv(head_f, 1, 1, act="linear", bn=0)
    yolo("3,4,5", 1.1)
    c.route(-4)
    c.conv(512, 3, 2)
    c.route(-1, -37)
    for _ in range(2):
        c.conv(512, 1, 1)
        c.conv(1024, 3, 1)
    c.conv(512, 1, 1)
    c.conv(1024, 3, 1)
    c.conv(head_f, 1, 1, act="linear", bn=0)
    yolo("6,7,8", 1.05)
    assert c.n == 162, c.n
    return c.text()


def micro(w=64, h=64, batch=1, subdivisions=1, classes=3, act="leaky", gaussian=False):
    """A small net touching every layer kind on the hot path (conv 3x3 s1/s2, 1x1, mish/leaky,
    shortcut, route with groups, multi-route, SPP maxpool, 2/2 maxpool, upsample, two heads).
    Used for fast parity tests and the committed golden fixtures."""
    c = _Cfg(w, h, batch, subdivisions)
    head_f = 3 * (classes + (9 if gaussian else 5))
    anchors = "4,5, 8,10, 13,16, 20,24, 30,35, 40,50"

    def head(mask, sxy):
        if gaussian:
            c.add("Gaussian_yolo", mask=mask, anchors=anchors, classes=classes, num=6, jitter=.3,
                  ignore_thresh=.7, truth_thresh=1, iou_thresh=0.213, uc_normalizer=1.0,
                  cls_normalizer=1.0, iou_normalizer=0.5, iou_loss="giou", scale_x_y=sxy, random=0)
        else:
            c.add("yolo", mask=mask, anchors=anchors, classes=classes, num=6, jitter=.3,
                  scale_x_y=sxy, cls_normalizer=1.0, iou_normalizer=0.07, iou_loss="ciou",
                  ignore_thresh=.7, truth_thresh=1, random=0, iou_thresh=0.213)

    c.conv(16, 3, 1, "mish")          # 0
    c.conv(32, 3, 2, act)             # 1   /2
    c.conv(16, 1, 1, act)             # 2
    c.conv(32, 3, 1, "mish")          # 3
    c.shortcut(-3)                    # 4
    c.route(-1, groups=2, group_id=1)  # 5   16 ch
    c.conv(32, 3, 1, act)             # 6
    c.route(-1, -3)                   # 7   48 ch
    c.maxpool(2, 2)                   # 8   /4
    c.conv(64, 3, 1, act)             # 9
    c.maxpool(5, 1)                   # 10
    c.route(-2)                       # 11
    c.maxpool(9, 1)                   # 12
    c.route(-1, -3, -4)               # 13  192 ch
    c.conv(64, 1, 1, act)             # 14
    c.conv(64, 3, 2, act)             # 15  /8
    c.conv(head_f, 1, 1, "linear", bn=0)  # 16
    head("3,4,5", 1.05)               # 17
    c.route(-3)                       # 18  (layer 15)
    c.conv(32, 1, 1, act)             # 19
    c.upsample(2)                     # 20  /4
    c.route(-1, 9)                    # 21  96 ch
    c.conv(64, 3, 1, act)             # 22
    c.conv(head_f, 1, 1, "linear", bn=0)  # 23
    head("0,1,2", 1.2)                # 24
    return c.text()


CONFIGS = {
    "yolov4-tiny": yolov4_tiny,
    "yolov3": yolov3,
    "yolov4": yolov4,
    "gaussian-yolov3": gaussian_yolov3,
    "micro": micro,
}

"""Host-side mirror of the reference's network API for the hot path, over the C ABI.

Same names and argument meaning as the reference (`src/yolo_core.h:625-663`, `src/network.h:8-37`,
`src/box.h:93-97`): LoadNetwork / NetworkPredict / GetNetworkBoxes / NmsSort / FreeNetwork.  All the
work happens in libyc_b200.so (csrc/network.cu + the kernels); this module only marshals numpy
buffers.  There is no CPU fallback: every call raises YcError if the library or the GPU is missing.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib as L


class Network:
    """yc_network handle (the reference's `Network`, yolo_core.h:453-546)."""

    def __init__(self, handle):
        self.h = handle
        lib = L.lib()
        self.n = lib.yc_network_info(self.h, 0)
        self.batch = lib.yc_network_info(self.h, 1)
        self.w = lib.yc_network_info(self.h, 2)
        self.hh = lib.yc_network_info(self.h, 3)
        self.c = lib.yc_network_info(self.h, 4)

    # ---- introspection --------------------------------------------------------------------------
    def info(self, what):
        return L.lib().yc_network_info(self.h, what)

    def layer_info(self, i, what):
        return L.lib().yc_network_layer_info(self.h, i, what)

    def layer_shape(self, i):
        return (self.batch, self.layer_info(i, 1), self.layer_info(i, 2), self.layer_info(i, 3))

    def layer_type(self, i):
        return ["convolutional", "maxpool", "upsample", "route", "shortcut", "yolo", "Gaussian_yolo"][self.layer_info(i, 0)]

    def has_output(self, i):
        return bool(self.layer_info(i, 4))

    def flops_per_image(self):
        return L.lib().yc_network_flops(self.h)

    def kernels_per_forward(self):
        return self.info(6)

    def layer_output(self, i):
        """fp32 NCHW copy of layer i's output (cuda_pull_array of l->output_gpu)."""
        out = np.empty(self.layer_shape(i), np.float32)
        L.check(L.lib().yc_network_layer_output(self.h, i, out.ctypes.data_as(C.c_void_p)), "yc_network_layer_output")
        return out

    def yolo_layers(self):
        return [i for i in range(self.n) if self.layer_info(i, 0) in (5, 6)]

    def close(self):
        if self.h:
            L.lib().yc_network_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def LoadNetwork(cfg, weights=None, batch=0, flags=L.NET_DEFAULT, cfg_is_text=False, finalize=True) -> Network:
    """LoadNetwork (parser.cpp:1852-1875): parse cfg, load .weights, fuse BN, plan kernels.
    `batch<=0` keeps the cfg's batch; the reference forces 1 for inference (parser.cpp:1114-1115)."""
    h = C.c_void_p()
    fn = L.lib().yc_network_load_text if cfg_is_text else L.lib().yc_network_load
    L.check(fn(C.byref(h), cfg.encode(), weights.encode() if weights else None, batch, flags), "LoadNetwork")
    net = Network(h)
    if not weights and finalize:
        L.check(L.lib().yc_network_finalize(h), "yc_network_finalize")
    return net


def SetConvWeights(net: Network, layer: int, weights, biases, scales=None, rolling_mean=None, rolling_variance=None):
    """Overwrite one conv layer's parameters from numpy (reference layouts: OIHW weights); call Finalize after."""
    def p(a):
        if a is None:
            return None
        a = np.ascontiguousarray(a, np.float32)
        keep.append(a)
        return a.ctypes.data_as(C.c_void_p)
    keep = []
    L.check(L.lib().yc_network_set_conv_weights(net.h, layer, p(weights), p(biases), p(scales), p(rolling_mean),
                                                p(rolling_variance)), "yc_network_set_conv_weights")


def Finalize(net: Network):
    L.check(L.lib().yc_network_finalize(net.h), "yc_network_finalize")


def ResizeNetwork(net: Network, w: int, h: int):
    """ResizeNetwork (network.cpp:255-383): new input size, parameters / momentum / counters kept."""
    L.check(L.lib().yc_network_resize(net.h, w, h), "ResizeNetwork")
    net.w = L.lib().yc_network_info(net.h, 2)
    net.hh = L.lib().yc_network_info(net.h, 3)


def LoadSynthetic(cfg_text, weights: dict, batch=0, flags=L.NET_DEFAULT) -> Network:
    """LoadNetwork from a cfg string and a {layer: {weights,biases,scales,rolling_mean,rolling_variance}} dict
    (darknet_b200.cfg.synth_weights) without going through a .weights file."""
    net = LoadNetwork(cfg_text, None, batch, flags, cfg_is_text=True, finalize=False)
    for i, d in weights.items():
        SetConvWeights(net, i, d["weights"], d["biases"], d.get("scales"), d.get("rolling_mean"), d.get("rolling_variance"))
    Finalize(net)
    return net


def InputBuffer(net: Network):
    """numpy view of the library's pinned input staging buffer [batch,c,h,w] (net->input_pinned_cpu)."""
    p = L.lib().yc_network_input_buffer(net.h)
    if not p:
        raise L.YcError("yc_network_input_buffer: " + L.lib().yc_last_error().decode())
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(net.batch, net.c, net.hh, net.w))


def FreeNetwork(net: Network):
    net.close()


def InputBufferSlot(net: Network, slot: int) -> np.ndarray:
    """One of the two pinned fp32 NCHW staging buffers (slot 0 == InputBuffer) to alternate between when prefetching."""
    p = L.lib().yc_network_input_buffer_slot(net.h, slot)
    if not p:
        raise L.YcError("yc_network_input_buffer_slot: " + L.lib().yc_last_error().decode())
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=(net.batch, net.c, net.hh, net.w))


def Prefetch(net: Network, x_next: np.ndarray):
    """Start the H2D copy of the NEXT batch while the current one computes; NetworkPredict(net, x_next) picks it up."""
    assert x_next.dtype == np.float32 and x_next.flags["C_CONTIGUOUS"]
    L.check(L.lib().yc_network_prefetch(net.h, x_next.ctypes.data_as(C.c_void_p)), "yc_network_prefetch")


def InputBufferU8(net: Network) -> np.ndarray:
    """The pinned u8 HWC staging buffer [batch,h,w,c] of NetworkPredictU8 (zero-copy producers write here)."""
    p = L.lib().yc_network_input_buffer_u8(net.h)
    if not p:
        raise L.YcError("yc_network_input_buffer_u8: " + L.lib().yc_last_error().decode())
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_ubyte)), shape=(net.batch, net.hh, net.w, net.c))


def NetworkPredictU8(net: Network, img_u8: np.ndarray, swap_rb=False):
    """ProcImage's Mat2Image + NetworkPredict (yolo_core.cpp:104-112): img_u8 = host u8 HWC [batch,h,w,c] at the network's
    resolution; swap_rb folds cvtColor(RGB2BGR).  Bit-identical to NetworkPredict(Mat2Image(img))."""
    if not (img_u8.dtype == np.uint8 and img_u8.flags["C_CONTIGUOUS"]):
        img_u8 = np.ascontiguousarray(img_u8, np.uint8)
    assert img_u8.size == net.batch * net.c * net.hh * net.w, "input size != batch*h*w*c"
    p = L.lib().yc_network_predict_u8(net.h, img_u8.ctypes.data_as(C.c_void_p), 1 if swap_rb else 0)
    if not p:
        raise L.YcError("NetworkPredictU8 failed: " + L.lib().yc_last_error().decode())
    last = net.n - 1
    if net.layer_info(last, 0) in (5, 6):
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=net.layer_shape(last))
    return None


def NetworkPredictFrames(net: Network, frames_u8: np.ndarray, swap_rb=True):
    """ProcImage from the frame on (yolo_core.cpp:104-112): frames_u8 = host u8 HWC [batch,src_h,src_w,c] of ANY size;
    cv::resize to the network's size, cvtColor(RGB2BGR) (swap_rb), Mat2Image and NetworkPredict, the first three in one
    device kernel.  Bit-identical to NetworkPredictU8 on cv::resize'd frames."""
    if not (frames_u8.dtype == np.uint8 and frames_u8.flags["C_CONTIGUOUS"]):
        frames_u8 = np.ascontiguousarray(frames_u8, np.uint8)
    assert frames_u8.ndim == 4 and frames_u8.shape[0] == net.batch and frames_u8.shape[3] == net.c, "frames must be [batch,h,w,c]"
    p = L.lib().yc_network_predict_u8_resized(net.h, frames_u8.ctypes.data_as(C.c_void_p), frames_u8.shape[1],
                                              frames_u8.shape[2], 1 if swap_rb else 0)
    if not p:
        raise L.YcError("NetworkPredictFrames failed: " + L.lib().yc_last_error().decode())
    last = net.n - 1
    if net.layer_info(last, 0) in (5, 6):
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=net.layer_shape(last))
    return None


def NetworkPredict(net: Network, x: np.ndarray):
    """NetworkPredict (network.cpp:412-430): x = host fp32 NCHW [batch,c,h,w]; synchronous.  Returns a numpy
    view of the last layer's host output (layer-owned, valid until the next call)."""
    if not (x.dtype == np.float32 and x.flags["C_CONTIGUOUS"]):
        x = np.ascontiguousarray(x, np.float32)
    assert x.size == net.batch * net.c * net.hh * net.w, "input size != batch*c*h*w"
    p = L.lib().yc_network_predict(net.h, x.ctypes.data_as(C.c_void_p))
    if not p:
        raise L.YcError("NetworkPredict failed: " + L.lib().yc_last_error().decode())
    last = net.n - 1
    shape = net.layer_shape(last)
    if net.layer_info(last, 0) in (5, 6):
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=shape)
    return None


def ForwardNetworkGpu(net: Network, input_dev_ptr: int, stream=None):
    """ForwardNetworkGpu (network_kernels.cu:45-114) on a device-resident fp32 NCHW input; asynchronous."""
    L.check(L.lib().yc_network_forward(net.h, C.c_void_p(input_dev_ptr), stream if stream is not None else L.stream_ptr()),
            "ForwardNetworkGpu")


def yolo_output(net: Network, i):
    """l->output of YOLO layer i after NetworkPredict (pinned host memory), as [batch, c, h, w]."""
    p = L.lib().yc_network_yolo_output(net.h, i)
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_float)), shape=net.layer_shape(i))


class Detections:
    """Owning wrapper over the Detection array returned by GetNetworkBoxes (freed with FreeDetections)."""

    def __init__(self, ptr, n, classes):
        self.ptr, self.n, self.classes = ptr, n, classes

    def boxes(self):
        """[n, 5]: x y w h objectness"""
        return np.array([[d.x, d.y, d.w, d.h, d.objectness] for d in (self.ptr[i] for i in range(self.n))],
                        np.float32).reshape(self.n, 5)

    def probs(self):
        out = np.zeros((self.n, self.classes), np.float32)
        for i in range(self.n):
            out[i] = np.ctypeslib.as_array(self.ptr[i].prob, shape=(self.classes,))
        return out

    def free(self):
        if self.ptr is not None:
            L.lib().yc_free_detections(self.ptr, self.n)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def DetectionsFromArrays(boxes4, obj, probs):
    """Build a Detection array from numpy (tests / callers with their own candidates).  Returns (array, keepalive)."""
    boxes4 = np.ascontiguousarray(boxes4, np.float32)
    probs = np.ascontiguousarray(probs, np.float32).copy()
    n, classes = probs.shape
    arr = (L.Detection * max(n, 1))()
    for i in range(n):
        arr[i].x, arr[i].y, arr[i].w, arr[i].h = (float(v) for v in boxes4[i])
        arr[i].objectness = float(obj[i])
        arr[i].classes = classes
        arr[i].prob = C.cast(probs[i].ctypes.data_as(C.c_void_p), C.POINTER(C.c_float))
        arr[i].sort_class = -1
    return arr, probs


def NmsSortArrays(boxes4, obj, probs, thresh, nms_kind=0, beta=0.6):
    """NmsSort on arrays: (probs after suppression in original order, final array order)."""
    arr, p = DetectionsFromArrays(boxes4, obj, probs)
    n, classes = p.shape
    L.lib().yc_nms_sort(arr, n, classes, thresh, nms_kind, beta)
    base = p.ctypes.data
    order = np.array([(C.cast(arr[i].prob, C.c_void_p).value - base) // (4 * classes) for i in range(n)], np.int32)
    return p, order


def GetNetworkBoxes(net: Network, thresh: float, b: int = 0) -> Detections:
    """GetNetworkBoxes (network.cpp:498-503) for batch element b (the reference reads b=0 only)."""
    n = C.c_int(0)
    ptr = L.lib().yc_network_get_boxes(net.h, b, thresh, C.byref(n))
    if not ptr:
        raise L.YcError("GetNetworkBoxes failed: " + L.lib().yc_last_error().decode())
    yl = net.yolo_layers()
    classes = net.layer_info(yl[-1], 1) // 3 - 5 if yl else 0
    return Detections(ptr, n.value, _classes(net))


def _classes(net):
    yl = net.yolo_layers()
    i = yl[-1]
    per = net.layer_info(i, 1)
    gaussian = net.layer_info(i, 0) == 6
    return per // 3 - (9 if gaussian else 5)  # heads on the path use 3 anchors per layer


def NmsSort(dets: Detections, thresh: float, nms_kind: int = 0, beta: float = 0.6):
    """NmsSort (box.cpp:393-419), in place."""
    L.lib().yc_nms_sort(dets.ptr, dets.n, dets.classes, thresh, nms_kind, beta)


def DetectBatch(net: Network, thresh: float, nms_thresh: float = 0.45):
    """Throughput path: on-device threshold + box decode for every image, host NmsSort per image.
    Returns (rows [total, 7+classes], counts [batch]); row = image, x, y, w, h, objectness, head_layer, probs."""
    rows = C.POINTER(C.c_float)()
    counts = np.zeros(net.batch, np.int32)
    total = C.c_int(0)
    L.check(L.lib().yc_network_detect(net.h, thresh, nms_thresh, C.byref(rows), counts.ctypes.data_as(C.c_void_p),
                                      C.byref(total)), "yc_network_detect")
    R = 7 + _classes(net)
    if total.value == 0:
        return np.zeros((0, R), np.float32), counts
    return np.ctypeslib.as_array(rows, shape=(total.value, R)).copy(), counts


# ---- training: TrainNetworkDatum / UpdateNetwork mirrors (network.cpp:116-208) -------------------------------------
def TrainNetworkDatum(net: Network, x: np.ndarray, truth: np.ndarray) -> float:
    """TrainNetworkDatum (network.cpp:188-208): forward + backward on host x [batch,c,h,w] and truth
    [batch, max_boxes*5]; returns GetNetworkCost.  The network must be loaded with NET_TRAIN."""
    x = np.ascontiguousarray(x, np.float32)
    truth = np.ascontiguousarray(truth, np.float32)
    tf = L.lib().yc_network_train_info(net.h, 0)
    assert x.size == net.batch * net.c * net.hh * net.w, "input size != batch*c*h*w"
    assert truth.size == net.batch * tf, f"truth size != batch*{tf}"
    cost = C.c_float(0)
    L.check(L.lib().yc_network_train_datum(net.h, x.ctypes.data_as(C.c_void_p), truth.ctypes.data_as(C.c_void_p),
                                           C.byref(cost)), "TrainNetworkDatum")
    return float(cost.value)


def ForwardBackwardNetworkGpu(net: Network, input_dev_ptr: int, truth_dev_ptr: int, stream=None):
    """ForwardBackwardNetworkGpu (network_kernels.cu:245-267) on device-resident input/truth; asynchronous."""
    L.check(L.lib().yc_network_forward_backward(net.h, C.c_void_p(input_dev_ptr), C.c_void_p(truth_dev_ptr),
                                                stream if stream is not None else L.stream_ptr()),
            "ForwardBackwardNetworkGpu")


def UpdateNetwork(net: Network, lr=-1.0, momentum=-1.0, decay=-1.0, stream=None):
    """UpdateNetwork (network.cpp:116-128): SGD on every conv layer; negative values take the cfg's [net] entries."""
    L.check(L.lib().yc_network_update(net.h, lr, momentum, decay, stream), "UpdateNetwork")


def layer_delta(net: Network, i):
    """fp32 NCHW copy of l->delta of layer i."""
    out = np.empty(net.layer_shape(i), np.float32)
    L.check(L.lib().yc_network_layer_delta(net.h, i, out.ctypes.data_as(C.c_void_p)), "yc_network_layer_delta")
    return out


def layer_x(net: Network, i):
    """fp32 NCHW copy of l->x (the convolution output before batchnorm) of a batch-normalised conv layer i."""
    out = np.empty(net.layer_shape(i), np.float32)
    L.check(L.lib().yc_network_layer_x(net.h, i, out.ctypes.data_as(C.c_void_p)), "yc_network_layer_x")
    return out


def layer_param(net: Network, i, key, capacity=None):
    """l-><key> of layer i (weights, biases, scales, rolling_mean, ..., weight_updates, cost) in reference layout."""
    if capacity is None:
        n, cin, k = net.layer_info(i, 1), net.layer_info(i, 5), net.layer_info(i, 6)
        capacity = max(n * cin * k * k, 8)
    out = np.empty(capacity, np.float32)
    r = L.lib().yc_network_get_param(net.h, i, key.encode(), out.ctypes.data_as(C.c_void_p), capacity)
    if r < 0:
        raise L.YcError(f"yc_network_get_param({i}, {key}): " + L.lib().yc_last_error().decode())
    return out[:r].copy()

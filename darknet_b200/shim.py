"""ctypes window onto libyolo_core_b200.so (darknet_b200/shim/yolo_core_b200.cpp): the reference's own entry points
with the reference's own signatures.  Tests use it to replay the reference's call sequence (ProcImage,
src/yolo_core.cpp:104-137) and to install the per-layer trampolines into `layer` structs the reference itself parsed."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "libyolo_core_b200.so")

# every symbol a yolo-core build would bind (yolo_core.h:625-663, network.h:8-18, box.h:93-94, layer function pointers)
REFERENCE_SYMBOLS = ["LoadNetwork", "FreeNetwork", "NetworkPredict", "NetworkPredictGpu", "ForwardNetworkGpu",
                     "GetNetworkBoxes", "FreeDetections", "FuseConvBatchNorm", "_Z7NmsSortP9Detectioniif8NMS_KINDf",
                     "_Z13ResizeNetworkP7Networkii",
                     "ForwardConvolutionalLayerGpu", "ForwardMaxpoolLayerGpu", "ForwardUpsampleLayerGpu",
                     "ForwardShortcutLayerGpu", "ForwardRouteLayerGpu", "ForwardYoloLayerGpu",
                     "ForwardGaussianYoloLayerGpu", "CudaGetDeviceCount", "cuda_set_device"]


class RefDetection(C.Structure):
    """struct Detection (box.h:69-86): Box bbox; int classes; float* prob; float* mask; float objectness; int sort_class;
    float* uc; int points."""
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("w", C.c_float), ("h", C.c_float), ("classes", C.c_int),
                ("prob", C.POINTER(C.c_float)), ("mask", C.POINTER(C.c_float)), ("objectness", C.c_float),
                ("sort_class", C.c_int), ("uc", C.POINTER(C.c_float)), ("points", C.c_int)]


_lib = None


def have():
    return os.path.exists(SO)


def lib():
    global _lib
    if _lib is None:
        if not have():
            raise RuntimeError(f"{SO} is missing: darknet_b200.build.build_shim() (needs the reference header tree)")
        L = C.CDLL(SO)
        L.LoadNetwork.restype = C.c_bool
        L.LoadNetwork.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_bool, C.c_bool]
        L.FreeNetwork.argtypes = [C.c_void_p]
        L.NetworkPredict.restype = C.POINTER(C.c_float)
        L.NetworkPredict.argtypes = [C.c_void_p, C.c_void_p]
        L.GetNetworkBoxes.restype = C.POINTER(RefDetection)
        L.GetNetworkBoxes.argtypes = [C.c_void_p, C.c_float, C.POINTER(C.c_int)]
        L.FreeDetections.argtypes = [C.POINTER(RefDetection), C.c_int]
        L.yc_shim_nms_sort.argtypes = [C.POINTER(RefDetection), C.c_int, C.c_int, C.c_float, C.c_int, C.c_float]
        L.yc_shim_resize_network.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.yc_shim_sizeof.argtypes = [C.c_int]
        L.yc_shim_install.argtypes = [C.c_void_p]
        L.yc_shim_release.argtypes = [C.c_void_p]
        L.yc_shim_call_forward_gpu.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.yc_shim_layer_output_gpu.restype = C.c_void_p
        L.yc_shim_layer_output_gpu.argtypes = [C.c_void_p]
        _lib = L
    return _lib


class ShimNetwork:
    """A reference `Network` struct filled by the shim's LoadNetwork."""

    def __init__(self, cfg_path, weights_path, train=False):
        L = lib()
        self.buf = C.create_string_buffer(L.yc_shim_sizeof(1))
        self.ptr = C.cast(self.buf, C.c_void_p)
        ok = L.LoadNetwork(self.ptr, cfg_path.encode(), weights_path.encode() if weights_path else None, train, False)
        if not ok:
            raise RuntimeError("LoadNetwork failed")

    def predict(self, x: np.ndarray):
        x = np.ascontiguousarray(x, np.float32)
        p = lib().NetworkPredict(self.ptr, x.ctypes.data_as(C.c_void_p))
        if not p:
            raise RuntimeError("NetworkPredict failed")
        return p

    def boxes(self, thresh, nms_thresh, classes, nms_kind=0, beta=0.6):
        """GetNetworkBoxes + NmsSort + FreeDetections, the ProcImage sequence: (rows [n,5] x y w h objectness, probs)."""
        L = lib()
        n = C.c_int(0)
        d = L.GetNetworkBoxes(self.ptr, thresh, C.byref(n))
        if nms_thresh > 0:
            L.yc_shim_nms_sort(d, n.value, classes, nms_thresh, nms_kind, beta)
        rows = np.array([[d[i].x, d[i].y, d[i].w, d[i].h, d[i].objectness] for i in range(n.value)], np.float32).reshape(-1, 5)
        probs = np.zeros((n.value, classes), np.float32)
        for i in range(n.value):
            probs[i] = np.ctypeslib.as_array(d[i].prob, shape=(classes,))
        L.FreeDetections(d, n.value)
        return rows, probs

    def resize(self, w, h):
        lib().yc_shim_resize_network(self.ptr, w, h)

    def close(self):
        if self.ptr:
            lib().FreeNetwork(self.ptr)
            self.ptr = None

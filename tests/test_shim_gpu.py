"""The reference's call sequences against libyolo_core_b200.so (the compiled drop-in boundary).

1. ProcImage (src/yolo_core.cpp:104-137): LoadNetwork -> NetworkPredict -> GetNetworkBoxes -> NmsSort -> FreeDetections
   with the reference's signatures and structs; detections == the library's own API bit for bit, and == the compiled
   reference inside the bf16 margin policy.
2. The operator seam (yolo_core.h:157-159): `layer` structs PARSED BY THE REFERENCE ITSELF (oracle/_ref, CPU build) get
   the shim's trampolines installed into l->forward_gpu and are called the way ForwardNetworkGpu calls them
   (l->forward_gpu(l, state), state.input = the previous layer's device output, fp32 NCHW); every layer's output is
   compared with the reference's CPU output of the same network."""
import ctypes as C

import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod, shim
from oracle import pyoracle
from tests.netutil import materialise, rel_l2, rel_to_max

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not shim.have(), reason="libyolo_core_b200.so not built")]


def test_procimage_sequence(gpu):
    from darknet_b200 import lib as L, network as N
    from tests.test_network_gpu import _flatten, _unmatched
    text, cpath, wpath = materialise("yolov4-tiny", head_bias=-3.0, head_gain=40.0)
    x = cfgmod.synth_images(1, 3, 416, 416)
    net = shim.ShimNetwork(cpath, wpath)
    net.predict(x)
    classes = 80
    for thresh in (0.25, 0.5):
        rows, probs = net.boxes(thresh, 0.45, classes)
        # the same through the library's own API: identical, array order included
        own = N.LoadNetwork(cpath, wpath, batch=1)
        N.NetworkPredict(own, x)
        d = N.GetNetworkBoxes(own, thresh, 0)
        N.NmsSort(d, 0.45)
        assert np.array_equal(rows, d.boxes()) and np.array_equal(probs, d.probs())
        own.close()
        # and against the compiled reference running the same sequence
        ref = pyoracle.RefNet(cpath, wpath)
        ref.predict(x)
        r7, rprobs = ref.boxes(thresh, 0.0)
        pre_rows, pre_probs = net.boxes(thresh, 0.0, classes)
        R, G = _flatten(r7[:, :5], rprobs), _flatten(pre_rows, pre_probs)
        assert R.shape[0] > 0
        bad = _unmatched(R, G, thresh, 0.03) + _unmatched(G, R, thresh, 0.03)
        assert bad <= 0.01 * R.shape[0], (bad, R.shape[0])
        ref.close()
    net.close()


@pytest.mark.parametrize("fuse", [True, False])
def test_trampolines_in_reference_parsed_layers(gpu, fuse):
    from darknet_b200 import lib as L
    lib, S = L.lib(), shim.lib()
    text, cpath, wpath = materialise("micro")
    x = cfgmod.synth_images(1, 3, 64, 64, seed=11)
    ref = pyoracle.RefNet(cpath, wpath, train=False, fuse=fuse)
    ref.predict(x)  # the reference's CPU outputs, layer by layer (copied before the trampolines touch l->output)
    wants = [ref.layer_output(i) for i in range(ref.n)]
    S.yc_shim_set_ref_cpu_maxpool(1)  # compare against the CPU inference path, border artefact of its maxpool included
    R = pyoracle.ref_lib()
    netp = R.ref_network_ptr(ref.h)
    for i in range(ref.n):
        assert S.yc_shim_install(R.ref_layer_ptr(ref.h, i)) == 1, f"layer {i} is not on the hot path"
    xin = lib.yc_malloc(x.nbytes)
    L.check(lib.yc_push(xin, x.ctypes.data_as(C.c_void_p), x.nbytes, None), "yc_push")
    L.check(lib.yc_stream_sync(None), "sync")
    cur, worst = xin, 0.0
    for i in range(ref.n):
        S.yc_shim_call_forward_gpu(netp, i, cur, 0)       # l->forward_gpu(l, state) through the installed pointer
        cur = S.yc_shim_layer_output_gpu(R.ref_layer_ptr(ref.h, i))
        want = wants[i]
        got = np.empty_like(want)
        L.check(lib.yc_pull(got.ctypes.data_as(C.c_void_p), cur, got.nbytes, None), "yc_pull")
        L.check(lib.yc_stream_sync(None), "sync")
        e2, em = rel_l2(got, want), rel_to_max(got, want)
        worst = max(worst, e2)
        assert e2 < 1e-2 and em < 2e-2, f"layer {i}: rel-L2 {e2:.3e} rel-max {em:.3e}"
    print("trampolines, fuse", fuse, "worst rel-L2", worst)
    S.yc_shim_set_ref_cpu_maxpool(0)
    for i in range(ref.n):
        S.yc_shim_release(R.ref_layer_ptr(ref.h, i))
    lib.yc_free(xin)
    ref.close()

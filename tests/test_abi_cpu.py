"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol include/yc_b200.h
declares, host logic (cfg parsing, error reporting) behaves like the reference, and the product refuses to run
without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C

import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod
from darknet_b200 import cfgs


def test_library_exports_every_declared_symbol(built):
    from darknet_b200 import lib as L
    lib = L.lib()
    names = L.declared_symbols()
    assert len(names) >= 45
    missing = [s for s in names if not hasattr(lib, s)]
    assert not missing, missing
    unbound = [s for s in names if s not in L._SIGS]
    assert not unbound, f"declared in the header but not bound in lib.py: {unbound}"


def test_no_cpu_fallback(built):
    """Without a CUDA device every compute entry point fails loudly with YC_ERR_NO_DEVICE / an error string."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from darknet_b200 import lib as L
    lib = L.lib()
    assert lib.yc_device_count() == 0
    assert lib.yc_set_device(0) == -3  # YC_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.yc_last_error()
    h = C.c_void_p()
    rc = lib.yc_network_load_text(C.byref(h), cfgs.micro().encode(), None, 1, L.NET_DEFAULT)
    assert rc == 0  # parsing is host logic and works
    assert lib.yc_network_finalize(h) == -3
    assert lib.yc_network_predict(h, np.zeros(3 * 64 * 64, np.float32).ctypes.data_as(C.c_void_p)) is None
    lib.yc_network_free(h)


def test_cfg_parse_errors_like_reference(built):
    from darknet_b200 import lib as L
    lib = L.lib()
    h = C.c_void_p()
    # first section must be [net] (parser.cpp:1091-1092)
    assert lib.yc_network_load_text(C.byref(h), b"[convolutional]\nfilters=1\n", None, 1, 0) == -6
    assert b"First section must be [net]" in lib.yc_last_error()
    # yolo filters mismatch (parser.cpp:323-330)
    bad = cfgs.micro().replace("filters=24", "filters=25")
    assert lib.yc_network_load_text(C.byref(h), bad.encode(), None, 1, 0) == -2
    assert b"does not correspond" in lib.yc_last_error()
    # layer kinds outside the hot path are refused, not silently skipped
    other = "[net]\nwidth=8\nheight=8\nchannels=3\n[convolutional]\nfilters=8\nsize=1\n[connected]\noutput=4\n"
    assert lib.yc_network_load_text(C.byref(h), other.encode(), None, 1, 0) == -2
    assert b"outside the conv/YOLO hot path" in lib.yc_last_error()
    # missing file
    assert lib.yc_network_load(C.byref(h), b"/nonexistent.cfg", None, 1, 0) == -5


def test_network_shapes_match_python_walker(built):
    """C++ parser (csrc/network.cu) and the Python shape walker agree layer by layer on all five configs."""
    from darknet_b200 import lib as L
    lib = L.lib()
    for name, fn in cfgs.CONFIGS.items():
        text = fn()
        _, layers = cfgmod.walk_shapes(text)
        h = C.c_void_p()
        assert lib.yc_network_load_text(C.byref(h), text.encode(), None, 0, L.NET_DEFAULT) == 0, lib.yc_last_error()
        assert lib.yc_network_info(h, 0) == len(layers)
        for Ls in layers:
            got = tuple(lib.yc_network_layer_info(h, Ls.index, w) for w in (1, 2, 3))
            assert got == (Ls.out_c, Ls.out_h, Ls.out_w), (name, Ls.index, got)
        flops = lib.yc_network_flops(h)
        want = sum(L_.bflops for L_ in layers if L_.kind == "convolutional") * 1e9
        assert abs(flops - want) / want < 1e-9
        lib.yc_network_free(h)


def test_bflops_match_upstream_totals():
    """Conv BFLOPs of the regenerated cfgs == the reference parser's printed totals minus the maxpool/shortcut
    pseudo-BFLOPs: 6.905 / 65.862 / 128.39."""
    for name, want in (("yolov4-tiny", 6.905), ("yolov3", 65.862), ("yolov4", 128.39)):
        _, layers = cfgmod.walk_shapes(cfgs.CONFIGS[name]())
        got = sum(L.bflops for L in layers if L.kind == "convolutional")
        assert abs(got - want) < 0.01, (name, got)


def test_weights_file_roundtrip(tmp_path):
    text = cfgs.micro()
    w = cfgmod.synth_weights(text)
    p = str(tmp_path / "m.weights")
    cfgmod.write_weights(p, text, w, seen=1234)
    r, seen = cfgmod.read_weights(p, text)
    assert seen == 1234
    for i, d in w.items():
        for k, v in d.items():
            assert np.array_equal(r[i][k], v)

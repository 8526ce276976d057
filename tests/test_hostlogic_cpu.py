"""Host-side logic of the C++ runtime that needs no GPU, checked against the compiled reference where it has a twin:
mini-batch = batch / subdivisions (ParseNetOptions), the learning-rate schedule (GetCurrLr), the darknet .weights writer
(SaveWeights) and reader incl. backbone files that end early (LoadWeightsUpTo), the initialisation law (FillConvLayer)."""
import ctypes as C
import os
import struct
import tempfile

import numpy as np
import pytest

from darknet_b200 import cfg as cfgmod, cfgs, lib as L
from oracle import pyoracle

needs_ref = pytest.mark.skipif(not pyoracle.have_ref(), reason="oracle/_ref not built")
TMP = tempfile.mkdtemp(prefix="yc_host_")


def _load_text(text, batch=0, flags=L.NET_DEFAULT):
    h = C.c_void_p()
    L.check(L.lib().yc_network_load_text(C.byref(h), text.encode(), None, batch, flags), "load_text")
    return h


def _write(name, text):
    p = os.path.join(TMP, name)
    with open(p, "w") as f:
        f.write(text)
    return p


def test_minibatch_is_batch_over_subdivisions(built):
    lib = L.lib()
    text = cfgs.micro(batch=64, subdivisions=16)
    h = _load_text(text)                       # batch <= 0: keep the cfg's -> 64 / 16 (parser.cpp:926-928)
    assert lib.yc_network_info(h, 1) == 4 and lib.yc_network_info(h, 8) == 16
    lib.yc_network_free(h)
    h = _load_text(text, batch=8)              # an explicit batch is the mini-batch itself
    assert lib.yc_network_info(h, 1) == 8 and lib.yc_network_info(h, 8) == 16
    lib.yc_network_free(h)
    if pyoracle.have_ref():
        ref = pyoracle.RefNet(_write("mb.cfg", text), None, train=True)
        assert ref.batch == 4
        ref.close()


POLICIES = {
    "constant": "",
    "burn": "burn_in=100\npower=4\n",
    "step": "policy=step\nstep=300\nscale=0.5\n",
    "steps": "policy=steps\nsteps=0.4,0.8,0.9\nscales=.1,.1,.5\nburn_in=50\n",
    "exp": "policy=exp\ngamma=0.999\n",
    "poly": "policy=poly\npower=2\n",
    "sigmoid": "policy=sigmoid\ngamma=0.01\nstep=500\n",
    "sgdr": "policy=sgdr\nsgdr_cycle=200\nsgdr_mult=2\nlearning_rate_min=0.0001\nsteps=1\nscales=1\n",  # (the reference parser dereferences steps= for sgdr too)
}


@needs_ref
@pytest.mark.parametrize("name", sorted(POLICIES))
def test_current_rate_equals_GetCurrLr(built, name):
    lib = L.lib()
    text = cfgs.micro(batch=4, subdivisions=1).replace("[net]\n", "[net]\nlearning_rate=0.00261\n" + POLICIES[name], 1)
    h = _load_text(text)
    ref = pyoracle.RefNet(_write(f"lr_{name}.cfg", text), None, train=True)
    R = pyoracle.ref_lib()
    for max_iter in (1000, 500500):
        for it in [0, 1, 7, 49, 50, 99, 100, 199, 200, 201, 399, 400, 401, 599, 600, 799, 800, 899, 900, 999, 1000, 200199,
                   400400, 450450, 500499]:
            want = R.ref_curr_lr(ref.h, it, max_iter)
            got = lib.yc_network_current_rate(h, it, max_iter)
            assert got == pytest.approx(want, rel=2e-6, abs=1e-12), (name, it, max_iter, got, want)
    ref.close()
    lib.yc_network_free(h)


def _set_weights(h, W):
    lib = L.lib()
    keep = []
    for i, d in W.items():
        ptrs = []
        for k in ("weights", "biases", "scales", "rolling_mean", "rolling_variance"):
            a = d.get(k)
            if a is None:
                ptrs.append(None)
            else:
                a = np.ascontiguousarray(a, np.float32)
                keep.append(a)
                ptrs.append(a.ctypes.data_as(C.c_void_p))
        L.check(lib.yc_network_set_conv_weights(h, i, *ptrs), "set_conv_weights")


def test_save_weights_format_and_roundtrip(built):
    """yc_network_save_weights writes the darknet layout byte for byte: equal to the Python writer the tests have used all
    along, and the reference's own LoadWeights reads it back."""
    lib = L.lib()
    text = cfgs.micro()
    W = cfgmod.synth_weights(text, seed=9)
    h = _load_text(text)
    _set_weights(h, W)
    path = os.path.join(TMP, "saved.weights")
    L.check(lib.yc_network_save_weights(h, path.encode(), 0), "save_weights")
    want = os.path.join(TMP, "py.weights")
    cfgmod.write_weights(want, text, W)
    assert open(path, "rb").read() == open(want, "rb").read()
    with open(path, "rb") as f:
        assert struct.unpack("<iii", f.read(12)) == (0, 2, 5)
    if pyoracle.have_ref():
        ref = pyoracle.RefNet(_write("rt.cfg", text), path, train=False, fuse=False)
        for i, d in W.items():
            assert np.array_equal(ref.layer_buf(i, "weights"), d["weights"])
            assert np.array_equal(ref.layer_buf(i, "biases"), d["biases"])
            if "scales" in d:
                assert np.array_equal(ref.layer_buf(i, "rolling_variance"), d["rolling_variance"])
        ref.close()
    # cutoff = SaveWeightsUpto: a shorter file that loads as a backbone
    cut = os.path.join(TMP, "cut.weights")
    L.check(lib.yc_network_save_weights(h, cut.encode(), 7), "save_weights cutoff")
    assert os.path.getsize(cut) < os.path.getsize(path)
    h2 = _load_text(text)
    L.check(lib.yc_network_load_weights(h2, cut.encode()), "load backbone")
    assert lib.yc_network_info(h2, 11) == 7   # layers 0..6 filled (the last conv below the cutoff is layer 6)
    resaved = os.path.join(TMP, "resaved.weights")
    L.check(lib.yc_network_save_weights(h2, resaved.encode(), 7), "save")
    assert open(resaved, "rb").read() == open(cut, "rb").read()
    # a file that ends inside a layer is an error
    bad = os.path.join(TMP, "bad.weights")
    with open(bad, "wb") as f:
        f.write(open(path, "rb").read()[:-10])
    assert lib.yc_network_load_weights(h2, bad.encode()) == -5 and b"truncated" in lib.yc_last_error()
    lib.yc_network_free(h)
    lib.yc_network_free(h2)


def test_seen_travels_through_the_weights_header(built):
    lib = L.lib()
    text = cfgs.micro(batch=8, subdivisions=2)
    W = cfgmod.synth_weights(text)
    p = os.path.join(TMP, "seen.weights")
    cfgmod.write_weights(p, text, W, seen=6400)
    h = _load_text(text)
    L.check(lib.yc_network_load_weights(h, p.encode()), "load")
    assert lib.yc_network_info(h, 10) == 6400 and lib.yc_network_info(h, 9) == 6400 // 8
    out = os.path.join(TMP, "seen2.weights")
    L.check(lib.yc_network_save_weights(h, out.encode(), 0), "save")
    assert cfgmod.read_weights(out, text)[1] == 6400
    lib.yc_network_free(h)


@needs_ref
def test_init_weights_is_the_reference_law(built):
    """srand(seed) + the reference's parse == yc_network_init_weights(seed) for every conv before the first [yolo]
    section (FillYoloLayer re-seeds from the clock), bit for bit; later layers follow the same law (range, scale)."""
    lib = L.lib()
    libc = C.CDLL(None)
    text = cfgs.micro()
    cpath = _write("init.cfg", text)
    libc.srand(4242)
    ref = pyoracle.RefNet(cpath, None, train=False, fuse=False)
    h = _load_text(text)
    L.check(lib.yc_network_init_weights(h, 4242, 0), "init_weights")
    out = os.path.join(TMP, "init.weights")
    L.check(lib.yc_network_save_weights(h, out.encode(), 0), "save")
    mine, _ = cfgmod.read_weights(out, text)
    _, layers = cfgmod.walk_shapes(text)
    first_yolo = min(Ly.index for Ly in layers if Ly.kind == "yolo")
    checked = 0
    for Ly in layers:
        if Ly.kind != "convolutional":
            continue
        w = mine[Ly.index]["weights"]
        bound = np.sqrt(2.0 / (Ly.size * Ly.size * Ly.c))
        assert np.abs(w).max() <= bound * (1 + 1e-6) and w.std() == pytest.approx(bound / np.sqrt(3), rel=0.1)
        assert not mine[Ly.index]["biases"].any()
        if Ly.bn:
            assert np.all(mine[Ly.index]["scales"] == 1) and not mine[Ly.index]["rolling_mean"].any()
        if Ly.index < first_yolo:
            assert np.array_equal(w, ref.layer_buf(Ly.index, "weights")), Ly.index
            checked += 1
    assert checked >= 9
    ref.close()
    lib.yc_network_free(h)


@needs_ref
@pytest.mark.parametrize("gen,size0,sizes", [("yolov4_tiny", 416, [(320, 320), (608, 352)]), ("yolov4", 608, [(416, 416), (320, 512)]),
                                             ("gaussian_yolov3", 608, [(352, 352)])])
def test_resize_rederives_every_shape_like_ResizeNetwork(built, gen, size0, sizes):
    """yc_network_resize on a parsed (not finalized: no device needed) network against ResizeNetwork of the compiled
    reference (network.cpp:255-383): input / output shape of every layer, for square and non-square sizes, and back."""
    lib = L.lib()
    text = getattr(cfgs, gen)(w=size0, h=size0)
    h = _load_text(text)
    ref = pyoracle.RefNet(_write(f"rs_{gen}.cfg", text), None, train=True, fuse=False)
    n = lib.yc_network_info(h, 0)
    assert n == ref.n
    for (w, hh) in sizes + [(size0, size0)]:
        L.check(lib.yc_network_resize(h, w, hh), "resize")
        ref.resize(w, hh)
        assert (lib.yc_network_info(h, 2), lib.yc_network_info(h, 3)) == (ref.w, ref.hh) == (w, hh)
        for i in range(n):
            mine = tuple(lib.yc_network_layer_info(h, i, k) for k in (1, 2, 3))
            # (ResizeYoloLayer updates l->h / l->w and leaves out_h / out_w stale, yolo_layer.cpp:89-95)
            yolo = lib.yc_network_layer_info(h, i, 0) >= 5
            theirs = tuple(ref.layer_int(i, k) for k in (("out_c", "h", "w") if yolo else ("out_c", "out_h", "out_w")))
            assert mine == theirs, (gen, w, hh, i, mine, theirs)
    # a size at which a layer would be empty is refused and the old shape stays
    assert lib.yc_network_resize(h, 8, 8) != 0
    assert (lib.yc_network_info(h, 2), lib.yc_network_info(h, 3)) == (size0, size0)
    assert lib.yc_network_layer_info(h, n - 1, 2) == ref.layer_int(n - 1, "h")
    lib.yc_network_free(h)
    ref.close()

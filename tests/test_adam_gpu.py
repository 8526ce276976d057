"""The GPU-only parts of UpdateConvolutionalLayerGpu (convolutional_kernels.cu:865-922) that the reference's CPU twin does
not have: Adam (adam_update_gpu, blas_kernels.cu:99-134, [net] adam=1) and the per-step NaN / Inf scrub (:881-882,
blas_kernels.cu:1157-1201).  PARITY UNPINNED: the reference has neither a CPU implementation nor a test vector for them
(its CUDA build needs cuDNN / OpenCV and cannot be compiled here), so the checker below is a numpy restatement of the CUDA
source, op for op in fp32; tolerance 2e-6 relative (the reference's axpy kernels may or may not contract to FMA).
Also: the k-th update of a run sees iteration k (TrainNetwork increments curr_iter before UpdateNetworkGpu,
network.cpp:227-231) — visible through Adam's bias correction and through the burn-in learning rate."""
import numpy as np
import pytest
import torch

from darknet_b200 import cfg as cfgmod, cfgs

pytestmark = pytest.mark.gpu

B, HW, CLASSES = 4, 64, 3
f32 = np.float32


def _adam_ref(w, d, m, v, B1, B2, eps, decay, rate, batch, t):
    w, d, m, v = (np.array(a, f32) for a in (w, d, m, v))
    m = m * f32(B1)
    v = v * f32(B2)
    d = d + f32(-decay * batch) * w
    m = m + f32(1 - B1) * d
    d = d * d
    v = v + f32(1 - B2) * d
    mhat = m / f32(1.0 - np.power(f32(B1), f32(t)))
    vhat = v / f32(1.0 - np.power(f32(B2), f32(t)))
    w = w + f32(rate) * mhat / (np.sqrt(vhat) + f32(eps))
    return w.astype(f32), m.astype(f32), v.astype(f32)


def test_adam_three_steps(gpu):
    from darknet_b200 import lib as L, network as N
    text = cfgs.micro(batch=B, subdivisions=1, classes=CLASSES).replace("policy=constant", "policy=constant\nadam=1\nB1=0.9\nB2=0.999\neps=0.000001")
    weights = cfgmod.synth_weights(text)
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    convs = [i for i in range(net.n) if net.layer_info(i, 0) == 0]
    lr, decay = 0.001, 0.0005
    keys = lambda i: [("weights", "weight_updates"), ("biases", "bias_updates")] + ([("scales", "scale_updates")] if net.layer_info(i, 9) else [])
    mom = {}
    for t in (1, 2, 3):
        x = cfgmod.synth_images(B, 3, HW, HW, seed=30 + t)
        truth = cfgmod.synth_truth(B, 90, classes=CLASSES, seed=40 + t)
        N.TrainNetworkDatum(net, x, truth)
        before = {(i, p): (N.layer_param(net, i, p), N.layer_param(net, i, g)) for i in convs for p, g in keys(i)}
        N.UpdateNetwork(net)
        for (i, p), (w0, g0) in before.items():
            m0, v0 = mom.get((i, p), (np.zeros_like(w0), np.zeros_like(w0)))
            want, m1, v1 = _adam_ref(w0, g0, m0, v0, 0.9, 0.999, 1e-6, decay, lr, B, t)
            mom[(i, p)] = (m1, v1)
            got = N.layer_param(net, i, p)
            err = np.abs(got - want).max() / (np.abs(want).max() + 1e-30)
            assert err < 2e-6, f"step {t} layer {i} {p}: rel err {err:.2e}"
            assert np.abs(got - w0).max() > 0, f"step {t} layer {i} {p}: not updated"
        for i in convs:
            for _, g in keys(i):
                assert np.abs(N.layer_param(net, i, g)).max() == 0.0, "adam_update_gpu zeroes the gradient (fill_ongpu)"
    net.close()


def test_nan_inf_scrub_of_the_update(gpu):
    """yc_sgd_update_conv on arrays with NaN / Inf planted: gradient -> 0, weight -> 1 / (OIHW index + 1), then the SGD step"""
    from darknet_b200 import lib as L
    lib = L.lib()
    n, c, k = 8, 16, 3
    rng = np.random.default_rng(9)
    w = rng.standard_normal((n, c, k, k)).astype(f32)
    ld = lib.yc_conv_wgrad_ld(c)
    g = rng.standard_normal((n, k * k, ld)).astype(f32)
    g[:, :, c:] = 0
    w.reshape(-1)[[5, 700]] = [np.nan, -np.inf]
    g[1, 2, 3], g[7, 8, 15] = np.inf, np.nan
    lr, batch, momentum, decay = 0.01, 4, 0.9, 0.0005
    wd, gd = torch.from_numpy(w.copy()).to(gpu), torch.from_numpy(g.copy()).to(gpu)
    L.check(lib.yc_sgd_update_conv(wd.data_ptr(), gd.data_ptr(), n, c, k, f32(lr), batch, f32(momentum), f32(decay), None, None,
                                   None, None, None, None, L.stream_ptr()), "yc_sgd_update_conv")
    torch.cuda.synchronize()
    w0 = w.copy().reshape(-1)
    bad = ~np.isfinite(w0)
    w0[bad] = f32(1.0) / (np.abs(np.nonzero(bad)[0].astype(f32)) + f32(1))
    w0 = w0.reshape(w.shape)
    g0 = np.where(np.isfinite(g), g, f32(0))
    g_oihw = g0[:, :, :c].reshape(n, k, k, c).transpose(0, 3, 1, 2)       # [n][tap][c] -> OIHW
    gg = g_oihw + f32(-decay * batch) * w0
    want_w = w0 + f32(lr / batch) * gg
    want_g = gg * f32(momentum)
    got_w = wd.cpu().numpy()
    got_g = gd.cpu().numpy()[:, :, :c].reshape(n, k, k, c).transpose(0, 3, 1, 2)
    assert np.isfinite(got_w).all() and np.isfinite(got_g).all()
    np.testing.assert_allclose(got_w, want_w, rtol=2e-6, atol=1e-7)
    np.testing.assert_allclose(got_g, want_g, rtol=2e-6, atol=1e-7)


def test_update_sees_the_incremented_iteration(gpu):
    """burn_in=4, power=1: the first update runs at iteration 1 with lr/4 (the reference increments before it updates); a run
    that used iteration 0 would leave the weights where they are"""
    from darknet_b200 import lib as L, network as N
    text = cfgs.micro(batch=B, subdivisions=1, classes=CLASSES).replace("burn_in=0", "burn_in=4\npower=1")
    weights = cfgmod.synth_weights(text)
    net = N.LoadSynthetic(text, weights, batch=B, flags=L.NET_DEFAULT | L.NET_TRAIN)
    assert abs(L.lib().yc_network_current_rate(net.h, 1, 0) - 0.001 / 4) < 1e-9
    i = [i for i in range(net.n) if net.layer_info(i, 0) == 0][-1]
    N.TrainNetworkDatum(net, cfgmod.synth_images(B, 3, HW, HW, seed=3), cfgmod.synth_truth(B, 90, classes=CLASSES, seed=4))
    w0, g0 = N.layer_param(net, i, "biases"), N.layer_param(net, i, "bias_updates")
    N.UpdateNetwork(net)
    want = w0 + f32(0.001 / 4 / B) * g0
    np.testing.assert_allclose(N.layer_param(net, i, "biases"), want, rtol=1e-6, atol=1e-9)
    assert net.layer_info(0, 0) == 0 and L.lib().yc_network_info(net.h, 9) == 1, "curr_iter after the first update"
    net.close()

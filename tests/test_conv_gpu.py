"""Parity of the tcgen05 implicit-GEMM conv (through the C ABI) against the CPU oracle.

Inputs/weights are rounded to bf16 first, so the only differences left are fp32 accumulation order
and the bf16 rounding of the output: tolerance 1e-2 relative to the layer's max (north_star bf16
bound) for bf16 outputs, 1e-4 for fp32 outputs (the "fp32-accumulate check").
"""
import numpy as np
import pytest
import torch

from oracle.pyoracle import O

pytestmark = pytest.mark.gpu


def _bf16_round(a):
    return torch.from_numpy(a).to(torch.bfloat16).float().numpy()


def _run(gpu, B, C, H, W, N, k, stride, act, out_f32=False, residual=False, seed=0, scale=False, plan_flags=0, split=None,
         want_variant=None):
    from darknet_b200 import ops
    rng = np.random.default_rng(seed)
    pad = k // 2
    x = _bf16_round(rng.standard_normal((B, C, H, W)).astype(np.float32))
    w = _bf16_round((rng.standard_normal((N, C, k, k)) * np.sqrt(2.0 / (C * k * k))).astype(np.float32))
    b = rng.uniform(-0.5, 0.5, N).astype(np.float32)
    sc = rng.uniform(0.5, 1.5, N).astype(np.float32) if scale else None
    wref = w * sc[:, None, None, None] if scale else w
    ref, pre = O.conv_forward(x, wref, b, k, stride, pad, act, want_pre=True)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
    res_t = None
    if residual:
        r = _bf16_round(rng.standard_normal(ref.shape).astype(np.float32))
        ref = ref + r
        res_t = ops.to_nhwc(torch.from_numpy(r).to(gpu))
    conv = ops.Conv(xd, (B, H, W, C), torch.from_numpy(w), torch.from_numpy(b), k, stride, pad, act,
                    scale=None if sc is None else torch.from_numpy(sc),
                    out_dtype=torch.float32 if out_f32 else torch.bfloat16, residual=res_t, plan_flags=plan_flags, split=split)
    if want_variant is not None:  # (mask, value) on yc_conv_plan_info(plan, 7): the case must cover the path it names
        assert conv.info()["variant"] & want_variant[0] == want_variant[1], conv.info()
    y = conv()
    torch.cuda.synchronize()
    if split is None:
        got = ops.to_nchw(y, N).cpu().numpy()
    else:
        got = np.concatenate([ops.to_nchw(y, split).cpu().numpy(), ops.to_nchw(conv.y2, N - split).cpu().numpy()], axis=1)
    assert got.shape == ref.shape
    denom = np.abs(ref).max() + 1e-12
    err = np.abs(got - ref).max() / denom
    tol = 2e-4 if out_f32 else 1e-2
    assert err < tol, f"rel-to-max err {err:.3e} (tol {tol}) info={conv.info()}"
    return err


@pytest.mark.parametrize("C,N,HW", [(64, 64, 16), (128, 256, 19), (256, 128, 13), (512, 255, 8), (64, 32, 24)])
def test_conv1x1_tiled_map(gpu, C, N, HW):
    _run(gpu, 2, C, HW, HW, N, 1, 1, "leaky", out_f32=(N == 255))


@pytest.mark.parametrize("C,N,HW,stride", [(64, 64, 16, 1), (128, 128, 19, 1), (64, 128, 32, 2), (256, 512, 13, 1),
                                           (128, 256, 19, 2), (64, 96, 7, 1)])
def test_conv3x3_im2col_map(gpu, C, N, HW, stride):
    _run(gpu, 2, C, HW, HW, N, 3, stride, "mish")


def test_conv_first_layer_c3(gpu):
    _run(gpu, 2, 3, 32, 32, 32, 3, 1, "mish")
    _run(gpu, 1, 3, 33, 47, 32, 3, 2, "leaky")


@pytest.mark.parametrize("C", [32, 48, 96, 192, 384])
def test_conv_small_k_chunks(gpu, C):
    _run(gpu, 1, C, 12, 12, 64, 3, 1, "leaky")
    _run(gpu, 1, C, 12, 12, 64, 1, 1, "linear")


def test_conv_fp32_out_and_scale(gpu):
    _run(gpu, 2, 128, 10, 10, 24, 1, 1, "linear", out_f32=True)
    _run(gpu, 2, 64, 10, 10, 64, 3, 1, "leaky", out_f32=True, scale=True)


def test_conv_residual_fused(gpu):
    _run(gpu, 2, 64, 19, 19, 128, 3, 1, "mish", residual=True)
    _run(gpu, 2, 128, 19, 19, 64, 1, 1, "leaky", residual=True)


def test_conv_many_tiles_persistent(gpu):
    # > 148 tiles per CTA wave and both TMEM buffers / several ring wraps
    _run(gpu, 8, 64, 76, 76, 64, 3, 1, "leaky")
    _run(gpu, 4, 128, 38, 38, 512, 3, 1, "leaky")


def test_conv_ragged_rect(gpu):
    _run(gpu, 3, 64, 5, 11, 72, 3, 1, "logistic")
    _run(gpu, 1, 64, 1, 1, 64, 1, 1, "swish")
    _run(gpu, 1, 64, 3, 2, 64, 3, 2, "relu")


def test_conv_fused_pair_two_outputs(gpu):
    """Two 1x1 layers reading the same tensor as one GEMM with two output tensors (yolov4's CSP branch pairs)."""
    for C, n1, n2, HW in [(64, 64, 64, 24), (128, 64, 64, 19), (256, 128, 128, 19), (512, 256, 256, 13), (64, 32, 40, 9)]:
        _run(gpu, 2, C, HW, HW, n1 + n2, 1, 1, "mish", split=n1, want_variant=(16, 16))
    _run(gpu, 2, 128, 10, 10, 64 + 24, 1, 1, "linear", out_f32=True, split=64, want_variant=(16, 16))


def test_conv_plan_switches(gpu):
    """Every plan switch selects the variant it names and stays parity-green (the defaults are covered by the tests above)."""
    from darknet_b200 import lib as L
    # halo patches on / off, tile pairs off, weights never resident, one staging buffer, round-1 tile widths
    _run(gpu, 4, 64, 38, 38, 64, 3, 1, "mish", plan_flags=L.PLAN_FORCE_HALO, want_variant=(3, 2))
    _run(gpu, 4, 64, 38, 38, 64, 3, 1, "mish", plan_flags=L.PLAN_NO_HALO, want_variant=(3, 1))
    _run(gpu, 4, 64, 38, 38, 64, 3, 1, "mish", plan_flags=L.PLAN_FORCE_HALO | L.PLAN_NO_HALO_PAIR, want_variant=(7, 2))
    _run(gpu, 4, 64, 38, 38, 64, 3, 1, "mish", residual=True, plan_flags=L.PLAN_FORCE_HALO, want_variant=(3, 2))
    _run(gpu, 4, 32, 38, 38, 64, 3, 2, "leaky", plan_flags=L.PLAN_NO_W_RES, want_variant=(8, 0))
    _run(gpu, 2, 128, 19, 19, 128, 1, 1, "mish", plan_flags=L.PLAN_ONE_OUT_BUF, want_variant=(32, 0))
    _run(gpu, 2, 128, 19, 19, 128, 1, 1, "mish", residual=True, plan_flags=L.PLAN_ONE_OUT_BUF, want_variant=(32, 0))
    _run(gpu, 2, 128, 19, 19, 128, 1, 1, "mish", residual=True, plan_flags=L.PLAN_TWO_OUT_BUFS, want_variant=(32, 32))
    _run(gpu, 2, 128, 19, 19, 128, 1, 1, "mish", plan_flags=L.PLAN_TWO_OUT_BUFS, want_variant=(32, 32))
    # the generic epilogue (what fp32 heads, unfused batchnorm and training statistics use) on a layer the specialised one
    # would take, and the specialised one with the generic warp-to-tile assignment
    _run(gpu, 2, 128, 19, 19, 128, 1, 1, "mish", residual=True, plan_flags=L.PLAN_GENERIC_EPI, want_variant=(128, 0))
    _run(gpu, 4, 64, 38, 38, 64, 3, 1, "leaky", plan_flags=L.PLAN_GENERIC_EPI | L.PLAN_FORCE_HALO, want_variant=(128 | 3, 2))
    _run(gpu, 2, 128, 19, 19, 256, 1, 1, "leaky", plan_flags=L.PLAN_EPI_SETS_V1, want_variant=(128, 128))
    _run(gpu, 2, 128, 19, 19, 256, 1, 1, "leaky", want_variant=(128, 128))
    _run(gpu, 8, 1024, 19, 19, 512, 1, 1, "leaky", want_variant=(128, 128))   # bias from global memory (no free shared memory)
    # tile pairs on the im2col / tiled path (two accumulators alternate), with and without resident weights, odd tile count
    _run(gpu, 64, 32, 152, 152, 64, 3, 2, "mish", want_variant=(256 | 8, 256 | 8))
    _run(gpu, 64, 64, 76, 76, 128, 3, 2, "mish", want_variant=(256 | 8, 256))
    _run(gpu, 37, 256, 38, 38, 64, 1, 1, "leaky", residual=True, want_variant=(256, 256))
    _run(gpu, 64, 64, 76, 76, 128, 3, 2, "mish", plan_flags=L.PLAN_NO_TILE_PAIR, want_variant=(256, 0))
    _run(gpu, 8, 512, 19, 19, 512, 3, 1, "leaky", plan_flags=L.PLAN_COST_V1)
    _run(gpu, 8, 512, 19, 19, 512, 3, 1, "leaky")


@pytest.mark.parametrize("entry", ["yc_stem_conv_direct", "yc_stem_conv_tc"])
@pytest.mark.parametrize("B,H,W,N,act", [(2, 40, 72, 32, "mish"), (1, 16, 32, 32, "leaky"), (3, 19, 33, 16, "linear"),
                                         (2, 64, 64, 24, "mish"), (1, 5, 7, 8, "leaky"), (5, 200, 328, 32, "mish")])
def test_stem_conv_direct(gpu, B, H, W, N, act, entry):
    """yc_stem_conv_direct (3 -> N channels, 3x3 / stride 1 / pad 1 straight from the fp32 NCHW image; mma.sync with register
    A fragments) and yc_stem_conv_tc (tcgen05: operand tiles built in shared memory by the kernel, accumulators in TMEM, TMA
    stores) — neither keeps an operand in global memory — against the oracle conv on the same bf16-rounded inputs; ragged
    sizes cover partial tiles in both directions, N < 32 the channel guard, a wider pixel pitch the concat placement, and
    the last case gives every persistent CTA several tiles (all A slots and accumulators wrap around)."""
    from darknet_b200 import lib as L, ops
    lib = L.lib()
    rng = np.random.default_rng(B * 100 + H)
    x = _bf16_round(rng.standard_normal((B, 3, H, W)).astype(np.float32))
    w = _bf16_round((rng.standard_normal((N, 3, 3, 3)) * np.sqrt(2.0 / 27)).astype(np.float32))
    b = rng.uniform(-0.5, 0.5, N).astype(np.float32)
    ref = O.conv_forward(x, w, b, 3, 1, 1, act)
    # filter in the stem order: [n][32], column (kh*3+kw)*3 + ci
    w2 = np.zeros((N, 32), np.float32)
    for ci in range(3):
        for t in range(9):
            w2[:, t * 3 + ci] = w[:, ci, t // 3, t % 3]
    wd = torch.from_numpy(w2).to(gpu)
    packed = torch.zeros(N * 32, dtype=torch.bfloat16, device=gpu)
    L.check(lib.yc_conv_pack(wd.data_ptr(), N, 32, 1, packed.data_ptr(), L.stream_ptr()), "yc_conv_pack")
    xd = torch.from_numpy(x).to(gpu)
    bd = torch.from_numpy(b).to(gpu)
    ldy = 40 if N == 24 else N        # N == 24: the output is a channel slice of a wider buffer
    y = torch.full((B, H, W, ldy), 3.0, dtype=torch.bfloat16, device=gpu)
    L.check(getattr(lib, entry)(xd.data_ptr(), B, H, W, packed.data_ptr(), bd.data_ptr(), L.ACT[act],
                                L.view(y, B, H, W, N, ldy), L.stream_ptr()), entry)
    torch.cuda.synchronize()
    got = y[..., :N].float().permute(0, 3, 1, 2).cpu().numpy()
    err = np.abs(got - ref).max() / (np.abs(ref).max() + 1e-12)
    assert err < 1e-2, f"rel-to-max err {err:.3e}"
    if ldy > N:
        assert float((y[..., N:].float() - 3.0).abs().max()) == 0.0, "wrote outside its channel slice"


def test_conv_cta_pairs(gpu):
    """conv_fprop_pair_kernel (tcgen05 cta_group::2: 256-row tiles over a cluster of two CTAs, each staging its own A rows
    and half of the weight rows): the default rule (256-wide tiles, specialised epilogue) and, forced, every operand shape it
    accepts — odd tile counts (the last pair's second CTA entirely past M), several N tiles, many tiles per pair, narrow
    tiles, 32-channel K chunks, stride 2, fused shortcut, two output tensors."""
    from darknet_b200 import lib as L
    PA = L.PLAN_CTA_PAIR_ALL
    _run(gpu, 8, 256, 38, 38, 256, 3, 1, "leaky", want_variant=(512 | 3, 512 | 1))          # default rule, im2col
    _run(gpu, 8, 256, 38, 38, 256, 3, 1, "leaky", plan_flags=L.PLAN_NO_CTA_PAIR, want_variant=(512, 0))
    _run(gpu, 64, 128, 19, 19, 512, 3, 1, "leaky", want_variant=(512, 512))                 # 2 N tiles, 182 pair tiles on 74 pairs
    _run(gpu, 64, 1024, 19, 19, 256, 1, 1, "mish", want_variant=(512 | 3, 512 | 0))         # default rule, 1x1 tiled map, 16 K blocks
    _run(gpu, 64, 512, 19, 19, 256, 1, 1, "mish", want_variant=(512, 512))                  # 8 K blocks: the threshold
    _run(gpu, 64, 256, 19, 19, 256, 1, 1, "mish", want_variant=(512, 0))                    # 4 K blocks: stays on one CTA per tile
    _run(gpu, 3, 512, 19, 19, 256, 3, 1, "mish", plan_flags=PA, want_variant=(512, 512))    # 9 row tiles: odd
    _run(gpu, 5, 512, 19, 19, 256, 1, 1, "leaky", residual=True, plan_flags=PA, want_variant=(512, 512))
    _run(gpu, 4, 64, 38, 38, 64, 1, 1, "mish", plan_flags=PA | L.PLAN_NO_HALO, want_variant=(512, 512))
    _run(gpu, 4, 64, 38, 38, 128, 3, 2, "mish", plan_flags=PA, want_variant=(512, 512))
    _run(gpu, 2, 96, 24, 24, 96, 3, 1, "linear", plan_flags=PA | L.PLAN_NO_HALO, want_variant=(512, 512))
    _run(gpu, 1, 64, 5, 5, 40, 3, 1, "leaky", plan_flags=PA | L.PLAN_NO_HALO, want_variant=(512, 512))   # one tile, ragged
    _run(gpu, 2, 256, 19, 19, 256, 1, 1, "mish", split=128, plan_flags=PA, want_variant=(16 | 512, 16 | 512))
    # halo patches on CTA pairs (each CTA its own 16 x 8 block, tile pairs alternate between two accumulators): the default
    # rule (128 -> 128), then forced: resident weights + shortcut, an odd number of blocks, 32-channel chunks, three chunks
    FH = L.PLAN_FORCE_HALO
    _run(gpu, 4, 128, 76, 76, 128, 3, 1, "mish", want_variant=(512 | 3 | 4, 512 | 2 | 4))
    _run(gpu, 4, 128, 76, 76, 128, 3, 1, "mish", residual=True, plan_flags=L.PLAN_NO_CTA_PAIR, want_variant=(512 | 3, 2))
    _run(gpu, 4, 64, 38, 38, 64, 3, 1, "mish", residual=True, plan_flags=PA | FH, want_variant=(512 | 3, 512 | 2))
    _run(gpu, 64, 64, 38, 38, 64, 3, 1, "mish", residual=True, plan_flags=PA | FH, want_variant=(512 | 3 | 8, 512 | 2 | 8))  # one N tile: resident weights
    _run(gpu, 3, 64, 38, 22, 64, 3, 1, "leaky", residual=True, plan_flags=PA | FH, want_variant=(512 | 3, 512 | 2))
    _run(gpu, 32, 64, 38, 38, 64, 3, 1, "linear", plan_flags=PA | FH, want_variant=(512 | 3, 512 | 2))  # 480 blocks on 74 pairs
    _run(gpu, 4, 32, 38, 38, 64, 3, 1, "mish", plan_flags=PA | FH, want_variant=(512 | 3, 512 | 2))
    _run(gpu, 2, 96, 24, 24, 128, 3, 1, "leaky", plan_flags=PA | FH, want_variant=(512 | 3, 512 | 2))
    _run(gpu, 2, 256, 19, 19, 256, 3, 1, "leaky", plan_flags=PA | FH, want_variant=(512 | 3, 512 | 2))  # two N tiles? no: one of 256
    # not legal -> stays on the one-CTA kernel whatever the flag says: fp32 output, unfused scale
    _run(gpu, 2, 128, 10, 10, 24, 1, 1, "linear", out_f32=True, plan_flags=PA, want_variant=(512, 0))
    _run(gpu, 2, 64, 10, 10, 64, 3, 1, "leaky", scale=True, plan_flags=PA | L.PLAN_NO_HALO, want_variant=(512, 0))


def test_mish_newton_reciprocal_variant(gpu):
    """The specialised epilogue's opt-in mish (one MUFU.EX2 + a Newton reciprocal on the FMA pipe, 9e-6 relative) against
    the default two-MUFU form on the same accumulators: the bf16 outputs may differ by one ulp on a handful of elements that
    sit on a rounding boundary (or by the forms' 2e-6 absolute error where the output is ~0), nowhere by more."""
    from darknet_b200 import lib as L, ops
    rng = np.random.default_rng(7)
    B, C, HW, N = 4, 64, 38, 128
    x = _bf16_round((3.0 * rng.standard_normal((B, C, HW, HW))).astype(np.float32))   # pre-activations spread over [-25, 25]
    w = _bf16_round((rng.standard_normal((N, C, 1, 1)) * 0.3).astype(np.float32))
    b = rng.uniform(-2, 2, N).astype(np.float32)
    xd = ops.to_nhwc(torch.from_numpy(x).to(gpu))
    outs = []
    for flags in (L.PLAN_MISH_NEWTON, 0):
        conv = ops.Conv(xd, (B, HW, HW, C), torch.from_numpy(w), torch.from_numpy(b), 1, 1, 0, "mish", plan_flags=flags)
        assert conv.info()["variant"] & 128
        outs.append(ops.to_nchw(conv(), N).cpu().numpy())
        torch.cuda.synchronize()
    a, m = outs
    ref = O.conv_forward(x, w, b, 1, 1, 0, "mish")
    assert np.abs(a - ref).max() <= 1e-2 * np.abs(ref).max() and np.abs(m - ref).max() <= 1e-2 * np.abs(ref).max()
    diff = a != m
    ulp = np.maximum(np.abs(m) * 2.0 ** -7, 4e-6)       # one bf16 ulp is at most 2^-7 relative
    assert np.all(np.abs(a - m)[diff] <= ulp[diff] * 1.001), "more than one bf16 ulp apart"
    assert diff.mean() < 0.15, f"{diff.mean():.4f} of the outputs differ"  # measured 0.06: one ulp, either way
    # and the error against the fp32 reference is no worse than the MUFU variant's
    assert np.abs(a - ref).mean() <= 1.02 * np.abs(m - ref).mean()

"""Per-kernel parity on the GPU, through the C-ABI (b200net.ops -> ctypes -> libb200net.so), against stock torch
in float64 on the same seeded inputs.  Shapes follow SURVEY.md 8(a') (conv classes S,a,b,c,f,h,l,...) at small N,
plus odd sizes.  Tolerances: fp32 path 2e-5 of the output scale (FFMA, different summation order only);
bf16 path 2e-2 of the output scale (bf16 storage of the result, fp32 accumulation)."""
import math

import pytest
import torch
import torch.nn.functional as TF

pytestmark = pytest.mark.gpu

DT = {"fp32": torch.float32, "bf16": torch.bfloat16}
TOL = {"fp32": 2e-5, "bf16": 2e-2}


def _dev():
    return torch.device("cuda:0")


def _rel(a, b):
    a, b = a.double(), b.double()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()


def _nhwc(t, dtype):
    return t.permute(0, 2, 3, 1).contiguous().to(dtype)


def _rand(shape, dtype, seed, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(shape, generator=g) * scale).to(dtype).to(_dev())


# ---------------------------------------------------------------------------------------------- layout
def test_layout_roundtrip():
    from b200net import ops
    for C, H, W in ((3, 17, 23), (64, 9, 10), (130, 5, 7)):
        x = torch.rand(2, C, H, W, device=_dev())
        for dt in DT.values():
            y = ops.to_nhwc(x, dt)
            assert y.shape == (2, H, W, C)
            assert torch.equal(y, x.permute(0, 2, 3, 1).to(dt))
            assert torch.equal(ops.to_nchw(y), x.to(dt).float())


# ---------------------------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (300, 200, 136), (33, 10, 520), (1000, 72, 4100)])
def test_gemm_f32(a_mn, b_mn, M, N, K):
    from b200net import ops
    A = _rand((M, K), torch.float32, 1)
    B = _rand((N, K), torch.float32, 2)
    bias = _rand((N,), torch.float32, 3)
    As = A.t().contiguous() if a_mn else A
    Bs = B.t().contiguous() if b_mn else B
    ref = A.double() @ B.double().t()
    C = ops.gemm_f32(As, Bs, M, N, K, As.shape[1], Bs.shape[1], a_mn, b_mn)
    assert _rel(C, ref) < 2e-6 * math.sqrt(K)
    C2 = ops.gemm_f32(As, Bs, M, N, K, As.shape[1], Bs.shape[1], a_mn, b_mn, bias=bias, relu=True)
    assert _rel(C2, torch.relu(ref + bias.double())) < 2e-6 * math.sqrt(K)
    C3 = C.clone()
    ops.gemm_f32(As, Bs, M, N, K, As.shape[1], Bs.shape[1], a_mn, b_mn, out=C3, accumulate=True)
    assert _rel(C3, 2 * ref) < 2e-6 * math.sqrt(K)


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("M,N,K,splits", [(256, 256, 512, 1), (300, 200, 136, 1), (64, 576, 6000, 9), (1000, 72, 264, 2)])
def test_gemm_bf16_tcgen05(a_mn, b_mn, M, N, K, splits):
    """tcgen05.mma + TMA + TMEM GEMM through the C-ABI, every operand major-ness, M/N/K tails, split-K."""
    import b200net
    from b200net._C import lib, check
    pad8 = lambda v: (v + 7) // 8 * 8
    A = _rand((M, K), torch.bfloat16, 1)
    B = _rand((N, K), torch.bfloat16, 2)
    ref = A.double() @ B.double().t()

    def store(X, mn):  # leading dimensions padded to multiples of 8 elements (16-byte TMA strides)
        rows, cols = (X.shape[1], X.shape[0]) if mn else X.shape
        buf = torch.zeros((rows, pad8(cols)), dtype=torch.bfloat16, device=_dev())
        buf[:, :cols] = X.t() if mn else X
        return buf

    As, Bs = store(A, a_mn), store(B, b_mn)
    ldc = pad8(N)
    st = torch.cuda.current_stream().cuda_stream
    for out_dtype, tdt, tol in ((0, torch.float32, 1e-5 * math.sqrt(K)), (1, torch.bfloat16, 1e-2)):
        C = torch.full((M, ldc), float("nan"), dtype=tdt, device=_dev())
        ws = torch.empty(max(1, splits) * M * ldc * 4 + 16, dtype=torch.uint8, device=_dev())
        check(lib.b200_gemm_bf16(As.data_ptr(), Bs.data_ptr(), C.data_ptr(), M, N, K, As.shape[1], Bs.shape[1], ldc,
                                 a_mn, b_mn, out_dtype, 0, splits, ws.data_ptr(), ws.numel(), st), "gemm_bf16")
        assert _rel(C[:, :N], ref) < tol
        check(lib.b200_gemm_bf16(As.data_ptr(), Bs.data_ptr(), C.data_ptr(), M, N, K, As.shape[1], Bs.shape[1], ldc,
                                 a_mn, b_mn, out_dtype, 1, splits, ws.data_ptr(), ws.numel(), st), "gemm_bf16 acc")
        assert _rel(C[:, :N], 2 * ref) < 2 * tol


# ---------------------------------------------------------------------------------------------- convolution
CONV_CASES = [
    # N, H, C, K, R, stride, pad        (SURVEY 8(a') class)
    (2, 40, 3, 64, 7, 2, 3),     # S  stem (C=3, K-dim 147 -> padded 152)
    (2, 20, 64, 64, 1, 1, 0),    # a  1x1
    (2, 20, 64, 64, 3, 1, 1),    # b  3x3
    (2, 20, 64, 256, 1, 1, 0),   # c
    (3, 20, 128, 128, 3, 2, 1),  # f  3x3 stride 2
    (2, 20, 256, 512, 1, 2, 0),  # h  1x1 stride 2 downsample
    (2, 10, 256, 256, 3, 1, 1),  # p
    (4, 5, 512, 2048, 1, 1, 0),  # s  tiny M
    (1, 21, 64, 72, 3, 2, 1),    # odd H, Cout not a multiple of 64
    (1, 13, 24, 40, 3, 1, 1),    # small odd channel counts
    (2, 14, 128, 128, 3, 1, 1),  # 224-px family: 14x14
    (1, 21, 64, 128, 3, 2, 1),   # odd H, stride 2 (implicit path: parity classes with ragged extents)
    (3, 7, 512, 512, 3, 1, 1),   # 7x7 map
    (1, 11, 64, 64, 1, 2, 0),    # odd H, 1x1 stride 2
    (5, 5, 512, 512, 3, 1, 1),   # 5x5 map, N not a multiple of the pixel box
    (2, 10, 512, 512, 3, 2, 1),  # r  10 -> 5
]


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_fprop_dgrad_wgrad(prec, case):
    from b200net import ops
    N, H, C, K, R, stride, pad = case
    dt = DT[prec]
    x = _rand((N, C, H, H), dt, 11)
    w = _rand((K, C, R, R), torch.float32, 12, scale=1.0 / math.sqrt(C * R * R))
    wq = w.to(dt)  # what the kernel sees after packing
    xd = x.double().requires_grad_(True)
    wd = wq.double().requires_grad_(True)
    yref = TF.conv2d(xd, wd, stride=stride, padding=pad)
    dy = _rand(tuple(yref.shape), dt, 13)
    yref.backward(dy.double())

    xh = _nhwc(x, dt)
    wp = ops.pack_weight(w, dt)
    y = ops.conv_fprop(xh, wp, K, R, R, stride, pad)
    assert y.shape == (N, yref.shape[2], yref.shape[3], K)
    tol = TOL[prec]
    assert _rel(y.permute(0, 3, 1, 2), yref) < tol

    dyh = _nhwc(dy, dt)
    dx = ops.conv_dgrad(dyh, wp, xh.shape, K, R, R, stride, pad)
    assert _rel(dx.permute(0, 3, 1, 2), xd.grad) < tol
    base = _rand(tuple(xh.shape), dt, 14)
    acc = base.clone()
    ops.conv_dgrad(dyh, wp, xh.shape, K, R, R, stride, pad, out=acc, accumulate=True)
    assert _rel(acc.permute(0, 3, 1, 2), xd.grad + base.double().permute(0, 3, 1, 2)) < tol

    dw = ops.conv_wgrad(xh, dyh, K, R, R, stride, pad)
    assert dw.shape == w.shape and dw.dtype == torch.float32
    assert _rel(dw, wd.grad) < (1e-4 if prec == "fp32" else 5e-3)


# Shapes of the BENCHMARKED configuration (bench.py: N=256 at 160 px, bf16): every case has several times more output
# tiles than the B200 has SMs, so each persistent CTA walks many tiles, both TMEM accumulators wrap around several
# times, the smem ring's phase bits flip repeatedly, and the per-CTA BatchNorm statistics accumulate over many tiles.
# (The small cases above never give a CTA a second tile.)  fp64 cuDNN/ATen on the GPU is the checker.
BIG_CONV_CASES = [
    # N, H, C, K, R, stride, pad     tiles (fprop, 128-pixel boxes x column tiles)
    (64, 40, 64, 256, 1, 1, 0),      # c: 800 x 1
    (64, 40, 64, 64, 3, 1, 1),       # b: 800 x 1
    (128, 40, 128, 128, 3, 2, 1),    # f: 400 x 1, dgrad = four parity classes
    (256, 10, 256, 1024, 1, 1, 0),   # m at the benchmark batch: 200 x 4
    (1024, 5, 512, 512, 3, 1, 1),    # v: 205 x 2, 125-row boxes (ragged M tile)
    (96, 20, 256, 512, 1, 2, 0),     # h: 1x1 stride-2 downsample
]


@pytest.mark.parametrize("case", BIG_CONV_CASES)
def test_conv_bf16_multi_tile_persistent_walk(case):
    from b200net import ops
    N, H, C, K, R, stride, pad = case
    dt = torch.bfloat16
    x = _rand((N, C, H, H), dt, 11)
    w = _rand((K, C, R, R), torch.float32, 12, scale=1.0 / math.sqrt(C * R * R))
    xd = x.double().requires_grad_(True)
    wd = w.to(dt).double().requires_grad_(True)
    yref = TF.conv2d(xd, wd, stride=stride, padding=pad)
    dy = _rand(tuple(yref.shape), dt, 13)
    yref.backward(dy.double())
    xh, dyh = _nhwc(x, dt), _nhwc(dy, dt)
    wp = ops.pack_weight(w, dt)
    tol = TOL["bf16"]

    y, part = ops.conv_fprop_bnstats(xh, wp, K, R, R, stride, pad)
    assert part is not None, "benchmark shapes must be on the fused-statistics path"
    ntiles = ((N * yref.shape[2] * yref.shape[3]) // 128) * ((K + 255) // 256)
    assert ntiles >= 148, ntiles
    assert _rel(y.permute(0, 3, 1, 2), yref) < tol
    assert torch.equal(y, ops.conv_fprop(xh, wp, K, R, R, stride, pad))
    yd = y.double().reshape(-1, K)
    assert _rel(part[:, 0].double().sum(0), yd.sum(0)) < 1e-5
    assert _rel(part[:, 1].double().sum(0), (yd * yd).sum(0)) < 1e-5

    dx = ops.conv_dgrad(dyh, wp, xh.shape, K, R, R, stride, pad)
    assert _rel(dx.permute(0, 3, 1, 2), xd.grad) < tol
    base = _rand(tuple(xh.shape), dt, 14)
    acc = base.clone()
    ops.conv_dgrad(dyh, wp, xh.shape, K, R, R, stride, pad, out=acc, accumulate=True)
    assert _rel(acc.permute(0, 3, 1, 2), xd.grad + base.double().permute(0, 3, 1, 2)) < tol

    dw = ops.conv_wgrad(xh, dyh, K, R, R, stride, pad)
    assert _rel(dw, wd.grad) < 5e-3
    # deterministic: a second launch gives the same bits (ordered split-K reduction, no atomics)
    assert torch.equal(dw, ops.conv_wgrad(xh, dyh, K, R, R, stride, pad))


@pytest.mark.parametrize("N,H,W", [(2, 40, 40), (3, 37, 42), (1, 224, 224)])
def test_stem_row_pair_staging(N, H, W):
    """The stem (Conv2d(3, 64, 7, stride 2, padding 3), pytorch_training_inference_on_image.ipynb:455) in bf16 mode:
    channel-padded image -> row-pair window tensor -> 4x1 stride-1 convolution over 64 channels on the implicit
    tcgen05 path, forward and weight gradient, against float64 conv2d (odd sizes included)."""
    from b200net import ops
    K = 64
    x = torch.rand(N, 3, H, W, generator=torch.Generator().manual_seed(5)).to(_dev())
    w = _rand((K, 3, 7, 7), torch.float32, 6, scale=1.0 / math.sqrt(147))
    xq = x.to(torch.bfloat16)
    xd = xq.double()
    wd = w.to(torch.bfloat16).double().requires_grad_(True)
    yref = TF.conv2d(xd, wd, stride=2, padding=3)
    dy = _rand(tuple(yref.shape), torch.bfloat16, 7)
    yref.backward(dy.double())
    x4 = ops.to_nhwc(x, torch.bfloat16, 4)
    st = ops.stem_stage(x4, 7, 7, 3)
    Ho, Wo = yref.shape[2], yref.shape[3]
    assert st.shape == (N, Ho + 3, Wo, 64)
    wp = ops.stem_pack_weight(w)
    y, part = ops.conv_fprop_bnstats(st, wp, K, 4, 1, 1, 0)
    assert y.shape == (N, Ho, Wo, K) and part is not None
    assert _rel(y.permute(0, 3, 1, 2), yref) < TOL["bf16"]
    dw = ops.conv_wgrad(st, _nhwc(dy, torch.bfloat16), K, 4, 1, 1, 0)          # [K][64 = (r2, s, c)][4 = j][1]
    dw = dw.reshape(K, 2, 8, 4, 4).permute(0, 3, 4, 1, 2).reshape(K, 4, 8, 8)[:, :3, :7, :7]
    assert _rel(dw, wd.grad) < 5e-3


@pytest.mark.parametrize("N,C,H,W", [(3, 3, 160, 160), (2, 3, 37, 45), (1, 1, 16, 20), (2, 4, 21, 18)])
def test_stem_stage_from_nchw_matches_two_pass(N, C, H, W):
    """b200_stem_stage_nchw (fp32 NCHW batch -> row-pair window tensor in one pass) is bit-identical to
    b200_nchw_to_nhwc (4-channel bf16) followed by b200_stem_stage."""
    from b200net import ops
    x = _rand((N, C, H, W), torch.float32, 71)
    two = ops.stem_stage(ops.to_nhwc(x, torch.bfloat16, 4), 7, 7, 3)
    one = ops.stem_stage_nchw(x, 7, 7, 3)
    assert one.shape == two.shape and torch.equal(one, two)



@pytest.mark.parametrize("case", [c for c in CONV_CASES if c[2] % 64 == 0 and c[3] % 64 == 0])
def test_conv_fprop_fused_bn_statistics(case):
    """The fprop epilogue also emits per-pixel-tile column sums / sums of squares of the bf16 output: they must add up
    to the statistics of the tensor it wrote, and feed b200_bn_finalize_stats to the same result as the stand-alone
    statistics kernel."""
    from b200net import ops
    N, H, C, K, R, stride, pad = case
    x = _nhwc(_rand((N, C, H, H), torch.bfloat16, 11), torch.bfloat16)
    wp = ops.pack_weight(_rand((K, C, R, R), torch.float32, 12, scale=1.0 / math.sqrt(C * R * R)), torch.bfloat16)
    y0 = ops.conv_fprop(x, wp, K, R, R, stride, pad)
    y, part = ops.conv_fprop_bnstats(x, wp, K, R, R, stride, pad)
    assert part is not None and torch.equal(y, y0)
    yd = y.double().reshape(-1, K)
    assert _rel(part[:, 0].double().sum(0), yd.sum(0)) < 1e-5
    assert _rel(part[:, 1].double().sum(0), (yd * yd).sum(0)) < 1e-5
    gamma, beta = _rand((K,), torch.float32, 13).abs() + 0.5, _rand((K,), torch.float32, 14)
    rm1, rv1, rm2, rv2 = (torch.zeros(K, device=_dev()), torch.ones(K, device=_dev()),
                          torch.zeros(K, device=_dev()), torch.ones(K, device=_dev()))
    a = ops.bn_finalize_stats(part, yd.shape[0], gamma, beta, rm1, rv1, 0.1, 1e-5)
    b = ops.bn_train_stats(y, gamma, beta, rm2, rv2, 0.1, 1e-5)
    for u, v in zip(a, b):
        assert _rel(u, v) < 1e-5
    assert _rel(rm1, rm2) < 1e-5 and _rel(rv1, rv2) < 1e-5


def test_conv_wgrad_deferred_reductions_match_immediate():
    """b200_wgrad_defer / b200_wgrad_flush: the split-K slabs of several layers are reduced by ONE launch; the
    gradients are bit-identical to the per-layer reduction (same summation order), layers that write dw directly
    are unaffected, and leaving the mode drops nothing silently into later calls."""
    from b200net import ops
    dt = torch.bfloat16
    cases = [(8, 20, 64, 64, 3, 1, 1), (8, 20, 64, 256, 1, 1, 0), (8, 10, 256, 256, 3, 1, 1), (8, 10, 512, 128, 1, 1, 0),
             (4, 20, 128, 128, 3, 2, 1)]
    data = []
    for i, (N, H, C, K, R, stride, pad) in enumerate(cases):
        Ho = (H + 2 * pad - R) // stride + 1
        x = _nhwc(_rand((N, C, H, H), dt, 300 + i), dt)
        dy = _nhwc(_rand((N, K, Ho, Ho), dt, 400 + i), dt)
        data.append((x, dy, K, R, stride, pad, ops.conv_wgrad(x, dy, K, R, R, stride, pad)))
    bufs, got = [], []
    ops.wgrad_defer(True)
    try:
        for x, dy, K, R, stride, pad, _ in data:
            got.append(ops.conv_wgrad(x, dy, K, R, R, stride, pad, own_ws=bufs))
        ops.wgrad_flush()
    finally:
        ops.wgrad_defer(False)
    torch.cuda.synchronize()
    for (x, dy, K, R, stride, pad, ref), g in zip(data, got):
        assert torch.equal(g, ref), (tuple(x.shape), K, R, stride)
    x, dy, K, R, stride, pad, ref = data[0]            # back in immediate mode
    assert torch.equal(ops.conv_wgrad(x, dy, K, R, R, stride, pad), ref)


# conv3-of-a-Bottleneck shapes whose CTAs walk several tiles: the residual arrives by TMA, prefetched one item ahead
RESIDUAL_CASES = [(64, 40, 64, 256, 1, 1, 0), (64, 20, 128, 512, 1, 1, 0), (256, 10, 256, 1024, 1, 1, 0),
                  (64, 5, 512, 2048, 1, 1, 0), (64, 20, 512, 128, 1, 1, 0), (3, 7, 64, 256, 1, 1, 0)]


@pytest.mark.parametrize("case", [c for c in CONV_CASES if c[2] % 64 == 0 and c[3] % 64 == 0][:6] + RESIDUAL_CASES)
@pytest.mark.parametrize("relu", [True, False])
def test_conv_fprop_affine_epilogue(case, relu):
    """Eval-mode BatchNorm (+ReLU) folded into the fprop epilogue == conv followed by the separate apply kernel,
    up to the one bf16 rounding the fused path saves."""
    from b200net import ops
    N, H, C, K, R, stride, pad = case
    x = _nhwc(_rand((N, C, H, H), torch.bfloat16, 11), torch.bfloat16)
    wp = ops.pack_weight(_rand((K, C, R, R), torch.float32, 12, scale=1.0 / math.sqrt(C * R * R)), torch.bfloat16)
    scale = _rand((K,), torch.float32, 13).abs() + 0.5
    shift = _rand((K,), torch.float32, 14)
    z = ops.conv_fprop_affine(x, wp, K, R, R, stride, pad, scale, shift, relu)
    assert z is not None
    y = ops.conv_fprop(x, wp, K, R, R, stride, pad)
    ref = y.double() * scale.double() + shift.double()
    res = _rand(tuple(y.shape), torch.bfloat16, 15)
    zr = ops.conv_fprop_affine(x, wp, K, R, R, stride, pad, scale, shift, relu, residual=res)
    refr = ref + res.double()
    if relu:
        ref, refr = torch.relu(ref), torch.relu(refr)
    assert _rel(z, ref) < 2e-2 and _rel(zr, refr) < 2e-2
    assert (z >= 0).all() if relu else True


# ---------------------------------------------------------------------------------------------- batch norm
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("rows,C", [(2 * 20 * 20, 64), (3 * 7 * 5, 2048), (1000, 24), (50, 130)])
@pytest.mark.parametrize("relu,residual", [(True, False), (True, True), (False, False)])
def test_batchnorm_train_fwd_bwd(prec, rows, C, relu, residual):
    from b200net import ops
    dt = DT[prec]
    x = _rand((rows, C), dt, 21, scale=2.0) + 0.5
    res = _rand((rows, C), dt, 22) if residual else None
    gamma = _rand((C,), torch.float32, 23).abs() + 0.5
    beta = _rand((C,), torch.float32, 24)
    rm = torch.zeros(C, device=_dev())
    rv = torch.ones(C, device=_dev())
    mean, invstd, scale, shift = ops.bn_train_stats(x, gamma, beta, rm, rv, 0.1, 1e-5)
    z = ops.bn_apply(x, scale, shift, relu, res)

    xd = x.double().requires_grad_(True)
    gd = gamma.double().requires_grad_(True)
    bd = beta.double().requires_grad_(True)
    rmd, rvd = torch.zeros(C, dtype=torch.float64, device=_dev()), torch.ones(C, dtype=torch.float64, device=_dev())
    resd = res.double().requires_grad_(True) if residual else None
    zr = TF.batch_norm(xd, rmd, rvd, gd, bd, True, 0.1, 1e-5)
    if residual:
        zr = zr + resd
    if relu:
        zr = torch.relu(zr)
    tol = TOL[prec] if prec == "bf16" else 1e-5
    assert _rel(z, zr) < tol
    assert _rel(mean, xd.mean(0)) < 1e-5 and _rel(invstd, 1 / torch.sqrt(xd.var(0, unbiased=False) + 1e-5)) < 1e-5
    assert _rel(rm, rmd) < 1e-5 and _rel(rv, rvd) < 1e-5

    dz = _rand((rows, C), dt, 25)
    # the kernel derives the ReLU mask from ITS OWN saved output z, as the backbone does
    mask = (z > 0).double() if relu else torch.ones_like(zr)
    (zr_lin,) = (TF.batch_norm(xd, None, None, gd, bd, True, 0.1, 1e-5),)
    g = dz.double() * mask
    zr_lin.backward(g)
    dx, dres, dgamma, dbeta = ops.bn_backward(dz, z if relu else None, x, gamma, mean, invstd, relu, want_dres=residual)
    assert _rel(dx, xd.grad) < (tol if prec == "bf16" else 2e-4)
    assert _rel(dgamma, gd.grad) < 2e-4 and _rel(dbeta, bd.grad) < 2e-4
    if residual:
        assert _rel(dres, g) < tol
    # the same backward driven by the one-byte-per-vector ReLU mask written by bn_apply instead of by z
    z2, mask = ops.bn_apply(x, scale, shift, relu, res, want_mask=True)
    assert torch.equal(z2, z)
    if relu and mask is not None:
        W = 8 if prec == "bf16" else 4
        bits = (mask.view(-1, 1).int() >> torch.arange(W, device=_dev()).int().view(1, W)) & 1
        assert torch.equal(bits.view(rows, C).bool(), z > 0)
        dx2, dres2, dgamma2, dbeta2 = ops.bn_backward(dz, None, x, gamma, mean, invstd, relu, want_dres=residual, mask=mask)
        assert torch.equal(dx2, dx) and torch.equal(dgamma2, dgamma) and torch.equal(dbeta2, dbeta)
        if residual:
            assert torch.equal(dres2, dres)
    else:
        assert mask is None


def test_batchnorm_eval_coeffs():
    from b200net import ops
    C = 96
    x = _rand((77, C), torch.float32, 31)
    gamma, beta = _rand((C,), torch.float32, 32), _rand((C,), torch.float32, 33)
    rm, rv = _rand((C,), torch.float32, 34), _rand((C,), torch.float32, 35).abs() + 0.1
    scale, shift = ops.bn_eval_coeffs(gamma, beta, rm, rv, 1e-5)
    z = ops.bn_apply(x, scale, shift, False)
    ref = TF.batch_norm(x.double(), rm.double(), rv.double(), gamma.double(), beta.double(), False, 0.1, 1e-5)
    assert _rel(z, ref) < 1e-5


# ---------------------------------------------------------------------------------------------- pooling
@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("N,H,W,C", [(2, 80, 80, 64), (1, 7, 9, 24), (3, 6, 6, 8)])
def test_maxpool_avgpool(prec, N, H, W, C):
    from b200net import ops
    dt = DT[prec]
    x = torch.relu(_rand((N, C, H, W), dt, 41))  # post-ReLU input: exact ties at zero, like the stem
    xd = x.double().requires_grad_(True)
    yr = TF.max_pool2d(xd, 3, 2, 1)
    dy = _rand(tuple(yr.shape), dt, 42)
    yr.backward(dy.double())
    xh = _nhwc(x, dt)
    y, idx = ops.maxpool_fwd(xh, True)
    assert torch.equal(y.permute(0, 3, 1, 2).double(), yr.detach())
    dx = ops.maxpool_bwd(_nhwc(dy, dt), idx, xh.shape)
    assert _rel(dx.permute(0, 3, 1, 2), xd.grad) < TOL[prec]
    a = ops.avgpool_fwd(xh)
    assert _rel(a, x.double().mean((2, 3))) < 1e-5
    da = _rand((N, C), torch.float32, 43)
    dxa = ops.avgpool_bwd(da, xh.shape, dt)
    assert _rel(dxa, (da.double() / (H * W))[:, None, None, :].expand(N, H, W, C)) < TOL[prec]


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("N,H,W,C", [(4, 80, 80, 64), (3, 15, 13, 64), (2, 9, 10, 32), (5, 8, 8, 128)])
def test_stem_bn_relu_maxpool_one_pass(prec, N, H, W, C):
    """b200_bn_relu_maxpool_fwd / b200_maxpool_bn_bwd (the stem's BatchNorm -> ReLU -> MaxPool in one pass each way):
    pooled values and argmax are bit-identical to the separate apply + pool kernels, the gradient matches both the
    separate kernels and (fp32) torch's fp64 autograd through batch_norm -> relu -> max_pool2d."""
    from b200net import ops
    dt = DT[prec]
    y = _rand((N, H, W, C), dt, 51, scale=2.0)
    gamma = (torch.rand(C, device=_dev()) + 0.5)
    beta = torch.randn(C, device=_dev()) * 0.3
    mean, invstd, scale, shift = ops.bn_train_stats(y, gamma, beta, None, None, 0.0, 1e-5)
    z, mask = ops.bn_apply(y, scale, shift, True, None, want_mask=True)
    p_ref, i_ref = ops.maxpool_fwd(z, True)
    p, i = ops.bn_relu_maxpool_fwd(y, scale, shift, True)
    assert torch.equal(p, p_ref)
    assert torch.equal(ops.bn_relu_maxpool_fwd(y, scale, shift, False)[0], p_ref)
    # the arg-max names an element of the window that holds the pooled value (where different raw values round to
    # the same output, the one-pass kernel picks the larger raw value, apply-then-pool the first by position)
    pp = torch.arange(p.shape[1], device=_dev())[None, :, None, None]
    qq = torch.arange(p.shape[2], device=_dev())[None, None, :, None]
    hh, ww = 2 * pp - 1 + (i // 3).long(), 2 * qq - 1 + (i % 3).long()
    assert bool(((hh >= 0) & (hh < H) & (ww >= 0) & (ww < W)).all())
    nn = torch.arange(N, device=_dev())[:, None, None, None].expand_as(hh)
    cc = torch.arange(C, device=_dev())[None, None, None, :].expand_as(hh)
    assert torch.equal(z[nn, hh, ww, cc], p)
    assert ((i != i_ref) & (p > 0)).float().mean().item() < (1e-4 if prec == "fp32" else 0.05)
    dpool = _rand(tuple(p.shape), dt, 52)
    dz = ops.maxpool_bwd(dpool, i, z.shape)
    dy_ref, _, dg_ref, db_ref = ops.bn_backward(dz, None, y, gamma, mean, invstd, True, mask=mask)
    dy, dg, db = ops.maxpool_bn_bwd(dpool, i, p, y, gamma, mean, invstd)
    # the separate kernels round the pooling gradient to bf16 before the BatchNorm reduction, the fused pass does not
    tol_pg = 1e-5 if prec == "fp32" else 5e-3
    assert _rel(dg, dg_ref) < tol_pg and _rel(db, db_ref) < tol_pg
    assert _rel(dy, dy_ref) < (1e-5 if prec == "fp32" else 1e-2)
    assert ops.maxpool_bn_bwd(dpool, i, p, y, gamma, mean, invstd, want_param_grads=False)[1] is None
    if prec == "fp32":
        yd = y.permute(0, 3, 1, 2).double().requires_grad_(True)
        gd, bd = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
        out = TF.max_pool2d(torch.relu(TF.batch_norm(yd, None, None, gd, bd, True, 0.0, 1e-5)), 3, 2, 1)
        out.backward(dpool.permute(0, 3, 1, 2).double())
        assert _rel(p.permute(0, 3, 1, 2), out.detach()) < 1e-5
        assert _rel(dy.permute(0, 3, 1, 2), yd.grad) < 1e-4
        assert _rel(dg, gd.grad) < 1e-4 and _rel(db, bd.grad) < 1e-4


# ---------------------------------------------------------------------------------------------- head
def test_head_functions_match_torch():
    import b200net
    F = b200net.functional
    x = _rand((37, 2048), torch.float32, 51).requires_grad_(True)
    lin1 = torch.nn.Linear(2048, 512).to(_dev())
    lin2 = torch.nn.Linear(512, 10).to(_dev())
    labels = torch.randint(0, 10, (37,), device=_dev())
    h = F.relu(F.linear(x, lin1.weight, lin1.bias))
    logp = F.log_softmax(F.linear(h, lin2.weight, lin2.bias), 1)
    loss = F.nll_loss(logp, labels)
    loss.backward()
    got = [x.grad.clone(), lin1.weight.grad.clone(), lin1.bias.grad.clone(), lin2.weight.grad.clone(), lin2.bias.grad.clone()]
    xd = x.detach().double().requires_grad_(True)
    l1, l2 = torch.nn.Linear(2048, 512).to(_dev()).double(), torch.nn.Linear(512, 10).to(_dev()).double()
    l1.load_state_dict({k: v.double() for k, v in lin1.state_dict().items()})
    l2.load_state_dict({k: v.double() for k, v in lin2.state_dict().items()})
    lr = TF.log_softmax(l2(torch.relu(l1(xd))), 1)
    lossr = TF.nll_loss(lr, labels)
    lossr.backward()
    assert _rel(logp, lr) < 1e-5 and abs(loss.item() - lossr.item()) < 1e-5
    for a, b in zip(got, [xd.grad, l1.weight.grad, l1.bias.grad, l2.weight.grad, l2.bias.grad]):
        assert _rel(a, b) < 1e-4


def test_dropout_mask_replay_and_rate():
    import b200net
    F = b200net.functional
    x = torch.ones(1 << 20, device=_dev(), requires_grad=True)
    y = F.dropout(x, 0.2, True)
    keep = (y > 0).float().mean().item()
    assert abs(keep - 0.8) < 5e-3
    assert torch.allclose(y[y > 0], torch.tensor(1.25, device=_dev()))
    y.sum().backward()
    assert torch.equal(x.grad > 0, y > 0)          # backward replays the forward mask
    y2 = F.dropout(x, 0.2, True)
    assert not torch.equal(y2 > 0, y > 0)          # the Philox offset advances between calls
    assert F.dropout(x, 0.2, False) is x           # eval mode: identity


# ---------------------------------------------------------------------------------------------- optimizers
def _param_set(seed):
    g = torch.Generator().manual_seed(seed)
    shapes = [(64, 3, 7, 7), (64,), (256, 64, 1, 1), (2048, 512), (10,), (1,)] + [(17,)] * 60
    return [torch.randn(s, generator=g) for s in shapes]


@pytest.mark.parametrize("name,kw", [
    ("SGD", dict(lr=0.001)),
    ("SGD", dict(lr=0.05, momentum=0.9, weight_decay=1e-4)),
    ("SGD", dict(lr=0.05, momentum=0.9, nesterov=True)),
    ("Adam", dict(lr=0.003)),
    ("Adam", dict(lr=0.003, weight_decay=1e-2)),
    ("AdamW", dict(lr=2e-5, eps=1e-8, weight_decay=0.0)),
    ("AdamW", dict(lr=1e-3, weight_decay=0.1)),
])
def test_optimizers_match_torch(name, kw):
    import b200net
    ours = [torch.nn.Parameter(t.clone().to(_dev())) for t in _param_set(61)]
    ref = [torch.nn.Parameter(t.clone().double().to(_dev())) for t in _param_set(61)]
    o1 = getattr(b200net.optim, name)(ours, **kw)
    o2 = getattr(torch.optim, name)(ref, **kw)
    for step in range(4):
        grads = _param_set(100 + step)
        for p, q, g in zip(ours, ref, grads):
            p.grad = g.to(_dev())
            q.grad = g.double().to(_dev())
        v0 = ours[0]._version
        o1.step()
        o2.step()
        assert ours[0]._version > v0          # packed-weight caches key on the version counter
    for p, q in zip(ours, ref):
        assert _rel(p, q) < 1e-5


def test_transformers_adamw_rule():
    """bert.AdamW follows the update rule transformers.AdamW published (the class the reference imports at
    pytorch_on_language_distr.py:151; transformers 5.x no longer ships it): eps is added to sqrt(v) BEFORE the bias
    correction.  Checked against a float64 restatement of that rule; torch.optim.AdamW differs measurably at these
    settings, which the last assertion shows."""
    from b200net import bert
    lr, b1, b2, eps = 1e-3, 0.9, 0.999, 1e-3          # large eps so the two rules separate clearly
    ours = [torch.nn.Parameter(t.clone().to(_dev())) for t in _param_set(81)]
    ref = [t.clone().double().to(_dev()) for t in _param_set(81)]
    m = [torch.zeros_like(t) for t in ref]
    v = [torch.zeros_like(t) for t in ref]
    stock = [torch.nn.Parameter(t.clone().double().to(_dev())) for t in _param_set(81)]
    o1 = bert.AdamW(ours, lr=lr, eps=eps)
    o2 = torch.optim.AdamW(stock, lr=lr, eps=eps, weight_decay=0.0)
    for step in range(1, 5):
        grads = _param_set(200 + step)
        for i, g in enumerate(grads):
            gd = g.double().to(_dev())
            ours[i].grad = g.to(_dev())
            stock[i].grad = gd
            m[i].mul_(b1).add_(gd, alpha=1 - b1)
            v[i].mul_(b2).addcmul_(gd, gd, value=1 - b2)
            step_size = lr * math.sqrt(1 - b2 ** step) / (1 - b1 ** step)
            ref[i].addcdiv_(m[i], v[i].sqrt().add_(eps), value=-step_size)
        o1.step()
        o2.step()
    for p, q in zip(ours, ref):
        assert _rel(p, q) < 1e-5
    assert max(_rel(p, q) for p, q in zip(stock, ref)) > 1e-4


def test_clip_grad_norm():
    import b200net
    ours = [torch.nn.Parameter(t.clone().to(_dev())) for t in _param_set(71)]
    ref = [torch.nn.Parameter(t.clone().to(_dev())) for t in _param_set(71)]
    for p, q, g in zip(ours, ref, _param_set(72)):
        p.grad = g.clone().to(_dev())
        q.grad = g.clone().to(_dev())
    n1 = b200net.utils.clip_grad_norm_(ours, 1.0)
    n2 = torch.nn.utils.clip_grad_norm_(ref, 1.0)
    assert abs(n1.item() - n2.item()) < 1e-4 * n2.item()
    for p, q in zip(ours, ref):
        assert _rel(p.grad, q.grad) < 1e-5
    n3 = b200net.utils.clip_grad_norm_(ours, 1e9)     # below the threshold: untouched
    assert abs(n3.item() - 1.0) < 1e-3


@pytest.mark.parametrize("env", [{"B200_2SM": "1"}, {"B200_DUAL_MIN_KB": "1"}])
def test_two_tile_kernel_variants(env):
    """The two-pixel-tile variants of the fprop / dgrad kernel -- CTA pairs (tcgen05 cta_group::2, opt-in) and the
    256 x 256 single-CTA tile (chosen by reduction depth; forced for every shape here).  The switches are read once
    per process, so the conv / Linear parity tests are re-run in a child process."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", "-x", "-q", "-p", "no:cacheprovider",
                        os.path.join(root, "tests", "test_kernels_gpu.py"), os.path.join(root, "tests", "test_bert_gpu.py"),
                        "-k", "conv or dense"], env=dict(os.environ, **env), cwd=root, capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout

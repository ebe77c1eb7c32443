"""Tensor-level wrappers over the C-ABI: torch owns every buffer (caching allocator), the library only sees
device pointers, sizes and the current stream.  Internal activations are NHWC tensors ([N,H,W,C], fp32 or bf16).

Nothing here falls back to ATen/cuDNN/cuBLAS: a CPU tensor or a missing library is an error.
"""
import torch

from . import _C
from ._C import lib, check, ConvDesc
import ctypes

_DT = {torch.float32: _C.F32, torch.bfloat16: _C.BF16}


def _dt(t):
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError("b200net: unsupported dtype %s (float32 / bfloat16 only)" % t.dtype)


def _need_cuda(t):
    if not t.is_cuda:
        raise RuntimeError("b200net kernels run on CUDA tensors only (no CPU fallback); got a %s tensor" % t.device)


def _stream():
    return torch.cuda.current_stream().cuda_stream


# ------------------------------------------------------------------ workspace (one growing buffer per device)
_WS = {}
NO_WORKSPACE_GROWTH = [False]   # set while a CUDA graph is being captured
WORKSPACE_SLOT = [0]            # 0 = the main stream's buffer; 1 = the side stream's (kernels that overlap with the
                                # main stream -- the weight gradients -- must not share scratch with it)


class side_workspace(object):
    """Context: scratch requests inside it come from the side stream's own buffer."""

    def __enter__(self):
        self.prev = WORKSPACE_SLOT[0]
        WORKSPACE_SLOT[0] = 1

    def __exit__(self, *a):
        WORKSPACE_SLOT[0] = self.prev


def workspace(nbytes, device):
    key = (device.index if device.index is not None else torch.cuda.current_device(), WORKSPACE_SLOT[0])
    buf = _WS.get(key)
    if buf is None or buf.numel() < nbytes:
        if NO_WORKSPACE_GROWTH[0]:
            raise RuntimeError("b200net: workspace would have to grow during CUDA-graph capture")
        # stream-ordered reuse: every consumer runs on the current stream
        buf = torch.empty(max(int(nbytes), 1 << 20), dtype=torch.uint8, device=device)
        _WS[key] = buf
    return buf


# ------------------------------------------------------------------ layout
def to_nhwc(x, dtype, cpad=None):
    """NCHW fp32 (what the reference's loaders produce, ...on_image.ipynb:667) -> NHWC `dtype`.
    cpad widens the channel dimension with zeros (3-channel images -> 8-byte bf16 pixels for the stem)."""
    _need_cuda(x)
    if x.dtype != torch.float32:
        x = x.float()
    x = x.contiguous()
    N, C, H, W = x.shape
    cp = C if cpad is None else cpad
    y = torch.empty((N, H, W, cp), dtype=dtype, device=x.device)
    check(lib.b200_nchw_to_nhwc(x.data_ptr(), y.data_ptr(), N, C, H, W, cp, _DT[dtype], _stream()), "nchw_to_nhwc")
    return y


def to_nchw(x):
    _need_cuda(x)
    N, H, W, C = x.shape
    y = torch.empty((N, C, H, W), dtype=torch.float32, device=x.device)
    check(lib.b200_nhwc_to_nchw(x.data_ptr(), y.data_ptr(), N, C, H, W, _dt(x), _stream()), "nhwc_to_nchw")
    return y


def weight_ld(R, S, C):
    return (R * S * C + 7) // 8 * 8


def pack_weight(w, dtype, cpad=None):
    """OIHW fp32 parameter -> [K][ld] `dtype`, (r,s,c) column order; cpad >= C zero-pads the channel dimension."""
    _need_cuda(w)
    K, Cw, R, S = w.shape
    C = Cw if cpad is None else cpad
    ld = weight_ld(R, S, C)
    out = torch.empty((K, ld), dtype=dtype, device=w.device)
    wd = w.detach()
    if wd.dtype != torch.float32 or not wd.is_contiguous():
        wd = wd.float().contiguous()
    check(lib.b200_pack_weight(wd.data_ptr(), out.data_ptr(), K, Cw, R, S, C, ld, _DT[dtype], _stream()),
          "pack_weight")
    return out


def pack_weights_multi(ws, dtype):
    """pack_weight for a list of OIHW fp32 parameters (channels not padded) in one launch per 48 tensors."""
    outs, meta = [], []
    for w in ws:
        K, C, R, S = w.shape
        ld = weight_ld(R, S, C)
        outs.append(torch.empty((K, ld), dtype=dtype, device=w.device))
        meta.append((K, C, C, R * S, ld))
    if not ws:
        return outs
    srcs = []
    for w in ws:
        wd = w.detach()
        if wd.dtype != torch.float32 or not wd.is_contiguous():
            wd = wd.float().contiguous()
        srcs.append(wd)
    check(lib.b200_pack_weight_multi(len(ws), _C.ptr_array([t.data_ptr() for t in srcs]),
                                     _C.ptr_array([t.data_ptr() for t in outs]), _C.i32_array([m[0] for m in meta]),
                                     _C.i32_array([m[1] for m in meta]), _C.i32_array([m[2] for m in meta]),
                                     _C.i32_array([m[3] for m in meta]), _C.i32_array([m[4] for m in meta]),
                                     _DT[dtype], _stream()), "pack_weight_multi")
    return outs


def cast_rows_multi(pairs, dtype):
    """[(fp32 source [rows][cols] or [n], destination of the same shape in `dtype`)]: all copies / casts in one launch
    per 48 tensors (b200_pack_weight_multi with R*S = 1: a matrix is a 1x1 convolution weight, a vector one row)."""
    if not pairs:
        return
    K = [int(s.shape[0]) if s.dim() == 2 else 1 for s, _ in pairs]
    C = [int(s.shape[-1]) for s, _ in pairs]
    check(lib.b200_pack_weight_multi(len(pairs), _C.ptr_array([s.data_ptr() for s, _ in pairs]),
                                     _C.ptr_array([d.data_ptr() for _, d in pairs]), _C.i32_array(K), _C.i32_array(C),
                                     _C.i32_array(C), _C.i32_array([1] * len(pairs)), _C.i32_array(C), _DT[dtype],
                                     _stream()), "pack_weight_multi")


def im2col(x, R, S, stride, pad, ld):
    """NHWC [N,H,W,C] -> staged matrix [N,Ho,Wo,ld] (column (r*S+s)*C+c, zero tail)."""
    N, H, W, C = x.shape
    Ho, Wo = (H + 2 * pad - R) // stride + 1, (W + 2 * pad - S) // stride + 1
    col = torch.empty((N, Ho, Wo, ld), dtype=x.dtype, device=x.device)
    check(lib.b200_im2col(x.data_ptr(), col.data_ptr(), N, H, W, C, R, S, stride, pad, ld, _dt(x), _stream()), "im2col")
    return col


def stem_stage(x4, R, S, pad):
    """[N,H,W,4] bf16 -> [N, Ho + R' - 1, Wo, 64] row-pair window tensor of a stride-2 stem (see b200_stem_stage)."""
    N, H, W, C = x4.shape
    assert C == 4 and x4.dtype == torch.bfloat16
    Ho, Wo = (H + 2 * pad - R) // 2 + 1, (W + 2 * pad - S) // 2 + 1
    out = torch.empty((N, Ho + (R + 1) // 2 - 1, Wo, 64), dtype=torch.bfloat16, device=x4.device)
    check(lib.b200_stem_stage(x4.data_ptr(), out.data_ptr(), N, H, W, R, S, pad, _stream()), "stem_stage")
    return out


def stem_stage_nchw(x, R, S, pad):
    """[N,C<=4,H,W] fp32 NCHW -> the row-pair window tensor of stem_stage, without the NHWC4 intermediate."""
    _need_cuda(x)
    if x.dtype != torch.float32:
        x = x.float()
    x = x.contiguous()
    N, C, H, W = x.shape
    Ho, Wo = (H + 2 * pad - R) // 2 + 1, (W + 2 * pad - S) // 2 + 1
    out = torch.empty((N, Ho + (R + 1) // 2 - 1, Wo, 64), dtype=torch.bfloat16, device=x.device)
    check(lib.b200_stem_stage_nchw(x.data_ptr(), out.data_ptr(), N, C, H, W, R, S, pad, _stream()), "stem_stage_nchw")
    return out


def stem_pack_weight(w):
    """OIHW fp32 stem filter -> [K][R'*64] bf16 in the staged column order."""
    K, Cw, R, S = w.shape
    out = torch.empty((K, ((R + 1) // 2) * 64), dtype=torch.bfloat16, device=w.device)
    wd = w.detach()
    if wd.dtype != torch.float32 or not wd.is_contiguous():
        wd = wd.float().contiguous()
    check(lib.b200_stem_pack_weight(wd.data_ptr(), out.data_ptr(), K, Cw, R, S, _stream()), "stem_pack_weight")
    return out


def pack_weight_ld(w, dtype, cpad, ld):
    """pack_weight with an explicit row stride (the stem's staged-1x1 form needs ld = 256)."""
    K, Cw, R, S = w.shape
    out = torch.empty((K, ld), dtype=dtype, device=w.device)
    wd = w.detach()
    if wd.dtype != torch.float32 or not wd.is_contiguous():
        wd = wd.float().contiguous()
    check(lib.b200_pack_weight(wd.data_ptr(), out.data_ptr(), K, Cw, R, S, cpad, ld, _DT[dtype], _stream()),
          "pack_weight")
    return out


# ------------------------------------------------------------------ convolution
_DESC = {}


def _desc(N, H, W, C, K, R, S, stride, pad, dt):
    key = (N, H, W, C, K, R, S, stride, pad, dt)
    e = _DESC.get(key)
    if e is None:
        d = ConvDesc(N, H, W, C, K, R, S, stride, pad, dt)
        ws = [int(lib.b200_conv_workspace_bytes(ctypes.byref(d), which)) for which in range(3)]
        if min(ws) < 0:
            raise RuntimeError("b200net conv: unsupported descriptor %s: %s" % (key, _C.last_error()))
        Ho = (H + 2 * pad - R) // stride + 1
        Wo = (W + 2 * pad - S) // stride + 1
        e = (d, ctypes.byref(d), ws, Ho, Wo)
        _DESC[key] = e
    return e


def conv_fprop(x, wp, K, R, S, stride, pad):
    N, H, W, C = x.shape
    d, dref, ws, Ho, Wo = _desc(N, H, W, C, K, R, S, stride, pad, _dt(x))
    y = torch.empty((N, Ho, Wo, K), dtype=x.dtype, device=x.device)
    w = workspace(ws[0], x.device) if ws[0] else None
    check(lib.b200_conv_fprop(dref, x.data_ptr(), wp.data_ptr(), y.data_ptr(), w.data_ptr() if w is not None else None,
                              ws[0], _stream()), "conv_fprop")
    return y


_STATS_ROWS = {}


def conv_fprop_bnstats(x, wp, K, R, S, stride, pad):
    """conv forward + per-pixel-tile column statistics of the output; returns (y, partials or None).
    partials is None when the shape is not on the fused path (caller then runs bn_train_stats on y)."""
    N, H, W, C = x.shape
    key = (N, H, W, C, K, R, S, stride, pad, _dt(x))
    d, dref, ws, Ho, Wo = _desc(*key)
    rows = _STATS_ROWS.get(key)
    if rows is None:
        rows = int(lib.b200_conv_fprop_bnstats_rows(dref))
        _STATS_ROWS[key] = rows
    if rows <= 0:
        return conv_fprop(x, wp, K, R, S, stride, pad), None
    y = torch.empty((N, Ho, Wo, K), dtype=x.dtype, device=x.device)
    part = torch.empty((rows, 2, K), dtype=torch.float32, device=x.device)
    check(lib.b200_conv_fprop_bnstats(dref, x.data_ptr(), wp.data_ptr(), y.data_ptr(), part.data_ptr(), rows,
                                      _stream()), "conv_fprop_bnstats")
    return y, part


def conv_fprop_affine(x, wp, K, R, S, stride, pad, scale, shift, relu, residual=None):
    """conv forward with y = act(conv * scale + shift (+ residual)) applied in the epilogue (eval-mode BN folded in).
    Returns None when the shape is not on the fused path."""
    N, H, W, C = x.shape
    key = (N, H, W, C, K, R, S, stride, pad, _dt(x))
    d, dref, ws, Ho, Wo = _desc(*key)
    rows = _STATS_ROWS.get(key)
    if rows is None:
        rows = int(lib.b200_conv_fprop_bnstats_rows(dref))
        _STATS_ROWS[key] = rows
    if rows <= 0:
        return None
    y = torch.empty((N, Ho, Wo, K), dtype=x.dtype, device=x.device)
    check(lib.b200_conv_fprop_affine(dref, x.data_ptr(), wp.data_ptr(), y.data_ptr(), scale.data_ptr(),
                                     shift.data_ptr(), residual.data_ptr() if residual is not None else None,
                                     1 if relu else 0, _stream()), "conv_fprop_affine")
    return y


def bn_finalize_stats(part, rows, gamma, beta, running_mean, running_var, momentum, eps):
    nblk, _, C = part.shape
    st = torch.empty((4, C), dtype=torch.float32, device=part.device)
    check(lib.b200_bn_finalize_stats(
        part.data_ptr(), nblk, rows, C, gamma.data_ptr() if gamma is not None else None,
        beta.data_ptr() if beta is not None else None,
        running_mean.data_ptr() if running_mean is not None else None,
        running_var.data_ptr() if running_var is not None else None,
        float(momentum), float(eps), st[0].data_ptr(), st[1].data_ptr(), st[2].data_ptr(), st[3].data_ptr(),
        _stream()), "bn_finalize_stats")
    return st[0], st[1], st[2], st[3]


def conv_dgrad(dy, wp, x_shape, K, R, S, stride, pad, out=None, accumulate=False):
    N, H, W, C = x_shape
    d, dref, ws, Ho, Wo = _desc(N, H, W, C, K, R, S, stride, pad, _dt(dy))
    if out is None:
        out = torch.empty((N, H, W, C), dtype=dy.dtype, device=dy.device)
        accumulate = False
    w = workspace(ws[1], dy.device) if ws[1] else None
    check(lib.b200_conv_dgrad(dref, dy.data_ptr(), wp.data_ptr(), out.data_ptr(), 1 if accumulate else 0,
                              w.data_ptr() if w is not None else None, ws[1], _stream()), "conv_dgrad")
    return out


def conv_wgrad(x, dy, K, R, S, stride, pad, own_ws=None):
    """own_ws: a list -> the call gets a workspace of its own, appended to the list (deferred reductions: the slabs
    must survive until wgrad_flush)."""
    N, H, W, C = x.shape
    d, dref, ws, Ho, Wo = _desc(N, H, W, C, K, R, S, stride, pad, _dt(x))
    dw = torch.empty((K, C, R, S), dtype=torch.float32, device=x.device)
    if own_ws is not None:
        w = torch.empty(max(int(ws[2]), 16), dtype=torch.uint8, device=x.device)
        own_ws.append(w)
    else:
        w = workspace(ws[2], x.device)
    check(lib.b200_conv_wgrad(dref, x.data_ptr(), dy.data_ptr(), dw.data_ptr(), w.data_ptr(), ws[2], _stream()),
          "conv_wgrad")
    return dw


def wgrad_defer(on):
    check(lib.b200_wgrad_defer(1 if on else 0), "wgrad_defer")


def wgrad_flush():
    check(lib.b200_wgrad_flush(_stream()), "wgrad_flush")


# ------------------------------------------------------------------ batch norm
def _bn_ws(rows, C, device):
    n = int(lib.b200_bn_workspace_bytes(rows, C))
    return workspace(n, device), n


def bn_train_stats(x, gamma, beta, running_mean, running_var, momentum, eps):
    """x: [..., C] NHWC. Returns (mean, invstd, scale, shift), all fp32 [C]; updates running stats in place."""
    C = x.shape[-1]
    rows = x.numel() // C
    st = torch.empty((4, C), dtype=torch.float32, device=x.device)
    ws, n = _bn_ws(rows, C, x.device)
    check(lib.b200_bn_train_stats(
        x.data_ptr(), rows, C, _dt(x), gamma.data_ptr() if gamma is not None else None,
        beta.data_ptr() if beta is not None else None,
        running_mean.data_ptr() if running_mean is not None else None,
        running_var.data_ptr() if running_var is not None else None,
        float(momentum), float(eps), st[0].data_ptr(), st[1].data_ptr(), st[2].data_ptr(), st[3].data_ptr(),
        ws.data_ptr(), n, _stream()), "bn_train_stats")
    return st[0], st[1], st[2], st[3]


def bn_eval_coeffs(gamma, beta, running_mean, running_var, eps):
    C = running_mean.numel()
    st = torch.empty((2, C), dtype=torch.float32, device=running_mean.device)
    check(lib.b200_bn_eval_coeffs(gamma.data_ptr() if gamma is not None else None,
                                  beta.data_ptr() if beta is not None else None, running_mean.data_ptr(),
                                  running_var.data_ptr(), float(eps), st[0].data_ptr(), st[1].data_ptr(), C, _stream()),
          "bn_eval_coeffs")
    return st[0], st[1]


def bn_apply(x, scale, shift, relu, residual=None, out=None, want_mask=False, reverse=False):
    """y = act(x*scale + shift (+residual)).  With want_mask the ReLU mask of the stored output is returned too
    (one byte per 16-byte vector; None when C does not allow it) -- the backward kernels then never re-read y."""
    C = x.shape[-1]
    rows = x.numel() // C
    y = torch.empty_like(x) if out is None else out
    mask = None
    if want_mask and relu:
        nb = int(lib.b200_bn_mask_bytes(rows, C, _dt(x)))
        if nb > 0:
            mask = torch.empty(nb, dtype=torch.uint8, device=x.device)
    check(lib.b200_bn_apply(x.data_ptr(), residual.data_ptr() if residual is not None else None, y.data_ptr(),
                            mask.data_ptr() if mask is not None else None, rows, C, _dt(x), scale.data_ptr(),
                            shift.data_ptr(), 1 if relu else 0, 1 if reverse else 0, _stream()), "bn_apply")
    return (y, mask) if want_mask else y


def bn_backward(dz, z, x, gamma, mean, invstd, relu, want_dres=False, want_param_grads=True, mask=None, walk=0):
    C = x.shape[-1]
    rows = x.numel() // C
    dx = torch.empty_like(x)
    dres = torch.empty_like(x) if want_dres else None
    pg = torch.empty((2, C), dtype=torch.float32, device=x.device) if want_param_grads else None
    ws, n = _bn_ws(rows, C, x.device)
    check(lib.b200_bn_backward(
        dz.data_ptr(), z.data_ptr() if (z is not None and mask is None) else None,
        mask.data_ptr() if mask is not None else None, x.data_ptr(), dx.data_ptr(),
        dres.data_ptr() if dres is not None else None, rows, C, _dt(x),
        gamma.data_ptr() if gamma is not None else None, mean.data_ptr(), invstd.data_ptr(), 1 if relu else 0,
        pg[0].data_ptr() if pg is not None else None, pg[1].data_ptr() if pg is not None else None,
        ws.data_ptr(), n, int(walk), _stream()), "bn_backward")
    if pg is None:
        return dx, dres, None, None
    return dx, dres, pg[0], pg[1]


def add(a, b):
    y = torch.empty_like(a)
    check(lib.b200_add(a.data_ptr(), b.data_ptr(), y.data_ptr(), a.numel(), _dt(a), _stream()), "add")
    return y


def relu_bwd(dz, z):
    dx = torch.empty_like(dz)
    check(lib.b200_relu_bwd(dz.data_ptr(), z.data_ptr(), dx.data_ptr(), dz.numel(), _dt(dz), _stream()), "relu_bwd")
    return dx


# ------------------------------------------------------------------ pooling
def maxpool_fwd(x, want_index):
    N, H, W, C = x.shape
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    y = torch.empty((N, Ho, Wo, C), dtype=x.dtype, device=x.device)
    idx = torch.empty((N, Ho, Wo, C), dtype=torch.uint8, device=x.device) if want_index else None
    check(lib.b200_maxpool3x3s2_fwd(x.data_ptr(), y.data_ptr(), idx.data_ptr() if idx is not None else None, N, H, W, C,
                                    _dt(x), _stream()), "maxpool_fwd")
    return y, idx


def maxpool_bwd(dy, idx, x_shape):
    N, H, W, C = x_shape
    dx = torch.empty((N, H, W, C), dtype=dy.dtype, device=dy.device)
    check(lib.b200_maxpool3x3s2_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C, _dt(dy), _stream()),
          "maxpool_bwd")
    return dx


def bn_relu_maxpool_fwd(y, scale, shift, want_index):
    """maxpool3x3s2(relu(y * scale + shift)) straight from the convolution output (b200_bn_relu_maxpool_fwd)."""
    N, H, W, C = y.shape
    Ho, Wo = (H - 1) // 2 + 1, (W - 1) // 2 + 1
    pooled = torch.empty((N, Ho, Wo, C), dtype=y.dtype, device=y.device)
    idx = torch.empty((N, Ho, Wo, C), dtype=torch.uint8, device=y.device) if want_index else None
    check(lib.b200_bn_relu_maxpool_fwd(y.data_ptr(), scale.data_ptr(), shift.data_ptr(), pooled.data_ptr(),
                                       idx.data_ptr() if idx is not None else None, N, H, W, C, _dt(y), _stream()),
          "bn_relu_maxpool_fwd")
    return pooled, idx


def maxpool_bn_bwd(dpool, idx, pooled, y, gamma, mean, invstd, want_param_grads=True):
    """Gradient of the convolution output y through BatchNorm (batch statistics) -> ReLU -> maxpool3x3s2, from the
    pooled gradient (b200_maxpool_bn_bwd).  Returns (dy, dgamma, dbeta)."""
    N, H, W, C = y.shape
    dy = torch.empty_like(y)
    pg = torch.empty((2, C), dtype=torch.float32, device=y.device) if want_param_grads else None
    ws, n = _bn_ws(N * H * W, C, y.device)
    check(lib.b200_maxpool_bn_bwd(dpool.data_ptr(), idx.data_ptr(), pooled.data_ptr(), y.data_ptr(), dy.data_ptr(),
                                  N, H, W, C, _dt(y), gamma.data_ptr() if gamma is not None else None, mean.data_ptr(),
                                  invstd.data_ptr(), pg[0].data_ptr() if pg is not None else None,
                                  pg[1].data_ptr() if pg is not None else None, ws.data_ptr(), n, _stream()),
          "maxpool_bn_bwd")
    if pg is None:
        return dy, None, None
    return dy, pg[0], pg[1]


def avgpool_fwd(x):
    N, H, W, C = x.shape
    y = torch.empty((N, C), dtype=torch.float32, device=x.device)
    check(lib.b200_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, H * W, C, _dt(x), _stream()), "avgpool_fwd")
    return y


def avgpool_bwd(dy, x_shape, dtype):
    N, H, W, C = x_shape
    dx = torch.empty((N, H, W, C), dtype=dtype, device=dy.device)
    check(lib.b200_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), N, H * W, C, _DT[dtype], _stream()), "avgpool_bwd")
    return dx


# ------------------------------------------------------------------ head (fp32 matrices)
def gemm_f32(A, B, M, N, K, lda, ldb, a_mn, b_mn, bias=None, relu=False, out=None, accumulate=False):
    C = torch.empty((M, N), dtype=torch.float32, device=A.device) if out is None else out
    check(lib.b200_gemm_f32(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, lda, ldb, N, 1 if a_mn else 0,
                            1 if b_mn else 0, bias.data_ptr() if bias is not None else None, 1 if relu else 0,
                            1 if accumulate else 0, _stream()), "gemm_f32")
    return C


def linear_fwd(x, w, b, relu=False):
    """y[M,N] = x[M,K] @ w[N,K]^T + b   (nn.Linear; head at ...on_image.ipynb:439-442)"""
    M, K = x.shape
    N = w.shape[0]
    return gemm_f32(x, w, M, N, K, K, K, False, False, bias=b, relu=relu)


def linear_bwd_input(dy, w):
    """dx[M,K] = dy[M,N] @ w[N,K]"""
    M, N = dy.shape
    K = w.shape[1]
    return gemm_f32(dy, w, M, K, N, N, K, False, True)


def linear_bwd_weight(dy, x):
    """dw[N,K] = dy[M,N]^T @ x[M,K]"""
    M, N = dy.shape
    K = x.shape[1]
    return gemm_f32(dy, x, N, K, M, N, K, True, True)


def colsum(dy):
    M, N = dy.shape
    db = torch.empty((N,), dtype=torch.float32, device=dy.device)
    check(lib.b200_colsum_f32(dy.data_ptr(), db.data_ptr(), M, N, _stream()), "colsum")
    return db


def relu_fwd(x):
    y = torch.empty_like(x)
    check(lib.b200_relu_fwd_f32(x.data_ptr(), y.data_ptr(), x.numel(), _stream()), "relu_fwd")
    return y


def dropout_fwd(x, p, seed, offset):
    y = torch.empty_like(x)
    mask = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    check(lib.b200_dropout_fwd(x.data_ptr(), y.data_ptr(), mask.data_ptr(), x.numel(), float(p), seed, offset,
                               _stream()), "dropout_fwd")
    return y, mask


def dropout_bwd(dy, mask, p):
    dx = torch.empty_like(dy)
    check(lib.b200_dropout_bwd(dy.data_ptr(), mask.data_ptr(), dx.data_ptr(), dy.numel(), float(p), _stream()),
          "dropout_bwd")
    return dx


def log_softmax_fwd(x):
    rows, cols = x.shape
    y = torch.empty_like(x)
    check(lib.b200_log_softmax_fwd(x.data_ptr(), y.data_ptr(), rows, cols, _stream()), "log_softmax_fwd")
    return y


def log_softmax_bwd(dy, y):
    rows, cols = y.shape
    dx = torch.empty_like(y)
    check(lib.b200_log_softmax_bwd(dy.data_ptr(), y.data_ptr(), dx.data_ptr(), rows, cols, _stream()),
          "log_softmax_bwd")
    return dx


def nll_loss_fwd(logp, labels):
    rows, cols = logp.shape
    loss = torch.empty((), dtype=torch.float32, device=logp.device)
    check(lib.b200_nll_loss_fwd(logp.data_ptr(), labels.data_ptr(), loss.data_ptr(), rows, cols, _stream()),
          "nll_loss_fwd")
    return loss


def nll_loss_bwd(dloss, labels, rows, cols):
    d = torch.empty((rows, cols), dtype=torch.float32, device=labels.device)
    check(lib.b200_nll_loss_bwd(dloss.data_ptr(), labels.data_ptr(), d.data_ptr(), rows, cols, _stream()),
          "nll_loss_bwd")
    return d


# ------------------------------------------------------------------ BERT encoder ops ([tokens][hidden] matrices)
def _p(t):
    return t.data_ptr() if t is not None else None


def embed_ln_fwd(ids, type_ids, word, pos, type_tab, gamma, beta, eps, p, seed, offset, dtype, status):
    """BertEmbeddings forward; returns (s, y, mean, rstd) with s the pre-LayerNorm sum (saved for backward)."""
    _need_cuda(ids)
    N, S = ids.shape
    V, H = word.shape
    s = torch.empty((N * S, H), dtype=dtype, device=ids.device)
    y = torch.empty_like(s)
    st = torch.empty((2, N * S), dtype=torch.float32, device=ids.device)
    check(lib.b200_embed_ln_fwd(ids.data_ptr(), _p(type_ids), word.data_ptr(), pos.data_ptr(), type_tab.data_ptr(),
                                gamma.data_ptr(), beta.data_ptr(), float(eps), N, S, H, V, pos.shape[0], float(p),
                                seed, offset, s.data_ptr(), y.data_ptr(), st[0].data_ptr(), st[1].data_ptr(),
                                status.data_ptr(), _DT[dtype], _stream()), "embed_ln_fwd")
    return s, y, st[0], st[1]


def embed_bwd(ids, type_ids, dsum, vocab, max_pos):
    N, S = ids.shape
    H = dsum.shape[-1]
    dword = torch.empty((vocab, H), dtype=torch.float32, device=dsum.device)
    dpos = torch.empty((max_pos, H), dtype=torch.float32, device=dsum.device)
    dtyp = torch.empty((2, H), dtype=torch.float32, device=dsum.device)
    n = int(lib.b200_colsum_workspace_bytes(N * S, H))
    ws = workspace(n, dsum.device)
    check(lib.b200_embed_bwd(ids.data_ptr(), _p(type_ids), dsum.data_ptr(), N, S, H, vocab, max_pos, dword.data_ptr(),
                             dpos.data_ptr(), dtyp.data_ptr(), ws.data_ptr(), n, _dt(dsum), _stream()), "embed_bwd")
    return dword, dpos, dtyp


def ln_fwd(x, residual, gamma, beta, eps, p=0.0, seed=0, offset=0, inplace_sum=True):
    """s = dropout(x) + residual ; y = LayerNorm(s).  Returns (s, y, mean, rstd); s overwrites x when inplace_sum."""
    rows, H = x.shape
    s = x if inplace_sum else torch.empty_like(x)
    y = torch.empty_like(x)
    st = torch.empty((2, rows), dtype=torch.float32, device=x.device)
    check(lib.b200_ln_fwd(x.data_ptr(), _p(residual), s.data_ptr(), y.data_ptr(), st[0].data_ptr(), st[1].data_ptr(),
                          gamma.data_ptr(), beta.data_ptr(), float(eps), rows, H, float(p), seed, offset, _dt(x),
                          _stream()), "ln_fwd")
    return s, y, st[0], st[1]


def ln_bwd(dy_a, dy_b, s, mean, rstd, gamma, p=0.0, seed=0, offset=0, drop_after=False, want_dbias=True):
    """Returns (dres, dx, dgamma, dbeta, dbias): dres = dL/ds, dx = gradient of the dropped branch (is dres when
    p == 0), dbias = column sums of dx."""
    rows, H = s.shape
    dres = torch.empty_like(s)
    dx = torch.empty_like(s) if (p > 0.0 and not drop_after) else None
    pg = torch.empty((3, H), dtype=torch.float32, device=s.device)
    n = int(lib.b200_ln_bwd_workspace_bytes(rows, H))
    ws = workspace(n, s.device)
    check(lib.b200_ln_bwd(dy_a.data_ptr(), _p(dy_b), s.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gamma.data_ptr(),
                          rows, H, float(p), seed, offset, 1 if drop_after else 0, dres.data_ptr(), _p(dx),
                          pg[0].data_ptr(), pg[1].data_ptr(), pg[2].data_ptr() if want_dbias else None,
                          ws.data_ptr(), n, _dt(s), _stream()), "ln_bwd")
    return dres, (dx if dx is not None else dres), pg[0], pg[1], (pg[2] if want_dbias else None)


def gelu_fwd(x):
    y = torch.empty_like(x)
    check(lib.b200_gelu_fwd(x.data_ptr(), y.data_ptr(), x.numel(), _dt(x), _stream()), "gelu_fwd")
    return y


def gelu_bwd(dy, x, out=None):
    dx = torch.empty_like(x) if out is None else out
    check(lib.b200_gelu_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), x.numel(), _dt(x), _stream()), "gelu_bwd")
    return dx


def gelu_bwd_bias(dy, x, out=None):
    """(dx, dbias) = (dy * gelu'(x), column sums of dx) in one pass over [rows][cols]"""
    rows, cols = x.shape
    dx = torch.empty_like(x) if out is None else out
    db = torch.empty((cols,), dtype=torch.float32, device=x.device)
    n = int(lib.b200_colsum_workspace_bytes(rows, cols))
    ws = workspace(n, x.device)
    check(lib.b200_gelu_bwd_bias(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), db.data_ptr(), rows, cols, _dt(x),
                                 ws.data_ptr(), n, _stream()), "gelu_bwd_bias")
    return dx, db


def colsum_any(x):
    """fp32 column sums of a [rows][cols] matrix in either dtype (bias gradients)."""
    rows, cols = x.shape
    out = torch.empty((cols,), dtype=torch.float32, device=x.device)
    n = int(lib.b200_colsum_workspace_bytes(rows, cols))
    ws = workspace(n, x.device)
    check(lib.b200_colsum(x.data_ptr(), out.data_ptr(), rows, cols, _dt(x), ws.data_ptr(), n, _stream()), "colsum")
    return out


def attention_fwd(qkv, mask, N, S, heads, p=0.0, seed=0, offset=0):
    """qkv [N*S][3*heads*64] -> (probs [N,heads,S,S] (pre-dropout, saved), ctx [N*S][heads*64])"""
    D = qkv.shape[1] // (3 * heads)
    probs = torch.empty((N, heads, S, S), dtype=qkv.dtype, device=qkv.device)
    ctx = torch.empty((N * S, heads * D), dtype=qkv.dtype, device=qkv.device)
    check(lib.b200_attention_fwd(qkv.data_ptr(), _p(mask), probs.data_ptr(), ctx.data_ptr(), N, S, heads, D, float(p),
                                 seed, offset, _dt(qkv), _stream()), "attention_fwd")
    return probs, ctx


def attention_bwd(qkv, probs, dctx, N, S, heads, p=0.0, seed=0, offset=0):
    D = qkv.shape[1] // (3 * heads)
    dqkv = torch.empty_like(qkv)
    ws = workspace(probs.numel() * probs.element_size(), qkv.device)
    check(lib.b200_attention_bwd(qkv.data_ptr(), probs.data_ptr(), dctx.data_ptr(), ws.data_ptr(), dqkv.data_ptr(),
                                 N, S, heads, D, float(p), seed, offset, _dt(qkv), _stream()), "attention_bwd")
    return dqkv


def attention_flash_ok(dtype, S, D=64):
    return dtype == torch.bfloat16 and bool(lib.b200_attention_flash_supported(S, D))


def attention_flash_fwd(qkv, mask, N, S, heads, p=0.0, seed=0, offset=0):
    """Fused tcgen05 attention: returns ((lse [N,heads,S] fp32, keep bits or None), ctx [N*S][heads*64]); no S x S
    tensor of probabilities is stored."""
    D = qkv.shape[1] // (3 * heads)
    lse = torch.empty((N, heads, S), dtype=torch.float32, device=qkv.device)
    bits = torch.empty((N, heads, S, S // 32), dtype=torch.int32, device=qkv.device) if p > 0.0 else None
    ctx = torch.empty((N * S, heads * D), dtype=qkv.dtype, device=qkv.device)
    check(lib.b200_attention_flash_fwd(qkv.data_ptr(), _p(mask), lse.data_ptr(), _p(bits), ctx.data_ptr(), N, S,
                                       heads, D, float(p), seed, offset, _stream()), "attention_flash_fwd")
    return (lse, bits), ctx


def attention_flash_bwd(qkv, mask, saved, ctx, dctx, N, S, heads, p=0.0):
    lse, bits = saved
    D = qkv.shape[1] // (3 * heads)
    dqkv = torch.empty_like(qkv)
    dvec = torch.empty_like(lse)
    check(lib.b200_attention_flash_bwd(qkv.data_ptr(), _p(mask), lse.data_ptr(), _p(bits), ctx.data_ptr(),
                                       dctx.data_ptr(), dvec.data_ptr(), dqkv.data_ptr(), N, S, heads, D, float(p),
                                       _stream()), "attention_flash_bwd")
    return dqkv


def tanh_fwd(x):
    y = torch.empty_like(x)
    check(lib.b200_tanh_fwd_f32(x.data_ptr(), y.data_ptr(), x.numel(), _stream()), "tanh_fwd")
    return y


def tanh_bwd(dy, y):
    dx = torch.empty_like(y)
    check(lib.b200_tanh_bwd_f32(dy.data_ptr(), y.data_ptr(), dx.data_ptr(), y.numel(), _stream()), "tanh_bwd")
    return dx


def gather_rows(x, N, row_stride):
    """out[n] = x[n*row_stride] as fp32 (the [CLS] row of every sequence)"""
    H = x.shape[-1]
    out = torch.empty((N, H), dtype=torch.float32, device=x.device)
    check(lib.b200_gather_rows(x.data_ptr(), out.data_ptr(), N, row_stride * H, H, _dt(x), _stream()), "gather_rows")
    return out


def scatter_rows(d, rows, row_stride, dtype):
    """zeros([rows][H]) with row n*row_stride = d[n]"""
    N, H = d.shape
    out = torch.zeros((rows, H), dtype=dtype, device=d.device)
    check(lib.b200_scatter_rows(d.data_ptr(), out.data_ptr(), N, row_stride * H, H, _DT[dtype], _stream()),
          "scatter_rows")
    return out


# Linear layers of the encoder: y[M][out] = x[M][in] @ w[out][in]^T + b.  bf16: flat 1x1 problems on the tcgen05
# implicit-GEMM kernel (bias added in its epilogue); fp32: the FFMA GEMM.
_ONES = {}


def _ones(n, device):
    key = (n, device)
    t = _ONES.get(key)
    if t is None:
        t = torch.ones(n, dtype=torch.float32, device=device)
        _ONES[key] = t
    return t


def dense_fwd(x, wp, bias, relu=False):
    """wp: packed [out][in] operand in x.dtype (for fp32 this is the parameter itself)."""
    M, K = x.shape
    N = wp.shape[0]
    if x.dtype == torch.float32:
        return gemm_f32(x, wp, M, N, K, K, K, False, False, bias=bias, relu=relu)
    y = conv_fprop_affine(x.view(1, 1, M, K), wp, N, 1, 1, 1, 0, _ones(N, x.device), bias, relu)
    if y is None:
        raise RuntimeError("b200net dense_fwd: shape [%d,%d]x[%d] is not on the tensor-core path" % (M, K, N))
    return y.view(M, N)


def dense_bwd_input(dy, wp, K, out=None, accumulate=False):
    """dx[M][in] (+)= dy[M][out] @ w[out][in]"""
    M, N = dy.shape
    if dy.dtype == torch.float32:
        return gemm_f32(dy, wp, M, K, N, N, K, False, True, out=out, accumulate=accumulate)
    o = None if out is None else out.view(1, 1, M, K)
    return conv_dgrad(dy.view(1, 1, M, N), wp, (1, 1, M, K), N, 1, 1, 1, 0, out=o, accumulate=accumulate).view(M, K)


def dense_bwd_weight(dy, x):
    """dw[out][in] fp32 = dy[M][out]^T @ x[M][in]"""
    M, N = dy.shape
    K = x.shape[1]
    if dy.dtype == torch.float32:
        return gemm_f32(dy, x, N, K, M, N, K, True, True)
    return conv_wgrad(x.view(1, 1, M, K), dy.view(1, 1, M, N), N, 1, 1, 1, 0).view(N, K)


# ------------------------------------------------------------------ VGG16 extras
def maxpool2_fwd(x, want_index):
    N, H, W, C = x.shape
    y = torch.empty((N, H // 2, W // 2, C), dtype=x.dtype, device=x.device)
    idx = torch.empty((N, H // 2, W // 2, C), dtype=torch.uint8, device=x.device) if want_index else None
    check(lib.b200_maxpool2x2_fwd(x.data_ptr(), y.data_ptr(), _p(idx), N, H, W, C, _dt(x), _stream()), "maxpool2x2_fwd")
    return y, idx


def maxpool2_bwd(dy, idx, x_shape):
    N, H, W, C = x_shape
    dx = torch.empty((N, H, W, C), dtype=dy.dtype, device=dy.device)
    check(lib.b200_maxpool2x2_bwd(dy.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, H, W, C, _dt(dy), _stream()),
          "maxpool2x2_bwd")
    return dx


def adaptive_avgpool_fwd(x, oh, ow):
    N, H, W, C = x.shape
    y = torch.empty((N, oh, ow, C), dtype=x.dtype, device=x.device)
    check(lib.b200_adaptive_avgpool_fwd(x.data_ptr(), y.data_ptr(), N, H, W, C, oh, ow, _dt(x), _stream()),
          "adaptive_avgpool_fwd")
    return y


def adaptive_avgpool_bwd(dy, x_shape):
    N, H, W, C = x_shape
    dx = torch.empty((N, H, W, C), dtype=dy.dtype, device=dy.device)
    check(lib.b200_adaptive_avgpool_bwd(dy.data_ptr(), dx.data_ptr(), N, H, W, C, dy.shape[1], dy.shape[2], _dt(dy),
                                        _stream()), "adaptive_avgpool_bwd")
    return dx


def dropout_act(x, p, seed, offset, out=None):
    """y = dropout(x) with the Philox mask of (seed, offset); the same call on a gradient is the backward."""
    y = torch.empty_like(x) if out is None else out
    check(lib.b200_dropout_act(x.data_ptr(), y.data_ptr(), x.numel(), float(p), seed, offset, _dt(x), _stream()),
          "dropout_act")
    return y

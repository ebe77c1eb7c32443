"""Autograd glue: torch.autograd.Function wrappers whose forward/backward enqueue b200net kernels.

Two granularities:
  * `resnet_backbone(model, x)` -- the hot path: ONE autograd node for the whole torchvision-shaped backbone
    (stem -> layer1..4 -> avgpool).  Forward and backward are hand-scheduled lists of kernel launches over NHWC
    activations; nothing goes through ATen convolution / batch_norm.
    Mirrors ResNet._forward_impl as called from `model.forward(inputs)`
    (pytorch_training_inference_on_image.ipynb:669) and `loss.backward()` (:671).
  * per-layer functions (conv2d, batch_norm, linear, relu, dropout, log_softmax, nll_loss ...) used by the
    stand-alone modules in nn.py and by the classifier head.
"""
import os
import weakref

import torch
from torch.autograd.function import once_differentiable

from . import ops

# ------------------------------------------------------------------ precision switch
_PRECISION = os.environ.get("B200_PRECISION", "bf16").lower()


def set_precision(p):
    """'bf16' (tcgen05 tensor cores, fp32 accumulate, fp32 statistics/params) or 'fp32' (exact FFMA parity mode)."""
    global _PRECISION
    p = p.lower()
    if p not in ("bf16", "fp32"):
        raise ValueError("precision must be 'bf16' or 'fp32'")
    _PRECISION = p


def get_precision():
    return _PRECISION


def act_dtype():
    return torch.bfloat16 if _PRECISION == "bf16" else torch.float32


# ------------------------------------------------------------------ gradient sink (set by parallel.DistributedDataParallel)
# Called as sink(param, grad) the moment a parameter gradient has been enqueued inside the fused backbone backward;
# returns the tensor to hand to autograd (DDP returns the parameter's bucket slice).
_GRAD_SINK = None


def set_grad_sink(fn):
    global _GRAD_SINK
    _GRAD_SINK = fn


def grad_sink_owner():
    """The DistributedDataParallel wrapper behind the current sink when it can take part in graph replay."""
    owner = getattr(_GRAD_SINK, "__self__", None)
    return owner if hasattr(owner, "_sink_capture") else None


# ------------------------------------------------------------------ packed-weight cache
# key: id(param) -> (data_ptr, version, dtype, packed).  Our optimizers bump the parameter version after the raw
# in-place update (torch.autograd.graph.increment_version), stock optimizers do it through ATen.
_PACKED = {}
_CAPTURE_PACKED = {}   # packed buffers of the graph being captured (one pack per weight per graph)


def _pack(w, dtype, cpad, ld):
    if cpad == "stem":                  # row-pair window order of the staged stride-2 stem
        return ops.stem_pack_weight(w)
    return ops.pack_weight(w, dtype, cpad) if ld is None else ops.pack_weight_ld(w, dtype, cpad, ld)


def packed_weight(w, dtype, cpad=None, ld=None):
    key = id(w)
    if _CAPTURING[0]:
        # inside a graph capture the repack has to be recorded (the weights change between replays) and must not
        # leak into the eager cache (the buffer belongs to the graph's memory pool)
        cached = _CAPTURE_PACKED.get((key, dtype, cpad, ld))
        if cached is None:
            cached = _pack(w, dtype, cpad, ld)
            _CAPTURE_PACKED[(key, dtype, cpad, ld)] = cached
        return cached
    e = _PACKED.get(key)
    # id() values are recycled: a parameter of a NEW model can land on the id (and storage address, and version
    # count) of a dead one, so the entry also remembers which object it was made for
    if (e is not None and e[5]() is w and e[0] == w.data_ptr() and e[1] == w._version and e[2] == dtype
            and e[4] == (cpad, ld)):
        return e[3]
    p = _pack(w, dtype, cpad, ld)
    _PACKED[key] = (w.data_ptr(), w._version, dtype, p, (cpad, ld), weakref.ref(w, lambda _r, k=key: _PACKED.pop(k, None)))
    return p


def prepack_weights(convs, dtype):
    """Packs every convolution weight of `convs` whose packed copy is missing or stale in ONE multi-tensor launch (per
    48 weights) and files the results where packed_weight() looks them up.  The training step calls this at the top of
    the forward: all 53 weights changed in the optimizer step, and 53 tiny pack launches cost ~0.2 ms per step."""
    todo = []
    for conv in convs:
        w = conv.weight
        if w.dim() != 4 or not w.is_cuda:
            continue
        key = id(w)
        if _CAPTURING[0]:
            if (key, dtype, None, None) not in _CAPTURE_PACKED:
                todo.append(w)
        else:
            e = _PACKED.get(key)
            if not (e is not None and e[5]() is w and e[0] == w.data_ptr() and e[1] == w._version and e[2] == dtype
                    and e[4] == (None, None)):
                todo.append(w)
    if not todo:
        return
    for w, p in zip(todo, ops.pack_weights_multi(todo, dtype)):
        key = id(w)
        if _CAPTURING[0]:
            _CAPTURE_PACKED[(key, dtype, None, None)] = p
        else:
            _PACKED[key] = (w.data_ptr(), w._version, dtype, p, (None, None),
                            weakref.ref(w, lambda _r, k=key: _PACKED.pop(k, None)))


def _stem_cpad(conv, dtype):
    """3-channel images are widened to 4 channels (8-byte pixels) in bf16 mode so the stem's staging copies move
    8 bytes at a time; the fourth channel and its weights are zero."""
    return 4 if (dtype == torch.bfloat16 and conv.in_channels == 3) else None


def _conv_geom(conv):
    if conv.groups != 1 or conv.dilation not in ((1, 1), 1):
        raise NotImplementedError("b200net Conv2d: groups=1, dilation=1 only")
    k = conv.kernel_size
    s = conv.stride
    p = conv.padding
    if s[0] != s[1] or p[0] != p[1]:
        raise NotImplementedError("b200net Conv2d: square stride / padding only")
    if conv.bias is not None:
        raise NotImplementedError("b200net Conv2d: bias=False only (ResNet convolutions carry no bias)")
    return conv.out_channels, k[0], k[1], s[0], p[0]


# ------------------------------------------------------------------ BN helpers shared by both granularities
def _bn_uses_batch_stats(bn):
    return bn.training or (bn.running_mean is None and bn.running_var is None)


def _bn_coeffs(bn, y, bump_nbt=True, partials=None):
    """Returns (scale, shift, mean, invstd); mean/invstd are None in eval mode.  Updates running stats in train mode.
    `partials` = per-tile column sums already produced by the convolution's epilogue (ops.conv_fprop_bnstats)."""
    use_batch = _bn_uses_batch_stats(bn)
    if use_batch:
        momentum = 0.0
        rm, rv = bn.running_mean, bn.running_var
        if bn.training and bn.track_running_stats and rm is not None:
            if bump_nbt and bn.num_batches_tracked is not None:
                bn.num_batches_tracked.add_(1)
            momentum = bn.momentum if bn.momentum is not None else 1.0 / float(bn.num_batches_tracked)
        else:
            rm = rv = None
        if partials is not None:
            mean, invstd, scale, shift = ops.bn_finalize_stats(partials, y.numel() // y.shape[-1], bn.weight, bn.bias,
                                                               rm, rv, momentum, bn.eps)
        else:
            mean, invstd, scale, shift = ops.bn_train_stats(y, bn.weight, bn.bias, rm, rv, momentum, bn.eps)
        return scale, shift, mean, invstd
    scale, shift = ops.bn_eval_coeffs(bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
    return scale, shift, None, None


# ------------------------------------------------------------------ the backbone as one autograd node
class _Unit(object):
    """Saved state of one conv -> BN (-> +residual) (-> ReLU) unit."""
    __slots__ = ("conv", "bn", "x", "y", "z", "mask", "mean", "invstd", "relu", "geom", "staged", "pool")


# L2-aware walk order (B200_L2_WALK=1; OFF by default): a bandwidth kernel that starts where its producer finished
# should find that end of the tensor still in the 126 MB L2 -- the BatchNorm apply walks last -> first after a
# convolution that walked first -> last, the backward reduction pass walks against the dgrad that produced its input,
# the dx pass against the reduction pass.  Measured on the N=256 step: 13.90 ms with, 13.93 ms without -- no gain (the
# tensors that matter are several times larger than L2), so the kernels keep the option but nothing uses it.
_L2_WALK = os.environ.get("B200_L2_WALK") is not None
_STEM_PAIRS = os.environ.get("B200_STEM_IM2COL") is None
_FOLD_EVAL_BN = os.environ.get("B200_NO_FOLD_EVAL_BN") is None
_FOLD_EVAL_RESIDUAL = os.environ.get("B200_NO_FOLD_EVAL_RESIDUAL") is None
_WGRAD_DEFER = os.environ.get("B200_NO_WGRAD_DEFER") is None
_STEM_FUSED_STAGE = os.environ.get("B200_NO_FUSED_STEM_STAGE") is None
# stem: BatchNorm -> ReLU -> MaxPool as one pass each way (csrc/stem_pool.cu); B200_NO_FUSED_STEM_POOL=1 = separate kernels
_FUSE_STEM_POOL = os.environ.get("B200_NO_FUSED_STEM_POOL") is None


def _unit_forward(conv, bn, x, relu, residual, save, stage=False, pool=False, nchw_src=None):
    """conv -> BN (-> + residual) (-> ReLU); pool=True (the stem) appends MaxPool2d(3, 2, 1).  Returns (output, unit)."""
    K, R, S, stride, pad = _conv_geom(conv)
    staged = None
    if stage and stride == 2 and S <= 8 and _STEM_PAIRS and (nchw_src is not None or x.shape[-1] == 4):
        # Stem (3 -> 64, 7x7/2) in bf16 mode: the channel-padded image is staged ONCE as a row-pair window tensor
        # [N][Ho+3][Wo][64] (128 bytes per output column: the 8-pixel windows of two input rows; a quarter of an
        # im2col matrix) and the convolution becomes a 4x1 stride-1 convolution over 64 "channels" on the implicit
        # tcgen05 path -- filter-row pairs are TMA coordinate shifts, the four taps re-read the staged tensor from L2;
        # fused BN statistics as everywhere; the weight gradient reuses the staged tensor.
        wp = packed_weight(conv.weight, torch.bfloat16 if nchw_src is not None else x.dtype, "stem")
        staged = ("pairs", R, S)
        # nchw_src: the loaders' fp32 NCHW batch is staged directly (no NHWC4 intermediate)
        x = ops.stem_stage_nchw(nchw_src, R, S, pad) if nchw_src is not None else ops.stem_stage(x, R, S, pad)
        R, S, stride, pad = (R + 1) // 2, 1, 1, 0
    elif stage:
        # other narrow-input first layers: the image is staged as an im2col matrix (columns padded to a multiple of
        # 64) and the convolution becomes a 1x1 convolution over those columns
        cp = x.shape[-1]
        ld = (R * S * cp + 63) // 64 * 64
        wp = packed_weight(conv.weight, x.dtype, cp, ld)
        staged = (R, S, stride, pad, cp, x.shape)
        x = ops.im2col(x, R, S, stride, pad, ld)
        R, S, stride, pad = 1, 1, 1, 0
    else:
        wp = packed_weight(conv.weight, x.dtype, x.shape[-1] if x.shape[-1] != conv.in_channels else None)
    if _bn_uses_batch_stats(bn):
        y, partials = ops.conv_fprop_bnstats(x, wp, K, R, S, stride, pad)   # BN statistics fused into the epilogue
    else:
        if not save and _FOLD_EVAL_BN and (residual is None or _FOLD_EVAL_RESIDUAL):
            # eval mode: BatchNorm (running statistics), the residual add and ReLU folded into the conv epilogue.  The
            # residual tile comes in by TMA, prefetched one item ahead into the staging tile the output leaves from
            # (conv_sm100.cuh, res_tma): batch-1024 inference 103.2k -> 117.6k img/s.  With per-thread residual loads
            # the epilogue was LSU-bound (75.6k vs 84.0k img/s at the time), B200_RES_TMA=0 brings that path back.
            scale, shift = ops.bn_eval_coeffs(bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.eps)
            z = ops.conv_fprop_affine(x, wp, K, R, S, stride, pad, scale, shift, relu, residual)
            if z is not None:
                return (ops.maxpool_fwd(z, False)[0] if pool else z), None
        y, partials = ops.conv_fprop(x, wp, K, R, S, stride, pad), None
    scale, shift, mean, invstd = _bn_coeffs(bn, y, bump_nbt=False, partials=partials)
    fuse_pool = (pool and relu and residual is None and _FUSE_STEM_POOL
                 and y.shape[-1] % (8 if y.dtype == torch.bfloat16 else 4) == 0
                 and 256 % (y.shape[-1] // (8 if y.dtype == torch.bfloat16 else 4)) == 0)
    if not save:
        if fuse_pool:
            return ops.bn_relu_maxpool_fwd(y, scale, shift, False)[0], None
        # y is dead after the apply -> write in place
        z = ops.bn_apply(y, scale, shift, relu, residual, out=y, reverse=_L2_WALK)
        return (ops.maxpool_fwd(z, False)[0] if pool else z), None
    u = _Unit()
    u.conv, u.bn, u.x, u.y, u.mean, u.invstd, u.relu = conv, bn, x, y, mean, invstd, relu
    u.staged = staged
    u.geom = (K, R, S, stride, pad)
    u.pool = None
    if fuse_pool:
        # the ReLU output is never materialised: pooled values + argmax are all the backward needs besides y
        out, idx = ops.bn_relu_maxpool_fwd(y, scale, shift, True)
        u.z = u.mask = None
        u.pool = ("fused", idx, out)
        return out, u
    z, mask = ops.bn_apply(y, scale, shift, relu, residual, want_mask=True, reverse=_L2_WALK)
    u.z, u.mask = z, mask
    if pool:
        out, idx = ops.maxpool_fwd(z, True)
        u.pool = ("plain", idx, z.shape)
        return out, u
    return z, u


# Parameter lists are cached per model: walking 53 convolutions and 53 BatchNorms through nn.Module.__getattr__ costs
# ~0.17 ms per call, and with the reference's `loss.item()` every step the GPU idles while the host gets from one step's
# read-back to the next forward's graph launch (tools/e2e_probe.py: 11.91 ms without the read-back, 12.46 ms with).
# The cache is dropped whenever ANY parameter is (re-)registered on any module (torch's global registration hook, which
# `module.weight = nn.Parameter(...)` goes through) and is keyed on the same structure signature as _bn_modules; tensors
# re-allocated in place (`model.to(...)`) keep their Parameter objects and are caught by the graph's pointer check.
_PARAM_EPOCH = [0]
_PARAM_CACHE = {}
_STAT_CACHE = {}      # model -> its BatchNorm running-statistics tensors (same validity rule; buffers have their own hook)


def _param_registered(module, name, param):
    _PARAM_EPOCH[0] += 1
    return None


torch.nn.modules.module.register_module_parameter_registration_hook(_param_registered)
torch.nn.modules.module.register_module_buffer_registration_hook(_param_registered)


def _structure_sig(model):
    layers = tuple(getattr(model, n, None) for n in ("layer1", "layer2", "layer3", "layer4"))
    return (id(getattr(model, "bn1", None)), id(getattr(model, "fc", None))) + tuple(
        (id(l), len(l) if l is not None else 0) for l in layers)


def _backbone_params(model):
    sig = (_PARAM_EPOCH[0], _structure_sig(model))
    e = _PARAM_CACHE.get(id(model))
    if e is not None and e[0] == sig and e[2]() is model:
        return e[1]
    ps = _backbone_params_walk(model)
    if len(_PARAM_CACHE) > 64:
        _PARAM_CACHE.clear()
    _PARAM_CACHE[id(model)] = (sig, ps, weakref.ref(model))
    return ps


def _backbone_params_walk(model):
    ps = [model.conv1.weight, model.bn1.weight, model.bn1.bias]
    for layer in (model.layer1, model.layer2, model.layer3, model.layer4):
        for blk in layer:
            ps += [blk.conv1.weight, blk.bn1.weight, blk.bn1.bias, blk.conv2.weight, blk.bn2.weight, blk.bn2.bias,
                   blk.conv3.weight, blk.bn3.weight, blk.bn3.bias]
            if blk.downsample is not None:
                ps += [blk.downsample[0].weight, blk.downsample[1].weight, blk.downsample[1].bias]
    return ps


def _backbone_convs(model):
    """Every convolution of the backbone except the stem (whose packed form depends on how it is staged)."""
    cs = []
    for layer in (model.layer1, model.layer2, model.layer3, model.layer4):
        for blk in layer:
            cs += [blk.conv1, blk.conv2, blk.conv3]
            if blk.downsample is not None:
                cs.append(blk.downsample[0])
    return cs


def _under(fc, m, model):
    """True when BatchNorm module m belongs to the user's head rather than to the backbone."""
    return any(m is x for x in fc.modules()) if isinstance(fc, torch.nn.Module) else False


class _BackboneState(object):
    """What the backbone's backward needs from its forward."""
    __slots__ = ("unit0", "blocks", "final_shape", "dtype", "pool")


def _backbone_fwd(model, x, save, x_needs_grad, pool=True):
    """Stem -> layer1..4 -> avgpool on NHWC activations.  Returns (features [N,2048] fp32, state or None).
    pool=False returns the layer4 map itself as NCHW fp32 [N,2048,h,w] (the `include_top=False` + Flatten + Dense
    head of resnet.py:17-23 has no average pool)."""
    dtype = act_dtype()
    # num_batches_tracked += 1 for every BatchNorm in train mode: one multi-tensor launch instead of 53
    nbt = [m.num_batches_tracked for m in model.modules()
           if isinstance(m, torch.nn.BatchNorm2d) and m.training and m.track_running_stats
           and m.num_batches_tracked is not None and not _under(model.fc, m, model)]
    if nbt:
        torch._foreach_add_(nbt, 1)
    cp = _stem_cpad(model.conv1, dtype)
    if dtype == torch.bfloat16:
        prepack_weights(_backbone_convs(model), dtype)
    if x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[-1] == 4 and cp == 4:
        xh = x.contiguous()               # already the stem's layout (b200net.utils.GpuPreprocess): nothing to convert
        nchw_src = None
    else:
        c1 = model.conv1
        stage_pairs = (cp == 4 and not x_needs_grad and _STEM_PAIRS and x.dim() == 4 and x.shape[1] <= 4
                       and c1.stride in ((2, 2), 2) and c1.kernel_size[1] <= 8 and _STEM_FUSED_STAGE)
        nchw_src = x if stage_pairs else None
        xh = None if stage_pairs else ops.to_nhwc(x, dtype, cp)
    cur, u0 = _unit_forward(model.conv1, model.bn1, xh, True, None, save, stage=cp is not None and not x_needs_grad,
                            pool=True, nchw_src=nchw_src)
    blocks = []
    for layer in (model.layer1, model.layer2, model.layer3, model.layer4):
        for blk in layer:
            z1, u1 = _unit_forward(blk.conv1, blk.bn1, cur, True, None, save)
            z2, u2 = _unit_forward(blk.conv2, blk.bn2, z1, True, None, save)
            ud = None
            if blk.downsample is not None:
                res, ud = _unit_forward(blk.downsample[0], blk.downsample[1], cur, False, None, save)
            else:
                res = cur
            cur, u3 = _unit_forward(blk.conv3, blk.bn3, z2, True, res, save)
            blocks.append((u1, u2, u3, ud))
    feat = ops.avgpool_fwd(cur) if pool else ops.to_nchw(cur)
    if not save:
        return feat, None
    st = _BackboneState()
    st.pool = pool
    st.unit0, st.blocks, st.final_shape, st.dtype = u0, blocks, cur.shape, dtype
    return feat, st


_SIDE_STREAMS = {}
_WGRAD_OVERLAP = os.environ.get("B200_WGRAD_OVERLAP") is not None


def _side_stream(device):
    key = device.index if device.index is not None else torch.cuda.current_device()
    s = _SIDE_STREAMS.get(key)
    if s is None:
        s = torch.cuda.Stream(device=device)
        _SIDE_STREAMS[key] = s
    return s


def _backbone_bwd(st, dfeat, params, param_needs, x_needs_grad, sink, cut=None, cut_after=()):
    try:
        return _backbone_bwd_impl(st, dfeat, params, param_needs, x_needs_grad, sink, cut, cut_after)
    finally:
        if dfeat.is_cuda:
            ops.wgrad_defer(False)       # leave the deferred-reduction mode of this host thread, whatever happened


def _backbone_bwd_impl(st, dfeat, params, param_needs, x_needs_grad, sink, cut=None, cut_after=()):
    """Reverse walk: BN backward (mask from the saved byte mask) -> wgrad -> dgrad; the block-input gradient is the
    residual gradient with conv1's dgrad accumulated in the GEMM epilogue.  Returns (dx or None, [grad per param]).

    B200_WGRAD_OVERLAP=1 (OFF by default): weight gradients are off the critical path (nothing in the backward reads
    them), so they can run on a SIDE STREAM -- wgrad(u) starts as soon as dy_u exists and overlaps the dgrad ->
    BatchNorm-backward chain of the main stream; both streams are joined before the function returns (and before every
    graph-segment cut), and inside a CUDA-graph capture the fork / join become parallel branches of the graph.
    Measured at N=256: 13.87 ms with, 13.85 ms without.  Every kernel of this step already runs against the same
    shared limits (L2 slice throughput for the GEMM operands, HBM for the BatchNorm passes, one ~200 KB-smem CTA per
    SM for both convolution kernels), so running two of them at once only interleaves them.  Kept as an option."""
    dfeat = dfeat.contiguous().float()
    grads = {}  # id(param) -> grad
    capturing = torch.cuda.is_current_stream_capturing()
    overlap = _WGRAD_OVERLAP and dfeat.is_cuda and (sink is None or capturing)
    main = torch.cuda.current_stream() if overlap else None
    side = _side_stream(dfeat.device) if overlap else None
    keep = []        # tensors the side stream reads: kept alive until the streams are joined
    forked = [False]

    # Deferred split-K reductions: the 53 per-layer reduce launches of the weight gradients (0.41 ms of the step when
    # launched one by one) become one launch per graph segment (ops.wgrad_flush).  Each deferred layer keeps its own
    # slab buffer until then, and its gradient is handed on (to DDP's bucket copy, to .grad) only after the flush.  Not in
    # eager DDP mode, where a bucket should start its all-reduce as soon as its last gradient exists.
    defer = _WGRAD_DEFER and dfeat.is_cuda and not overlap and (sink is None or capturing)
    slab_bufs, pending = [], []
    if defer:
        ops.wgrad_defer(True)

    def flush():
        if defer and pending:
            ops.wgrad_flush()
            for p, g in pending:
                grads[id(p)] = sink(p, g) if sink is not None else g
            del pending[:]
            del slab_bufs[:]

    def join():
        flush()
        if overlap and forked[0]:
            main.wait_stream(side)
            forked[0] = False
            del keep[:]

    def emit(p, g, deferred=False):
        if deferred:
            pending.append((p, g))
        else:
            grads[id(p)] = sink(p, g) if sink is not None else g

    def unit_backward(u, dz, want_dres):
        if u.mean is None:
            raise RuntimeError("b200net: backward through BatchNorm in eval mode is not supported on this path")
        need_bn = u.bn.weight is not None and (u.bn.weight.requires_grad or u.bn.bias.requires_grad)
        if u.pool is not None and u.pool[0] == "fused":
            # dz is the gradient of the POOLED output: maxpool / ReLU / BatchNorm backward in one reduction + one dx pass
            dy, dgamma, dbeta = ops.maxpool_bn_bwd(dz, u.pool[1], u.pool[2], u.y, u.bn.weight, u.mean, u.invstd,
                                                   want_param_grads=need_bn)
            dres = None
        else:
            if u.pool is not None:
                dz = ops.maxpool_bwd(dz, u.pool[1], u.pool[2])
            dy, dres, dgamma, dbeta = ops.bn_backward(dz, u.z if u.relu else None, u.y, u.bn.weight, u.mean, u.invstd,
                                                      u.relu, want_dres=want_dres, want_param_grads=need_bn,
                                                      mask=u.mask, walk=1 if _L2_WALK else 0)
        if need_bn:
            if u.bn.bias.requires_grad:
                emit(u.bn.bias, dbeta)
            if u.bn.weight.requires_grad:
                emit(u.bn.weight, dgamma)
        K, R, S, stride, pad = u.geom
        if u.conv.weight.requires_grad:
            def wgrad():
                # layers whose gradient is post-processed by torch ops right away (the staged / channel-padded stem)
                # cannot wait for the flush
                plain = u.staged is None and u.x.shape[-1] == u.conv.in_channels
                if defer and not plain:
                    flush()
                    ops.wgrad_defer(False)
                dw = ops.conv_wgrad(u.x, dy, K, R, S, stride, pad, own_ws=slab_bufs if (defer and plain) else None)
                if defer and not plain:
                    ops.wgrad_defer(True)
                if defer and plain:
                    emit(u.conv.weight, dw, deferred=True)
                    return
                if u.staged is not None and u.staged[0] == "pairs":
                    # row-pair staged stem: dw is [K][64 = (r2, s, c)][R' = j][1]; filter row r = 2j + r2
                    _, r0, s0 = u.staged
                    dw = dw.reshape(K, 2, 8, 4, dw.shape[2]).permute(0, 3, 4, 1, 2).reshape(K, 4, 2 * dw.shape[2], 8)
                    dw = dw[:, :u.conv.in_channels, :r0, :s0].contiguous()
                elif u.staged is not None:                 # im2col-staged first layer: [K][R*S*cp (+pad)] -> OIHW
                    r0, s0, _, _, cp0, _ = u.staged
                    dw = dw.reshape(K, -1)[:, :r0 * s0 * cp0].reshape(K, r0, s0, cp0)
                    dw = dw[..., :u.conv.in_channels].permute(0, 3, 1, 2).contiguous()
                elif dw.shape[1] != u.conv.in_channels:    # channel-padded stem: drop the zero channel
                    dw = dw[:, :u.conv.in_channels].contiguous()
                emit(u.conv.weight, dw)
            if overlap:
                ev = torch.cuda.Event()
                ev.record(main)                            # dy is complete here
                side.wait_event(ev)
                keep.append(dy)
                forked[0] = True
                with torch.cuda.stream(side), ops.side_workspace():
                    wgrad()
            else:
                wgrad()
        return dy, dres

    def dgrad(u, dy, out=None, accumulate=False):
        K, R, S, stride, pad = u.geom
        cin = u.x.shape[-1]
        wp = packed_weight(u.conv.weight, dy.dtype, cin if cin != u.conv.in_channels else None)
        return ops.conv_dgrad(dy, wp, u.x.shape, K, R, S, stride, pad, out=out, accumulate=accumulate)

    dcur = ops.avgpool_bwd(dfeat, st.final_shape, st.dtype) if st.pool else ops.to_nhwc(dfeat, st.dtype)
    for bi, (u1, u2, u3, ud) in enumerate(reversed(st.blocks)):
        if cut is not None and bi in cut_after:
            join()
            cut()         # graph capture under DDP: a new segment starts here (see _SegmentedCapture)
        dy3, g = unit_backward(u3, dcur, want_dres=True)
        dz2 = dgrad(u3, dy3)
        dy2, _ = unit_backward(u2, dz2, want_dres=False)
        dz1 = dgrad(u2, dy2)
        dy1, _ = unit_backward(u1, dz1, want_dres=False)
        if ud is not None:
            # dense branch first (plain store), then the strided downsample branch accumulates: for a 1x1
            # stride-2 convolution only one of the four output-parity classes receives anything
            dyd, _ = unit_backward(ud, g, want_dres=False)
            dcur = dgrad(u1, dy1)
            dgrad(ud, dyd, out=dcur, accumulate=True)
        else:
            dcur = dgrad(u1, dy1, out=g, accumulate=True)
    u0 = st.unit0
    dy0, _ = unit_backward(u0, dcur, want_dres=False)
    dx = None
    if x_needs_grad:
        dx = ops.to_nchw(dgrad(u0, dy0))[:, :u0.conv.in_channels].contiguous()
    join()
    return dx, [grads.get(id(p)) if need else None for p, need in zip(params, param_needs)]


class _ResNetBackbone(torch.autograd.Function):
    """Eager path: every kernel is enqueued from Python on each call."""

    @staticmethod
    def forward(ctx, model, grad_on, x, *params):
        # forward() itself always runs with grad mode off and needs_input_grad ignores torch.no_grad(), so the
        # caller's grad mode comes in explicitly.  grad_on may carry the pool flag as (grad_on, pool).
        pool = True
        if isinstance(grad_on, tuple):
            grad_on, pool = grad_on
        nig = ctx.needs_input_grad[2:]
        save = grad_on and any(nig)
        feat, st = _backbone_fwd(model, x, save, grad_on and nig[0], pool)
        if save:
            ctx.state = st
            ctx.params = params
            ctx.param_needs = list(nig[1:])
            ctx.x_needs_grad = nig[0]
        return feat

    @staticmethod
    @once_differentiable
    def backward(ctx, dfeat):
        dx, grads = _backbone_bwd(ctx.state, dfeat, ctx.params, ctx.param_needs, ctx.x_needs_grad, _GRAD_SINK)
        ctx.state = None
        ctx.params = None
        return (None, None, dx) + tuple(grads)


# ------------------------------------------------------------------ CUDA-graph replay of the backbone
# The backbone is ~170 (forward) + ~400 (backward) kernel launches; enqueuing them from Python costs about as much
# CPU time as the GPU needs to run them, and at batch 1 far more.  After GRAPH_WARMUP eager calls with the same
# (model, input shape, mode, precision, set of trainable parameters) the forward and the backward are each captured
# into a CUDA graph once and replayed from then on: the reference's loop
# (pytorch_training_inference_on_image.ipynb:664-673, :855-861) is unchanged, only what happens inside
# model.forward / loss.backward differs.  B200_CUDA_GRAPHS=0 disables it.  Under DistributedDataParallel the copies
# into the gradient buckets are recorded in the backward graph and the all-reduces start right after the replay.
GRAPH_WARMUP = 3
GRAPH_CACHE = 6
_GRAPHS_ON = os.environ.get("B200_CUDA_GRAPHS", "1") != "0"
_GRAPHS = {}     # key -> _GraphRecord | "failed"
_GRAPH_SEEN = {}  # key -> number of eager calls so far
_CAPTURING = [False]


_GRAPH_CACHES = []   # other modules' graph caches (bert.py registers its own), cleared together with ours


def set_cuda_graphs(on):
    global _GRAPHS_ON
    _GRAPHS_ON = bool(on)
    if not on:
        _GRAPHS.clear()
        _GRAPH_SEEN.clear()
        for c in _GRAPH_CACHES:
            c.clear()


class _SegmentedCapture(object):
    """Captures one launch sequence into SEVERAL CUDA graphs (one private pool), split wherever the sequence calls
    cut().  Under DistributedDataParallel the backward is replayed segment by segment and the gradient buckets that
    a segment completed are all-reduced on the communication stream while the next segment runs."""

    def __init__(self, pool):
        self.pool, self.graphs, self.cur, self.marks = pool, [], None, []
        self.stream = torch.cuda.Stream()

    def __enter__(self):
        torch.cuda.synchronize()
        self.stream.wait_stream(torch.cuda.current_stream())
        self._ctx = torch.cuda.stream(self.stream)
        self._ctx.__enter__()
        self._begin()
        return self

    def _begin(self):
        self.cur = torch.cuda.CUDAGraph()
        self.cur.capture_begin(pool=self.pool)

    def cut(self, mark=None):
        self.cur.capture_end()
        self.graphs.append(self.cur)
        self.marks.append(mark() if mark is not None else None)
        self._begin()

    def __exit__(self, et, ev, tb):
        try:
            if self.cur is not None:
                self.cur.capture_end()
                if et is None:
                    self.graphs.append(self.cur)
                self.cur = None
        finally:
            self._ctx.__exit__(et, ev, tb)
            torch.cuda.current_stream().wait_stream(self.stream)
        return False


class _GraphRecord(object):
    __slots__ = ("fwd", "bwd", "x", "feat", "dfeat", "grads", "dx", "state", "ptrs", "ws", "uses", "ddp", "order",
                 "segments", "gen", "live")


class _ReplayToken(object):
    """Lives exactly as long as the autograd node of one replayed forward (it hangs off the node's ctx)."""
    __slots__ = ("__weakref__",)


def graph_record_busy(rec):
    """A graph record holds ONE set of saved activations.  While a forward replayed from it still waits for its
    backward (its autograd node is alive), a second forward must not replay the same record -- it would overwrite
    what the first backward reads (two micro-batches whose losses are summed, a siamese use, forward twice then
    backward).  Such calls take the eager node, which keeps its state per call."""
    live = getattr(rec, "live", None)
    return live is not None and live() is not None


def graph_replay_begin(rec, ctx):
    """Forward side of the guard: stamps the record's generation into ctx and registers the node as pending."""
    rec.gen = (getattr(rec, "gen", 0) or 0) + 1
    ctx.gen = rec.gen
    if rec.bwd is not None:
        ctx.token = _ReplayToken()
        rec.live = weakref.ref(ctx.token)


def graph_replay_check(rec, ctx):
    """Backward side: the saved state must still be the one this node's forward produced."""
    if getattr(ctx, "gen", None) != getattr(rec, "gen", None):
        raise RuntimeError("b200net: the CUDA-graph record of this forward was replayed again before its backward ran; "
                           "its saved activations are gone (set B200_CUDA_GRAPHS=0 for this usage)")
    rec.live = None
    ctx.token = None


_BN_MODULES = {}   # id(model) -> (structure signature, [BatchNorm2d modules])


def _bn_modules(model):
    """The model's BatchNorm2d modules, cached: walking model.modules() on every call cost ~0.5 ms of the ~1.5 ms of
    host time per training step that the GPU waits for when the loop reads loss.item() every step."""
    layers = tuple(getattr(model, n, None) for n in ("layer1", "layer2", "layer3", "layer4"))
    sig = (id(getattr(model, "bn1", None)), id(getattr(model, "fc", None))) + tuple(
        (id(l), len(l) if l is not None else 0) for l in layers)
    e = _BN_MODULES.get(id(model))
    if e is None or e[0] != sig:
        e = (sig, [m for m in model.modules() if isinstance(m, torch.nn.BatchNorm2d)])
        _BN_MODULES[id(model)] = e
    return e[1]


def _graph_key(model, x, params, needs_x, pool=True):
    bn_flags = tuple(m.training for m in _bn_modules(model))
    needs = tuple(p.requires_grad for p in params) if torch.is_grad_enabled() else ()
    # forward-only graphs (inference, frozen backbone) use weights packed OUTSIDE the graph, so they are only valid
    # for the weight versions they were captured with
    wver = -1 if (any(needs) or needs_x) else sum(p._version for p in params)
    return (id(model), tuple(x.shape), x.device.index, _PRECISION, bn_flags, needs, bool(needs_x), wver,
            id(grad_sink_owner()), bool(pool))


def _tensor_ptrs(model, params=None):
    """Addresses of everything a captured graph reads or writes in place: the backbone parameters and the BatchNorm
    running statistics (a re-allocated tensor invalidates the graph)."""
    if params is None:
        params = _backbone_params(model)
    sig = (_PARAM_EPOCH[0], _structure_sig(model))
    e = _STAT_CACHE.get(id(model))
    if e is None or e[0] != sig or e[2]() is not model:
        stats = []
        for m in _bn_modules(model):
            if m.running_mean is not None:
                stats += [m.running_mean, m.running_var]
        if len(_STAT_CACHE) > 64:
            _STAT_CACHE.clear()
        e = (sig, stats, weakref.ref(model))
        _STAT_CACHE[id(model)] = e
    return tuple([p.data_ptr() for p in params] + [t.data_ptr() for t in e[1]])


def _capture(model, x, params, key):
    needs = key[5]
    needs_x = key[6]
    save = any(needs) or needs_x
    rec = _GraphRecord()
    rec.x = x.detach().clone()
    rec.ptrs = _tensor_ptrs(model, params)
    rec.uses = 0
    rec.gen, rec.live = 0, None
    pool = torch.cuda.graph_pool_handle()
    _CAPTURE_PACKED.clear()
    # training graphs re-pack the weights on every replay (they change every step); forward-only graphs read the
    # packed copies the warm-up calls left in the eager cache (their validity is part of the key)
    _CAPTURING[0] = save
    ops.NO_WORKSPACE_GROWTH[0] = True   # the warm-up calls sized the workspace; growing it now would move it into
    try:                                # the graph's private pool
        torch.cuda.synchronize()
        rec.fwd = torch.cuda.CUDAGraph()
        with torch.no_grad():
            with torch.cuda.graph(rec.fwd, pool=pool):
                rec.feat, rec.state = _backbone_fwd(model, rec.x, save, needs_x, key[9])
        rec.bwd, rec.ddp, rec.order, rec.segments = None, None, None, None
        if save:
            rec.dfeat = torch.zeros_like(rec.feat)
            rec.ddp, rec.order = grad_sink_owner(), []
            if rec.ddp is None:
                rec.bwd = torch.cuda.CUDAGraph()
                with torch.no_grad():
                    with torch.cuda.graph(rec.bwd, pool=pool):
                        rec.dx, rec.grads = _backbone_bwd(rec.state, rec.dfeat, params, list(needs), needs_x, None)
            else:
                # DDP: one graph per stage group (layer4 | layer3 | the rest).  layer4 + layer3 hold 87 % of the
                # parameters and finish first, so their buckets travel while layer2 / layer1 / stem (most of the
                # backward's time at this resolution) are still running.
                pending = []
                sink = lambda p, g: rec.ddp._sink_capture(p, g, rec.order, pending)   # noqa: E731
                n4 = len(getattr(model, "layer4", ()))
                n3 = len(getattr(model, "layer3", ()))
                cuts = tuple(c for c in (n4, n4 + n3) if 0 < c < len(rec.state.blocks))

                def cut():
                    rec.ddp.flush_pending(pending)      # this segment's gradients -> bucket slices, one launch per 48
                    sc.cut(lambda: len(rec.order))

                with torch.no_grad():
                    with _SegmentedCapture(pool) as sc:
                        rec.dx, rec.grads = _backbone_bwd(rec.state, rec.dfeat, params, list(needs), needs_x, sink,
                                                          cut=cut, cut_after=cuts)
                        rec.ddp.flush_pending(pending)
                bounds = [0] + list(sc.marks) + [len(rec.order)]
                rec.segments = [(g, rec.order[bounds[i]:bounds[i + 1]]) for i, g in enumerate(sc.graphs)]
                rec.bwd = sc.graphs[0]
        rec.ws = (dict(ops._WS), [e[3] for e in _PACKED.values()])   # raw pointers captured: keep the memory alive
    finally:
        _CAPTURING[0] = False
        ops.NO_WORKSPACE_GROWTH[0] = False
        _CAPTURE_PACKED.clear()
    return rec


class _GraphedBackbone(torch.autograd.Function):
    @staticmethod
    def forward(ctx, rec, x, *params):
        rec.x.copy_(x)
        rec.fwd.replay()
        graph_replay_begin(rec, ctx)
        ctx.rec = rec
        ctx.saved_params = params
        return rec.feat.clone()

    @staticmethod
    @once_differentiable
    def backward(ctx, dfeat):
        rec = ctx.rec
        if rec.bwd is None:
            raise RuntimeError("b200net: this backbone graph was captured without a backward pass")
        graph_replay_check(rec, ctx)
        rec.dfeat.copy_(dfeat)
        if rec.ddp is not None:
            for g, ready in rec.segments:
                g.replay()
                rec.ddp._replayed(ready)      # completed buckets start their all-reduce while the next segment runs
            return (None, rec.dx.clone() if rec.dx is not None else None) + tuple(rec.grads)
        rec.bwd.replay()
        # The gradients live in graph-owned memory that the next replay overwrites.  With the usual
        # zero_grad() -> backward() -> step() cycle they are consumed before that, so they are handed out as they
        # are; a parameter that still holds a .grad (gradient accumulation) gets a private copy instead, because
        # autograd would otherwise add the buffer to itself.
        params = ctx.saved_params
        out = [None if g is None else (g if p.grad is None else g.clone()) for p, g in zip(params, rec.grads)]
        return (None, rec.dx.clone() if rec.dx is not None else None) + tuple(out)


def resnet_backbone(model, x, pool=True):
    """[N,3,H,W] fp32 NCHW -> [N,2048] fp32 features (everything up to and including avgpool + flatten); with
    pool=False the layer4 map [N,2048,h,w] fp32 instead."""
    params = _backbone_params(model)
    owner = grad_sink_owner()
    if not (_GRAPHS_ON and x.is_cuda and not torch.cuda.is_current_stream_capturing()
            and (_GRAD_SINK is None or (owner is not None and owner.graph_ok()))):
        return _ResNetBackbone.apply(model, (torch.is_grad_enabled(), pool), x, *params)
    needs_x = x.requires_grad and torch.is_grad_enabled()
    key = _graph_key(model, x, params, needs_x, pool)
    rec = _GRAPHS.get(key)
    if rec is not None and rec != "failed":
        if rec.ptrs == _tensor_ptrs(model, params):
            if graph_record_busy(rec):     # an earlier replayed forward still waits for its backward
                return _ResNetBackbone.apply(model, (torch.is_grad_enabled(), pool), x, *params)
            rec.uses += 1
            return _GraphedBackbone.apply(rec, x, *params)
        del _GRAPHS[key]          # parameters were re-allocated (model.to(), new tensors): recapture
        _GRAPH_SEEN[key] = 0
        rec = None
    if rec is None:
        seen = _GRAPH_SEEN.get(key, 0)
        if seen >= GRAPH_WARMUP:
            try:
                if len(_GRAPHS) >= GRAPH_CACHE:      # bounded cache: drop the least used graph
                    victim = min((k for k in _GRAPHS if _GRAPHS[k] != "failed"), key=lambda k: _GRAPHS[k].uses,
                                 default=None)
                    if victim is not None:
                        del _GRAPHS[victim]
                rec = _capture(model, x, params, key)
                _GRAPHS[key] = rec
                rec.uses = 1
                return _GraphedBackbone.apply(rec, x, *params)
            except Exception as e:                    # capture is an optimisation: never fatal
                import warnings
                warnings.warn("b200net: CUDA-graph capture of the backbone failed (%s); staying on the eager path" % e)
                _GRAPHS[key] = "failed"
        else:
            if len(_GRAPH_SEEN) > 256:               # variable-size inference: do not grow without bound
                _GRAPH_SEEN.clear()
            _GRAPH_SEEN[key] = seen + 1
    return _ResNetBackbone.apply(model, (torch.is_grad_enabled(), pool), x, *params)


# ------------------------------------------------------------------ classifier head (fp32, [rows][features])
def _c2d(t):
    if not t.is_cuda:
        raise RuntimeError("b200net kernels run on CUDA tensors only (no CPU fallback)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class _Linear(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        x, w = _c2d(x), _c2d(w)
        ctx.save_for_backward(x, w)
        ctx.has_bias = b is not None
        return ops.linear_fwd(x, w, _c2d(b) if b is not None else None)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = _c2d(dy)
        dx = ops.linear_bwd_input(dy, w) if ctx.needs_input_grad[0] else None
        dw = ops.linear_bwd_weight(dy, x) if ctx.needs_input_grad[1] else None
        db = ops.colsum(dy) if (ctx.has_bias and ctx.needs_input_grad[2]) else None
        return dx, dw, db


def linear(x, w, b=None):
    lead = x.shape[:-1]
    y = _Linear.apply(x.reshape(-1, x.shape[-1]), w, b)
    return y.reshape(*lead, w.shape[0])


class _ReLU(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = ops.relu_fwd(_c2d(x))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return ops.relu_bwd(_c2d(dy), y)


def relu(x):
    return _ReLU.apply(x)


class _Tanh(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = ops.tanh_fwd(_c2d(x))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return ops.tanh_bwd(_c2d(dy), y)


def tanh(x):
    """BertPooler activation (fp32)"""
    return _Tanh.apply(x)


_DROPOUT_CALLS = [0]


class _Dropout(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p):
        x = _c2d(x)
        seed = torch.cuda.initial_seed() & 0xFFFFFFFFFFFFFFFF
        offset = _DROPOUT_CALLS[0]
        _DROPOUT_CALLS[0] += (x.numel() + 3) // 4
        y, mask = ops.dropout_fwd(x, p, seed, offset)
        ctx.save_for_backward(mask)
        ctx.p = p
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        return ops.dropout_bwd(_c2d(dy), mask, ctx.p), None


def dropout(x, p, training):
    if not training or p == 0.0:
        return x
    return _Dropout.apply(x, p)


class _LogSoftmax(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = ops.log_softmax_fwd(_c2d(x))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return ops.log_softmax_bwd(_c2d(dy), y)


def log_softmax(x, dim=-1):
    if x.dim() != 2 or dim not in (1, -1):
        raise NotImplementedError("b200net log_softmax: 2-D input, dim=1")
    return _LogSoftmax.apply(x)


class _NLLLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logp, labels):
        logp = _c2d(logp)
        labels = labels.contiguous()
        if labels.dtype != torch.int64:
            labels = labels.long()
        ctx.save_for_backward(labels)
        ctx.shape = logp.shape
        return ops.nll_loss_fwd(logp, labels)

    @staticmethod
    @once_differentiable
    def backward(ctx, dloss):
        (labels,) = ctx.saved_tensors
        return ops.nll_loss_bwd(_c2d(dloss).reshape(1), labels, ctx.shape[0], ctx.shape[1]), None


def nll_loss(logp, labels):
    return _NLLLoss.apply(logp, labels)


# ------------------------------------------------------------------ stand-alone NCHW layers (module-level use)
class _Conv2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, geom):
        K, R, S, stride, pad = geom
        dtype = act_dtype()
        xh = ops.to_nhwc(x, dtype)
        y = ops.conv_fprop(xh, packed_weight(w, dtype), K, R, S, stride, pad)
        ctx.save_for_backward(xh, w)
        ctx.geom = geom
        return ops.to_nchw(y)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        xh, w = ctx.saved_tensors
        K, R, S, stride, pad = ctx.geom
        dyh = ops.to_nhwc(dy, xh.dtype)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dx = ops.to_nchw(ops.conv_dgrad(dyh, packed_weight(w, xh.dtype), xh.shape, K, R, S, stride, pad))
        if ctx.needs_input_grad[1]:
            dw = ops.conv_wgrad(xh, dyh, K, R, S, stride, pad)
        return dx, dw, None


def conv2d_module(conv, x):
    return _Conv2d.apply(x, conv.weight, _conv_geom(conv))


class _BatchNorm2d(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, bn):
        xh = ops.to_nhwc(x, act_dtype())
        scale, shift, mean, invstd = _bn_coeffs(bn, xh)
        z = ops.bn_apply(xh, scale, shift, False)
        ctx.bn = bn
        ctx.stats = (mean, invstd)
        ctx.save_for_backward(xh)
        return ops.to_nchw(z)

    @staticmethod
    @once_differentiable
    def backward(ctx, dz):
        (xh,) = ctx.saved_tensors
        mean, invstd = ctx.stats
        if mean is None:
            raise RuntimeError("b200net: backward through BatchNorm in eval mode is not supported")
        dzh = ops.to_nhwc(dz, xh.dtype)
        dx, _, dgamma, dbeta = ops.bn_backward(dzh, None, xh, ctx.bn.weight, mean, invstd, False)
        return ops.to_nchw(dx), dgamma, dbeta, None


def batch_norm_module(bn, x):
    return _BatchNorm2d.apply(x, bn.weight, bn.bias, bn)


class _MaxPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xh = ops.to_nhwc(x, act_dtype())
        y, idx = ops.maxpool_fwd(xh, True)
        ctx.save_for_backward(idx)
        ctx.shape = xh.shape
        ctx.dtype = xh.dtype
        return ops.to_nchw(y)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        return ops.to_nchw(ops.maxpool_bwd(ops.to_nhwc(dy, ctx.dtype), idx, ctx.shape))


def maxpool3x3s2(x):
    return _MaxPool.apply(x)


class _AvgPool(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        xh = ops.to_nhwc(x, act_dtype())
        ctx.shape = xh.shape
        ctx.dtype = xh.dtype
        return ops.avgpool_fwd(xh).reshape(x.shape[0], x.shape[1], 1, 1)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        N, H, W, C = ctx.shape
        return ops.to_nchw(ops.avgpool_bwd(dy.reshape(N, C).contiguous().float(), ctx.shape, ctx.dtype))


def global_avgpool(x):
    return _AvgPool.apply(x)

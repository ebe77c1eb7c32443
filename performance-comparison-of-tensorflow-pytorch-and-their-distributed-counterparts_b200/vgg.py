"""VGG16 with torchvision's object protocol -- the second image model of the reference
(SURVEY.md 8(f) rank 1):

    model = models.vgg16(pretrained=True)                                   another_neural_net.py:244
    for param in model.parameters(): param.requires_grad = False            :247-248
    model.classifier[6] = nn.Sequential(nn.Linear(4096, 256), nn.ReLU(), nn.Dropout(0.4),
                                        nn.Linear(256, 10), nn.LogSoftmax(dim=1))          :250-255
    (same cells: pytorch_training_inference_on_image.ipynb:1968-2041, Standalone_Inference_Imagenette_trial.ipynb:355-374)

Module tree, repr and state_dict keys are torchvision's (`features.0.weight` ... `classifier.6.bias`).  `features` +
`avgpool` + flatten run as ONE autograd node of hand-scheduled launches: every conv3x3+bias+ReLU is the tcgen05
implicit-GEMM kernel with the bias and the ReLU in its epilogue (the 3-channel first layer is staged once as a
[pixels][64] matrix, like the ResNet stem), MaxPool2d(2,2) and AdaptiveAvgPool2d(7) are the kernels of csrc/vgg.cu.
The two 4096-wide classifier Linears run on the same GEMM kernel (bias + ReLU in the epilogue, Philox dropout);
whatever sits at classifier[6] (the reference's small head) runs on the fp32 head kernels.
"""
import torch
import torch.nn as _nn
from torch.autograd.function import once_differentiable

from . import functional as F
from . import nn
from . import ops

_CALLS = [0]
_M64 = 0xFFFFFFFFFFFFFFFF


def _site(call, site):
    return ((call * 64 + site) << 34) & _M64


def _feature_plan(features):
    """[(conv, relu_follows) | 'pool'] from the nn.Sequential; anything else is outside what this path implements."""
    mods = list(features)
    plan, i = [], 0
    while i < len(mods):
        m = mods[i]
        if isinstance(m, _nn.Conv2d):
            if (m.kernel_size, m.stride, m.padding, m.dilation, m.groups) != ((3, 3), (1, 1), (1, 1), (1, 1), 1):
                raise NotImplementedError("b200net VGG: 3x3 / stride 1 / padding 1 convolutions only, got %r" % (m,))
            relu = i + 1 < len(mods) and isinstance(mods[i + 1], _nn.ReLU)
            plan.append((m, relu))
            i += 2 if relu else 1
        elif isinstance(m, _nn.MaxPool2d):
            k = m.kernel_size if isinstance(m.kernel_size, int) else m.kernel_size[0]
            s = m.stride if isinstance(m.stride, int) else m.stride[0]
            p = m.padding if isinstance(m.padding, int) else m.padding[0]
            if (k, s, p) != (2, 2, 0) or m.ceil_mode:
                raise NotImplementedError("b200net VGG: MaxPool2d(kernel_size=2, stride=2) only, got %r" % (m,))
            plan.append("pool")
            i += 1
        else:
            raise NotImplementedError("b200net VGG: unsupported module in features: %r" % (m,))
    return plan


def _feature_params(features):
    ps = []
    for m in features:
        if isinstance(m, _nn.Conv2d):
            ps.append(m.weight)
            if m.bias is not None:
                ps.append(m.bias)
    return ps


class _FeatState(object):
    pass


def _bias_of(conv, device):
    if conv.bias is not None:
        return conv.bias.detach()
    return torch.zeros(conv.out_channels, dtype=torch.float32, device=device)


def _features_fwd(model, x, save, out_hw):
    dt = F.act_dtype()
    plan = _feature_plan(model.features)
    first = plan[0][0] if plan and plan[0] != "pool" else None
    stage = dt == torch.bfloat16 and first is not None and first.in_channels % 64 != 0
    cur = ops.to_nhwc(x, dt, cpad=4 if (stage and first.in_channels == 3) else None)
    st = _FeatState()
    st.ops = []
    for item in plan:
        if item == "pool":
            shape = cur.shape
            cur, idx = ops.maxpool2_fwd(cur, want_index=save)
            if save:
                st.ops.append(("pool", idx, shape))
            continue
        conv, relu = item
        K = conv.out_channels
        bias = _bias_of(conv, cur.device)
        ones = ops._ones(K, cur.device)
        staged = None
        xin = cur
        if conv is first and stage:
            # 3-channel input: stage once as [pixels][64] (27 real columns, channel-padded taps) and run as a 1x1 problem
            cp = cur.shape[-1]
            ld = (9 * cp + 63) // 64 * 64
            wp = F.packed_weight(conv.weight, dt, cp, ld)
            staged = (cp, cur.shape)
            xin = ops.im2col(cur, 3, 3, 1, 1, ld)
            geom = (K, 1, 1, 1, 0)
        else:
            wp = F.packed_weight(conv.weight, dt)
            geom = (K, 3, 3, 1, 1)
        z = ops.conv_fprop_affine(xin, wp, *geom, ones, bias, relu) if dt == torch.bfloat16 else None
        if z is None:
            y = ops.conv_fprop(xin, wp, *geom)
            z = ops.bn_apply(y, ones, bias, relu, out=y)       # y = act(conv + bias)
        if save:
            st.ops.append(("conv", conv, relu, xin, z, geom, staged))
        cur = z
    st.pre_pool_shape = cur.shape
    if tuple(cur.shape[1:3]) != tuple(out_hw):
        cur = ops.adaptive_avgpool_fwd(cur, out_hw[0], out_hw[1])
    st.dtype = dt
    flat = ops.to_nchw(cur).reshape(cur.shape[0], -1)           # torchvision flattens NCHW: c*49 + h*7 + w
    st.pooled_shape = cur.shape
    return flat, st


def _features_bwd(st, dflat, emit, x_needs_grad):
    N, OH, OW, C = st.pooled_shape
    d = ops.to_nhwc(dflat.contiguous().float().view(N, C, OH, OW), st.dtype)
    if tuple(st.pre_pool_shape) != tuple(st.pooled_shape):
        d = ops.adaptive_avgpool_bwd(d, st.pre_pool_shape)
    n_ops = len(st.ops)
    for oi in range(n_ops - 1, -1, -1):
        op = st.ops[oi]
        st.ops[oi] = None
        if op[0] == "pool":
            d = ops.maxpool2_bwd(d, op[1], op[2])
            continue
        _, conv, relu, xin, z, geom, staged = op
        K, R, S, stride, pad = geom
        dz = ops.relu_bwd(d, z) if relu else d
        if conv.bias is not None and conv.bias.requires_grad:
            emit(conv.bias, ops.colsum_any(dz.reshape(-1, K)))
        if conv.weight.requires_grad:
            dw = ops.conv_wgrad(xin, dz, K, R, S, stride, pad)
            if staged is not None:
                cp, _ = staged
                dw = dw.reshape(K, -1)[:, :9 * cp].reshape(K, 3, 3, cp)[..., :conv.in_channels]
                dw = dw.permute(0, 3, 1, 2).contiguous()
            emit(conv.weight, dw)
        last = oi == 0
        if last and not x_needs_grad:
            return None
        if staged is not None:
            raise NotImplementedError("b200net VGG: gradient w.r.t. the input image is not implemented")
        wp = F.packed_weight(conv.weight, dz.dtype)
        d = ops.conv_dgrad(dz, wp, xin.shape, K, R, S, stride, pad)
    return ops.to_nchw(d)


class _VGGFeatures(torch.autograd.Function):
    """features + avgpool + flatten as one autograd node: [N,3,H,W] fp32 -> [N, 512*7*7] fp32."""

    @staticmethod
    def forward(ctx, model, grad_on, x, *params):
        nig = ctx.needs_input_grad[2:]
        save = grad_on and any(nig)
        osz = model.avgpool.output_size
        out_hw = (osz, osz) if isinstance(osz, int) else tuple(osz)
        flat, st = _features_fwd(model, x, save, out_hw)
        if save:
            ctx.state, ctx.params, ctx.x_needs_grad = st, params, nig[0]
        return flat

    @staticmethod
    @once_differentiable
    def backward(ctx, dflat):
        grads = {}
        sink = F._GRAD_SINK

        def emit(p, g):
            grads[id(p)] = sink(p, g) if sink is not None else g

        dx = _features_bwd(ctx.state, dflat, emit, ctx.x_needs_grad)
        out = tuple(grads.get(id(p)) for p in ctx.params)
        ctx.state = None
        ctx.params = None
        return (None, None, dx) + out


# ------------------------------------------------------------------ classifier: Linear (+ReLU) (+Dropout) on the GEMM kernel
_WCAST = {}


def _cast(src, dtype):
    out = torch.empty(src.shape, dtype=dtype, device=src.device)
    ops.check(ops.lib.b200_cast(src.data_ptr(), out.data_ptr(), src.numel(), ops._dt(src), ops._DT[dtype],
                                ops._stream()), "cast")
    return out


def _weight_operand(w, dtype):
    """bf16 copy of a Linear weight, re-cast when the parameter's storage or version counter changes."""
    if dtype == torch.float32:
        return w.detach()
    e = _WCAST.get(id(w))
    if e is not None and e[0] == (w.data_ptr(), w._version):
        return e[1]
    out = _cast(w.detach(), dtype)
    _WCAST[id(w)] = ((w.data_ptr(), w._version), out)
    return out


def _to_act(x, dtype):
    return x if x.dtype == dtype else _cast(x, dtype)


class _DenseReLUDrop(torch.autograd.Function):
    """y = dropout(relu(x @ w^T + b)) for the 25088->4096 and 4096->4096 layers: tcgen05 GEMM with bias+ReLU in the
    epilogue, Philox dropout regenerated in the backward.  Input / output are fp32 [N][features] at the boundary."""

    @staticmethod
    def forward(ctx, x, w, b, relu, p, seed, offset):
        dt = F.act_dtype()
        xa = _to_act(x.contiguous(), dt)
        bias = b.detach() if b is not None else torch.zeros(w.shape[0], dtype=torch.float32, device=w.device)
        z = ops.dense_fwd(xa, _weight_operand(w, dt), bias, relu=relu)
        y = ops.dropout_act(z, p, seed, offset) if p > 0.0 else z
        ctx.save_for_backward(xa, w, z)
        ctx.cfg = (relu, p, seed, offset, b is not None)
        return _to_act(y, torch.float32)

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        xa, w, z = ctx.saved_tensors
        relu, p, seed, offset, has_bias = ctx.cfg
        d = _to_act(dy.contiguous().float(), xa.dtype)
        if d.data_ptr() == dy.data_ptr():
            d = d.clone()
        if p > 0.0:
            d = ops.dropout_act(d, p, seed, offset, out=d)
        if relu:
            d = ops.relu_bwd(d, z)
        dx = ops.dense_bwd_input(d, _weight_operand(w, xa.dtype), w.shape[1]) if ctx.needs_input_grad[0] else None
        dw = ops.dense_bwd_weight(d, xa) if ctx.needs_input_grad[1] else None
        db = ops.colsum_any(d) if (has_bias and ctx.needs_input_grad[2]) else None
        return (_to_act(dx, torch.float32) if dx is not None else None), dw, db, None, None, None, None


def _run_classifier(classifier, x, training):
    """Runs model.classifier: Linear layers with >= 1024 inputs (the two 4096-wide ones) take the tensor-core path
    together with the ReLU / Dropout that follow them; the rest (the reference's classifier[6] head) runs on the fp32
    head kernels exactly like ResNet's `fc`."""
    mods = list(classifier) if isinstance(classifier, _nn.Sequential) else [classifier]
    seed = torch.cuda.initial_seed() & _M64
    call = _CALLS[0]
    _CALLS[0] += 1
    i, site = 0, 0
    while i < len(mods):
        m = mods[i]
        big = (isinstance(m, _nn.Linear) and m.in_features >= 1024 and m.in_features % 64 == 0
               and m.out_features % 64 == 0 and F.get_precision() == "bf16")
        if big:
            relu = i + 1 < len(mods) and isinstance(mods[i + 1], _nn.ReLU)
            j = i + (2 if relu else 1)
            p = 0.0
            if j < len(mods) and isinstance(mods[j], _nn.Dropout):
                p = float(mods[j].p) if training else 0.0
                j += 1
            x = _DenseReLUDrop.apply(x, m.weight, m.bias, relu, p, seed, _site(call, site))
            site += 1
            i = j
        else:
            x = nn.run_head(m, x)
            i += 1
    return x


class VGG(_nn.Module):
    """torchvision.models.VGG: same constructor, attribute names, repr and state_dict."""

    def __init__(self, features, num_classes=1000, init_weights=True, dropout=0.5):
        super().__init__()
        self.features = features
        self.avgpool = nn.AdaptiveAvgPool2d((7, 7))
        self.classifier = nn.Sequential(
            nn.Linear(512 * 7 * 7, 4096), nn.ReLU(True), nn.Dropout(p=dropout),
            nn.Linear(4096, 4096), nn.ReLU(True), nn.Dropout(p=dropout),
            nn.Linear(4096, num_classes))
        if init_weights:
            for m in self.modules():
                if isinstance(m, _nn.Conv2d):
                    _nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                    if m.bias is not None:
                        _nn.init.constant_(m.bias, 0)
                elif isinstance(m, _nn.Linear):
                    _nn.init.normal_(m.weight, 0, 0.01)
                    _nn.init.constant_(m.bias, 0)

    def forward(self, x):
        if not x.is_cuda:
            raise RuntimeError("b200net.models.VGG runs on CUDA tensors only (no CPU fallback); "
                               "call model.to('cuda') and inputs.to('cuda') as the reference loop does")
        params = _feature_params(self.features)
        flat = _VGGFeatures.apply(self, torch.is_grad_enabled(), x, *params)
        return _run_classifier(self.classifier, flat, self.training)


_CFG_D = [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"]


def _make_layers(cfg):
    layers, c = [], 3
    for v in cfg:
        if v == "M":
            layers.append(_nn.MaxPool2d(kernel_size=2, stride=2))
        else:
            layers += [_nn.Conv2d(c, v, kernel_size=3, padding=1), _nn.ReLU(inplace=True)]
            c = v
    return _nn.Sequential(*layers)


def vgg16(pretrained=False, progress=True, weights=None, **kwargs):
    """Same call as `models.vgg16(pretrained=True)` (another_neural_net.py:244).  No network on the build / GPU boxes:
    `pretrained=True` raises unless torchvision finds the checkpoint in its local cache; load one with
    `model.load_state_dict(torch.load(path))` (keys are torchvision's)."""
    model = VGG(_make_layers(_CFG_D), **kwargs)
    if pretrained or weights is not None:
        import torchvision
        ref = torchvision.models.vgg16(weights=weights if weights is not None else "IMAGENET1K_V1")
        model.load_state_dict(ref.state_dict())
    return model

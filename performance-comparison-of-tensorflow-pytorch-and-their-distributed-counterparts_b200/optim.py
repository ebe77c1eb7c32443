"""torch.optim-compatible optimizers whose `step()` is one multi-tensor kernel launch per <=48 tensors.

Same constructor signatures and `param_groups` protocol as torch.optim.SGD / Adam / AdamW, so the reference's
statements -- `optim.Adam(model.fc.parameters(), lr=0.003)` (pytorch_training_inference_on_image.ipynb:445),
Keras `SGD(lr=0.001)` (resnet.py:24), `AdamW(model.parameters(), lr=2e-5, eps=1e-8)` +
`get_linear_schedule_with_warmup` (pytorch_on_language_distr.py:167-183) -- work unchanged, including LR schedulers
that rewrite `param_groups[i]['lr']`.
"""
import torch
from torch.optim.optimizer import Optimizer

from . import _C
from ._C import lib, check, ptr_array, i64_array


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _bump(params):
    inc = torch.autograd.graph.increment_version
    for p in params:
        inc(p)


def _collect(group):
    ps, gs = [], []
    for p in group["params"]:
        if p.grad is None:
            continue
        if not p.is_cuda:
            raise RuntimeError("b200net.optim: parameters must live on a CUDA device (no CPU fallback)")
        if p.dtype != torch.float32 or not p.is_contiguous():
            raise RuntimeError("b200net.optim: parameters must be contiguous float32")
        g = p.grad
        if g.is_sparse:
            raise RuntimeError("b200net.optim: sparse gradients are not supported")
        if g.dtype != torch.float32 or not g.is_contiguous():
            g = g.float().contiguous()
        ps.append(p)
        gs.append(g)
    return ps, gs


class _FastZeroGrad(object):
    """`optimizer.zero_grad()` (set_to_none, the default) as a plain loop: torch's version spends ~0.17 ms per call on
    profiler ranges and per-tensor bookkeeping for ResNet-50's 161 tensors, host time that sits between one step's
    `loss.item()` and the next forward's first kernel."""

    def zero_grad(self, set_to_none=True):
        if not set_to_none:
            return super().zero_grad(set_to_none=False)
        for group in self.param_groups:
            for p in group["params"]:
                p.grad = None


class SGD(_FastZeroGrad, Optimizer):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(params, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                      nesterov=nesterov))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            ps, gs = _collect(group)
            if not ps:
                continue
            mom = group["momentum"]
            if mom != 0:
                # torch semantics: the momentum buffer starts as a copy of the first gradient, per parameter
                fresh = [(p, g) for p, g in zip(ps, gs) if "momentum_buffer" not in self.state[p]]
                warm = [(p, g) for p, g in zip(ps, gs) if "momentum_buffer" in self.state[p]]
                for p, _ in fresh:
                    self.state[p]["momentum_buffer"] = torch.empty_like(p)
                parts = [(fresh, 1), (warm, 0)]
            else:
                parts = [(list(zip(ps, gs)), 0)]
            for sel, first in parts:
                if not sel:
                    continue
                bufs = ptr_array([self.state[p]["momentum_buffer"].data_ptr() for p, _ in sel]) if mom != 0 else None
                check(lib.b200_multi_tensor_sgd(
                    len(sel), ptr_array([p.data_ptr() for p, _ in sel]), ptr_array([g.data_ptr() for _, g in sel]),
                    bufs, i64_array([p.numel() for p, _ in sel]), float(group["lr"]), float(mom),
                    float(group["dampening"]), float(group["weight_decay"]), 1 if group["nesterov"] else 0,
                    first, 1.0, _stream()), "multi_tensor_sgd")
            _bump(ps)
        return loss


class Adam(_FastZeroGrad, Optimizer):
    _decoupled = False
    # transformers.AdamW divides by (sqrt(v) + eps) with the step size lr*sqrt(bc2)/bc1; torch divides by
    # (sqrt(v)/sqrt(bc2) + eps) with lr/bc1.  The two coincide when torch's rule is given eps/sqrt(bc2) at that step,
    # which is how bert.AdamW selects the published transformers rule on the same kernel.
    _eps_before_bias_correction = False

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, amsgrad=False):
        if amsgrad:
            raise NotImplementedError("b200net.optim: amsgrad is not implemented")
        if lr < 0.0:
            raise ValueError("Invalid learning rate: {}".format(lr))
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            ps, gs = _collect(group)
            if not ps:
                continue
            by_step = {}
            for p, g in zip(ps, gs):
                st = self.state[p]
                if not st:
                    st["step"] = 0
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
                st["step"] += 1
                by_step.setdefault(int(st["step"]), []).append((p, g, st))
            b1, b2 = group["betas"]
            for step, sel in by_step.items():
                eps = float(group["eps"])
                if self._eps_before_bias_correction:
                    eps = eps / (1.0 - float(b2) ** step) ** 0.5
                check(lib.b200_multi_tensor_adam(
                    len(sel), ptr_array([p.data_ptr() for p, _, _ in sel]), ptr_array([g.data_ptr() for _, g, _ in sel]),
                    ptr_array([s["exp_avg"].data_ptr() for _, _, s in sel]),
                    ptr_array([s["exp_avg_sq"].data_ptr() for _, _, s in sel]),
                    i64_array([p.numel() for p, _, _ in sel]), float(group["lr"]), float(b1), float(b2),
                    eps, float(group["weight_decay"]), 1 if self._decoupled else 0, step, 1.0,
                    _stream()), "multi_tensor_adam")
            _bump(ps)
        return loss


class AdamW(Adam):
    _decoupled = True

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2, amsgrad=False):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad)

"""Warm CUDA-graph-replay timing of the ResNet-50 bandwidth kernels at the headline shapes (N=256, 160x160 input,
bf16): BatchNorm statistics / apply / backward for each distinct activation shape of the network, max-pool, optimizer.
usage: python tools/resnet_kernel_bench.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from b200net import ops  # noqa: E402

dev = torch.device("cuda:0")
dt = torch.bfloat16
N = 256


def timeit(name, fn, nbytes, iters=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        for _ in range(iters):
            fn()
    graph.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    graph.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    print("%-44s %8.1f us  %6.2f TB/s" % (name, us, nbytes / us / 1e6), flush=True)
    return us


total = 0.0
# (H=W, C, count of BN layers with this output shape) for a 160x160 input
shapes = [(80, 64, 1), (40, 64, 6), (40, 256, 4), (20, 128, 8), (20, 512, 5), (10, 256, 12), (10, 1024, 7),
          (5, 512, 6), (5, 2048, 4), (40, 128, 1), (20, 256, 1), (10, 512, 1)]
for hw, C, cnt in shapes:
    x = torch.randn(N, hw, hw, C, device=dev).to(dt)
    res = torch.randn(N, hw, hw, C, device=dev).to(dt)
    dz = torch.randn(N, hw, hw, C, device=dev).to(dt)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    el = x.numel() * 2
    mean, invstd, scale, shift = ops.bn_train_stats(x, gamma, beta, rm, rv, 0.1, 1e-5)
    y, mask = ops.bn_apply(x, scale, shift, True, want_mask=True)
    tag = "[%d,%d,%d,%d] x%d" % (N, hw, hw, C, cnt)
    a = timeit("bn_train_stats " + tag, lambda: ops.bn_train_stats(x, gamma, beta, rm, rv, 0.1, 1e-5), el)
    b = timeit("bn_apply+relu+mask " + tag, lambda: ops.bn_apply(x, scale, shift, True, want_mask=True), 2 * el)
    c = timeit("bn_apply+res+relu " + tag, lambda: ops.bn_apply(x, scale, shift, True, residual=res, want_mask=True), 3 * el)
    d = timeit("bn_backward (mask) " + tag, lambda: ops.bn_backward(dz, None, x, gamma, mean, invstd, True, mask=mask), 3 * el)
    e = timeit("bn_backward +dres " + tag, lambda: ops.bn_backward(dz, None, x, gamma, mean, invstd, True, want_dres=True, mask=mask), 4 * el)
    total += cnt * (b + d)
print("sum over the network of (apply + backward) ~ %.2f ms" % (total / 1e3))

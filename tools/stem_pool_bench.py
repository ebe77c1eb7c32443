"""Times the stem's BatchNorm -> ReLU -> MaxPool tail at the benchmark shape (N=256, 80x80x64, bf16): the separate
kernels (bn_apply + maxpool_fwd; maxpool_bwd + bn_backward) against the one-pass kernels of csrc/stem_pool.cu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from b200net import ops

dev = torch.device("cuda:0")
N, H, W, C = int(os.environ.get("N", "256")), 80, 80, 64
y = torch.randn(N, H, W, C, device=dev).to(torch.bfloat16)
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, invstd, scale, shift = ops.bn_train_stats(y, gamma, beta, None, None, 0.0, 1e-5)
z, mask = ops.bn_apply(y, scale, shift, True, None, want_mask=True)
p, idx = ops.maxpool_fwd(z, True)
dpool = torch.randn_like(p)
dz = ops.maxpool_bwd(dpool, idx, z.shape)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(name, fn, bytes_):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    t = sorted(ts)[len(ts) // 2]
    print("%-44s %8.1f us  %6.0f GB/s (algorithmic %.0f MB)" % (name, t, bytes_ / t / 1e3, bytes_ / 1e6))


e, pe = y.numel(), p.numel()
timeit("bn_apply (+mask)", lambda: ops.bn_apply(y, scale, shift, True, None, want_mask=True), e * 4.125)
timeit("maxpool_fwd (+argmax)", lambda: ops.maxpool_fwd(z, True), e * 2 + pe * 3)
timeit("fused bn_relu_maxpool_fwd", lambda: ops.bn_relu_maxpool_fwd(y, scale, shift, True), e * 2 + pe * 3)
timeit("maxpool_bwd", lambda: ops.maxpool_bwd(dpool, idx, z.shape), e * 2 + pe * 3)
timeit("bn_backward (reduce+finalize+dx)", lambda: ops.bn_backward(dz, None, y, gamma, mean, invstd, True, mask=mask), e * 10.25)
timeit("fused maxpool_bn_bwd (reduce+finalize+dx)", lambda: ops.maxpool_bn_bwd(dpool, idx, p, y, gamma, mean, invstd), e * 6 + pe * 10)

"""Times the BatchNorm bandwidth kernels (apply +mask, apply +residual, backward reduce+finalize+dx) on the activation
shapes of the N=256 / 160 px ResNet-50 step, CUDA events, L2 flushed between launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from b200net import ops

dev = torch.device("cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[len(ts) // 2]


tot = {"apply": 0.0, "apply_res": 0.0, "bwd": 0.0, "bwd_dres": 0.0}
print("%-16s %22s %22s %22s %22s" % ("rows x C", "apply+mask", "apply+res+mask", "backward", "backward+dres"))
for rows, C, cnt in [(409600, 64, 7), (409600, 256, 4), (102400, 128, 8), (102400, 512, 5), (25600, 256, 12),
                     (25600, 1024, 7), (6400, 512, 6), (6400, 2048, 4)]:
    x = torch.randn(rows, C, device=dev).to(torch.bfloat16)
    r = torch.randn(rows, C, device=dev).to(torch.bfloat16)
    dz = torch.randn(rows, C, device=dev).to(torch.bfloat16)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    mean, invstd, scale, shift = ops.bn_train_stats(x, gamma, beta, None, None, 0.0, 1e-5)
    z, mask = ops.bn_apply(x, scale, shift, True, None, want_mask=True)
    e = rows * C
    t1 = timeit(lambda: ops.bn_apply(x, scale, shift, True, None, want_mask=True))
    t2 = timeit(lambda: ops.bn_apply(x, scale, shift, True, r, want_mask=True))
    t3 = timeit(lambda: ops.bn_backward(dz, None, x, gamma, mean, invstd, True, mask=mask))
    t4 = timeit(lambda: ops.bn_backward(dz, None, x, gamma, mean, invstd, True, mask=mask, want_dres=True))
    print("%-16s %8.1f us %6.0f GB/s %8.1f us %6.0f GB/s %8.1f us %6.0f GB/s %8.1f us %6.0f GB/s" % (
        "%d x %d" % (rows, C), t1, e * 4.125 / t1 / 1e3, t2, e * 6.125 / t2 / 1e3, t3, e * 10.25 / t3 / 1e3, t4, e * 12.25 / t4 / 1e3))
print("BN_STREAM_BENCH DONE")

"""Warm, back-to-back CUDA-event timing of the BERT-step kernels at the C5 per-GPU shapes (N=32, S=256, bf16).
Unlike an ncu launch list (cold caches, serialised) this is what a kernel costs inside the step.
usage: python tools/bert_kernel_bench.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from b200net import ops  # noqa: E402

dev = torch.device("cuda:0")
N, S, H, heads, FF = 32, 256, 768, 12, 3072
M = N * S
dt = torch.bfloat16
g = torch.Generator(device="cpu").manual_seed(0)


def rnd(*shape, dtype=dt, scale=1.0):
    return (torch.randn(*shape, generator=g) * scale).to(dtype).to(dev)


def timeit(name, fn, flops=None, nbytes=None, iters=20):
    """The op is captured `iters` times into one CUDA graph and the replay is timed: no Python / allocator time."""
    for _ in range(3):
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
    extra = ""
    if flops:
        extra += "  %7.1f TFLOP/s" % (flops / us / 1e6)
    if nbytes:
        extra += "  %6.2f TB/s" % (nbytes / us / 1e6)
    print("%-34s %8.1f us%s" % (name, us, extra), flush=True)


x = rnd(M, H)
res = rnd(M, H)
gamma, beta = torch.ones(H, device=dev), torch.zeros(H, device=dev)
wqkv, bqkv = rnd(3 * H, H, scale=0.02), torch.zeros(3 * H, device=dev)
wo, bo = rnd(H, H, scale=0.02), torch.zeros(H, device=dev)
w1, b1 = rnd(FF, H, scale=0.02), torch.zeros(FF, device=dev)
w2, b2 = rnd(H, FF, scale=0.02), torch.zeros(H, device=dev)
mask = torch.ones(N, S, dtype=torch.int64, device=dev)
qkv = ops.dense_fwd(x, wqkv, bqkv)
dctx = rnd(M, H)
pre = ops.dense_fwd(x, w1, b1)
dy_ff = rnd(M, FF)
dqkv = rnd(M, 3 * H)

timeit("dense fwd QKV  768->2304", lambda: ops.dense_fwd(x, wqkv, bqkv), 2.0 * M * H * 3 * H)
timeit("dense fwd out  768->768", lambda: ops.dense_fwd(x, wo, bo), 2.0 * M * H * H)
timeit("dense fwd FFN1 768->3072", lambda: ops.dense_fwd(x, w1, b1), 2.0 * M * H * FF)
timeit("dense fwd FFN2 3072->768", lambda: ops.dense_fwd(pre, w2, b2), 2.0 * M * H * FF)
timeit("dense dgrad FFN2 (->3072)", lambda: ops.dense_bwd_input(x, w2, FF), 2.0 * M * H * FF)
timeit("dense dgrad FFN1 (->768)", lambda: ops.dense_bwd_input(dy_ff, w1, H), 2.0 * M * H * FF)
timeit("dense dgrad QKV (->768)", lambda: ops.dense_bwd_input(dqkv, wqkv, H), 2.0 * M * H * 3 * H)
timeit("dense wgrad FFN2 [768x3072]", lambda: ops.dense_bwd_weight(x, pre), 2.0 * M * H * FF)
timeit("dense wgrad FFN1 [3072x768]", lambda: ops.dense_bwd_weight(dy_ff, x), 2.0 * M * H * FF)
timeit("dense wgrad QKV [2304x768]", lambda: ops.dense_bwd_weight(dqkv, x), 2.0 * M * H * 3 * H)
timeit("dense wgrad out [768x768]", lambda: ops.dense_bwd_weight(dctx, x), 2.0 * M * H * H)
for p in (0.0, 0.1):
    att_flops = 4.0 * S * S * 64 * heads * N
    saved, ctx = ops.attention_flash_fwd(qkv, mask, N, S, heads, p, 1, 0)
    timeit("attention fwd  p=%.1f" % p, lambda: ops.attention_flash_fwd(qkv, mask, N, S, heads, p, 1, 0), att_flops)
    timeit("attention bwd (dQ + dK,dV) p=%.1f" % p,
           lambda: ops.attention_flash_bwd(qkv, mask, saved, ctx, dctx, N, S, heads, p), 2.5 * att_flops)
    s_, y_, mean, rstd = ops.ln_fwd(x.clone(), res, gamma, beta, 1e-12, p, 1, 0)
    timeit("ln fwd  p=%.1f" % p, lambda: ops.ln_fwd(x, res, gamma, beta, 1e-12, p, 1, 0, inplace_sum=False),
           nbytes=4.0 * M * H * 2)
    timeit("ln bwd  p=%.1f" % p, lambda: ops.ln_bwd(dctx, res, s_, mean, rstd, gamma, p, 1, 0),
           nbytes=(5.0 if p > 0 else 4.0) * M * H * 2)
timeit("gelu fwd [M x 3072]", lambda: ops.gelu_fwd(pre), nbytes=2.0 * M * FF * 2)
timeit("gelu bwd + bias grad", lambda: ops.gelu_bwd_bias(dy_ff, pre), nbytes=3.0 * M * FF * 2)
timeit("colsum [M x 2304]", lambda: ops.colsum_any(dqkv), nbytes=1.0 * M * 3 * H * 2)

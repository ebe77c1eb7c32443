"""Times the staged stem convolution (4x1 over 64 "channels", N=256, 80x80 outputs) forward with fused statistics."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from b200net import ops
dev = torch.device("cuda:0")
N = int(os.environ.get("N", "256"))
img = torch.rand(N, 3, 160, 160, device=dev)
w0 = torch.randn(64, 3, 7, 7, device=dev) / 12.0
x4 = ops.to_nhwc(img, torch.bfloat16, 4)
st = ops.stem_stage(x4, 7, 7, 3)
wp0 = ops.stem_pack_weight(w0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
def timeit(fn):
    for _ in range(3): fn()
    ts = []
    for _ in range(10):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[len(ts) // 2]
print("staged tensor", tuple(st.shape), "B200_WS =", os.environ.get("B200_WS"))
print("stem fprop + statistics: %.1f us" % timeit(lambda: ops.conv_fprop_bnstats(st, wp0, 64, 4, 1, 1, 0)))
y, part = ops.conv_fprop_bnstats(st, wp0, 64, 4, 1, 1, 0)
print("checksum", float(y.float().abs().sum()), float(part.sum()))

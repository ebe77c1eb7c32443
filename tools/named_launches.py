"""Runs ONE measured launch of each named kernel of the ResNet-50 training step at the benchmark shapes (N=256, 160 px,
bf16) between cudaProfilerStart/Stop, so that `ncu --set full --profile-from-start off` captures exactly these:

    conv classes S (stem), b, c, f, j, p, v of SURVEY.md 8(a'): fprop (with fused BN statistics), dgrad, wgrad
    bn_apply (+ReLU mask), bn_backward (reduction + finalize + dx), bn_train_stats, mt_sgd, maxpool fwd / bwd,
    the stem's fused BatchNorm -> ReLU -> MaxPool forward / backward

and writes gpurun_out/named_launches_labels.json = [(label, kernels launched, algorithmic bytes, flops)] in launch order
(counted with b200_launch_count), which tools/ncu_named_summary.py joins with the ncu CSV.
usage: ncu --set full --clock-control none --profile-from-start off --csv --page raw ... python tools/named_launches.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from b200net import ops

N = int(os.environ.get("NAMED_N", "256"))
dev = torch.device("cuda:0")
lib = b200net._C.lib
CLASSES = [  # name, H, C, K, R, stride, pad
    ("b", 40, 64, 64, 3, 1, 1), ("c", 40, 64, 256, 1, 1, 0), ("f", 40, 128, 128, 3, 2, 1), ("j", 20, 128, 128, 3, 1, 1),
    ("p", 10, 256, 256, 3, 1, 1), ("v", 5, 512, 512, 3, 1, 1), ("m", 10, 256, 1024, 1, 1, 0),
]
ops_list = []   # (label, fn, algorithmic bytes, flops)

# stem: channel-padded image -> row-pair staging -> 4x1 conv
img = torch.rand(N, 3, 160, 160, device=dev)
w0 = torch.randn(64, 3, 7, 7, device=dev) / 12.0
x4 = ops.to_nhwc(img, torch.bfloat16, 4)
st = ops.stem_stage(x4, 7, 7, 3)
wp0 = ops.stem_pack_weight(w0)
y0 = torch.empty(N, 80, 80, 64, device=dev, dtype=torch.bfloat16)
dy0 = torch.randn(N, 80, 80, 64, device=dev).to(torch.bfloat16)
fl0 = 2.0 * N * 80 * 80 * 64 * 147
ops_list += [("S_stage", lambda: ops.stem_stage(x4, 7, 7, 3), 2.0 * (x4.numel() + st.numel()), 0),
             ("S_fprop", lambda: ops.conv_fprop_bnstats(st, wp0, 64, 4, 1, 1, 0), 2.0 * (st.numel() + y0.numel()), fl0),
             ("S_wgrad", lambda: ops.conv_wgrad(st, dy0, 64, 4, 1, 1, 0), 2.0 * (st.numel() + dy0.numel()), fl0)]
keep = []
for name, H, C, K, R, s, pad in CLASSES:
    Ho = (H + 2 * pad - R) // s + 1
    x = torch.randn(N, H, H, C, device=dev).to(torch.bfloat16)
    w = torch.randn(K, C, R, R, device=dev) / (C * R * R) ** 0.5
    wp = ops.pack_weight(w, torch.bfloat16)
    dy = torch.randn(N, Ho, Ho, K, device=dev).to(torch.bfloat16)
    dx = torch.empty_like(x)
    keep += [x, wp, dy, dx]
    fl = 2.0 * N * Ho * Ho * K * C * R * R
    by = 2.0 * (x.numel() + dy.numel() + wp.numel())
    byw = 2.0 * (x.numel() + dy.numel()) + 4.0 * w.numel()
    ops_list += [("%s_fprop" % name, (lambda x=x, wp=wp, K=K, R=R, s=s, pad=pad: ops.conv_fprop_bnstats(x, wp, K, R, R, s, pad)), by, fl),
                 ("%s_dgrad" % name, (lambda x=x, wp=wp, dy=dy, dx=dx, K=K, R=R, s=s, pad=pad: ops.conv_dgrad(dy, wp, x.shape, K, R, R, s, pad, out=dx)), by, fl),
                 ("%s_wgrad" % name, (lambda x=x, dy=dy, K=K, R=R, s=s, pad=pad: ops.conv_wgrad(x, dy, K, R, R, s, pad)), byw, fl)]
# bandwidth kernels on layer1's block output
rows, C = N * 40 * 40, 256
xb = torch.randn(rows, C, device=dev).to(torch.bfloat16)
dzb = torch.randn(rows, C, device=dev).to(torch.bfloat16)
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, invstd, scale, shift = ops.bn_train_stats(xb, gamma, beta, None, None, 0.0, 1e-5)
zb, maskb = ops.bn_apply(xb, scale, shift, True, None, want_mask=True)
e = rows * C
ops_list += [("bn_apply", lambda: ops.bn_apply(xb, scale, shift, True, None, want_mask=True), e * 4.125, 0),
             ("bn_backward", lambda: ops.bn_backward(dzb, None, xb, gamma, mean, invstd, True, mask=maskb), e * 10.25, 0),
             ("bn_stats", lambda: ops.bn_train_stats(xb, gamma, beta, None, None, 0.0, 1e-5), e * 2.0, 0)]
zs = torch.relu(torch.randn(N, 80, 80, 64, device=dev)).to(torch.bfloat16)
pooled, idx = ops.maxpool_fwd(zs, True)
dpool = torch.randn_like(pooled)
ops_list += [("maxpool_fwd", lambda: ops.maxpool_fwd(zs, True), 2.0 * (zs.numel() + pooled.numel()) + idx.numel(), 0),
             ("maxpool_bwd", lambda: ops.maxpool_bwd(dpool, idx, zs.shape), 2.0 * (zs.numel() + pooled.numel()) + idx.numel(), 0)]
# the stem's BatchNorm -> ReLU -> MaxPool as one pass each way (csrc/stem_pool.cu), on the stem output shape
ys = torch.randn(N, 80, 80, 64, device=dev).to(torch.bfloat16)
g64, b64 = torch.ones(64, device=dev), torch.zeros(64, device=dev)
mean_s, invstd_s, scale_s, shift_s = ops.bn_train_stats(ys, g64, b64, None, None, 0.0, 1e-5)
pooled_s, idx_s = ops.bn_relu_maxpool_fwd(ys, scale_s, shift_s, True)
dpool_s = torch.randn_like(pooled_s)
es, ps_ = ys.numel(), pooled_s.numel()
ops_list += [("stem_pool_fwd", lambda: ops.bn_relu_maxpool_fwd(ys, scale_s, shift_s, True), 2.0 * es + 3.0 * ps_, 0),
             ("stem_pool_bwd", lambda: ops.maxpool_bn_bwd(dpool_s, idx_s, pooled_s, ys, g64, mean_s, invstd_s),
              6.0 * es + 10.0 * ps_, 0)]
params = [torch.nn.Parameter(torch.randn(n, device=dev)) for n in (2048 * 512 * 9 // 4, 512 * 2048, 1024 * 256, 64 * 64)]
for p in params:
    p.grad = torch.randn_like(p)
opt = b200net.optim.SGD(params, lr=0.001)
ops_list += [("mt_sgd", lambda: opt.step(), 12.0 * sum(p.numel() for p in params), 0)]

if os.environ.get("NAMED_ONLY"):     # e.g. NAMED_ONLY=b_fprop,b_dgrad for a source-level capture of one kernel
    want = set(os.environ["NAMED_ONLY"].split(","))
    ops_list = [o for o in ops_list if o[0] in want]
for _, fn, _, _ in ops_list:      # warm-up (sizes the workspaces, sets function attributes)
    fn(); fn()
torch.cuda.synchronize()
labels = []
torch.cuda.profiler.start()
for label, fn, by, fl in ops_list:
    c0 = lib.b200_launch_count()
    fn()
    labels.append((label, int(lib.b200_launch_count() - c0), by, fl))
torch.cuda.synchronize()
torch.cuda.profiler.stop()
os.makedirs("gpurun_out", exist_ok=True)
json.dump(labels, open("gpurun_out/named_launches_labels.json", "w"))
print("measured %d ops, %d kernels" % (len(labels), sum(l[1] for l in labels)))

"""One launch of each stem pooling kernel between cudaProfilerStart/Stop (for ncu --profile-from-start off)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from b200net import ops
dev = torch.device("cuda:0")
N, H, W, C = 256, 80, 80, 64
y = torch.randn(N, H, W, C, device=dev).to(torch.bfloat16)
gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
mean, invstd, scale, shift = ops.bn_train_stats(y, gamma, beta, None, None, 0.0, 1e-5)
p, idx = ops.bn_relu_maxpool_fwd(y, scale, shift, True)
dpool = torch.randn_like(p)
ops.maxpool_bn_bwd(dpool, idx, p, y, gamma, mean, invstd)
torch.cuda.synchronize()
torch.cuda.profiler.start()
ops.bn_relu_maxpool_fwd(y, scale, shift, True)
ops.maxpool_bn_bwd(dpool, idx, p, y, gamma, mean, invstd)
torch.cuda.synchronize()
torch.cuda.profiler.stop()

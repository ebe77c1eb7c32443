"""Diagnostic (not a test): loss curves of ours(fp32), ours(bf16), oracle fp32 CPU and oracle fp64 CPU."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn
import b200net
from oracle import torch_oracle as O

def ours(prec, n, hw, steps):
    b200net.set_precision(prec)
    ref = O.build_reference_model(dropout=False)
    m = b200net.models.resnet50(); m.fc = O.make_head(); m.fc[2].p = 0.0
    m.load_state_dict(ref.state_dict()); m.to("cuda:0").train()
    opt = b200net.optim.SGD(m.parameters(), lr=0.001); crit = nn.NLLLoss()
    x, y = O.make_batch(n, hw, 0, "cuda:0")
    return [O.train_step(m, opt, crit, x, y)[1] for _ in range(steps)]

def oracle(dtype, n, hw, steps):
    m = O.build_reference_model(dropout=False).to(dtype)
    opt, crit = O.mode_a(m)
    x, y = O.make_batch(n, hw, 0); x = x.to(dtype)
    m.train()
    return [O.train_step(m, opt, crit, x, y)[1] for _ in range(steps)]

torch.set_num_threads(O.cpu_threads())
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
for n, hw in ((8, 64), (16, 96), (32, 128)):
    t=time.time(); o32 = oracle(torch.float32, n, hw, steps); t32=time.time()-t
    o64 = oracle(torch.float64, n, hw, steps)
    f32 = ours("fp32", n, hw, steps); b16 = ours("bf16", n, hw, steps)
    rel = lambda a, b: max(abs(x - y) / max(1.0, abs(y)) for x, y in zip(a, b))
    print("N=%d hw=%d steps=%d (oracle fp32 %.1fs): max rel dev vs fp64: oracle32 %.2e ours32 %.2e oursbf16 %.2e | oursbf16 vs oracle32 %.2e ours32 vs oracle32 %.2e" %
          (n, hw, steps, t32, rel(o32, o64), rel(f32, o64), rel(b16, o64), rel(b16, o32), rel(f32, o32)))
    for s in (0, 1, 2, 5, 10, 20, 50, steps - 1):
        if s < steps: print("   step %3d  o64 %.5f o32 %.5f f32 %.5f bf16 %.5f" % (s, o64[s], o32[s], f32[s], b16[s]))

"""Experiment: the whole training step (zero_grad, forward, loss, backward, SGD) captured in ONE CUDA graph and
replayed, against the eager step.  Tells how much of the eager step is launch/Python gaps."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from oracle import torch_oracle as O

dev = torch.device("cuda:0")
torch.manual_seed(42)
def build():
    torch.manual_seed(42)
    m = b200net.models.resnet50()
    m.fc = b200net.nn.Sequential(b200net.nn.Linear(2048, 512), b200net.nn.ReLU(), b200net.nn.Dropout(0.0),
                                 b200net.nn.Linear(512, 10), b200net.nn.LogSoftmax(dim=1))
    return m.to(dev).train()
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
x, y = O.make_batch(B, 160, 7, dev)
crit = b200net.nn.NLLLoss()

def timed(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n

# eager
m1 = build(); o1 = b200net.optim.SGD(m1.parameters(), lr=0.001)
def eager():
    o1.zero_grad(); l = crit(m1(x), y); l.backward(); o1.step(); return l
losses_e = [eager().item() for _ in range(5)]
t_e = timed(eager)
t0 = time.perf_counter(); 
for _ in range(20): eager()
cpu_enqueue = (time.perf_counter() - t0) / 20 * 1e3
torch.cuda.synchronize()

# graphed
m2 = build(); o2 = b200net.optim.SGD(m2.parameters(), lr=0.001)
sx, sy = x.clone(), y.clone()
s = torch.cuda.Stream()
s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3):
        o2.zero_grad(set_to_none=True); l = crit(m2(sx), sy); l.backward(); o2.step()
torch.cuda.current_stream().wait_stream(s)
m2 = build(); o2 = b200net.optim.SGD(m2.parameters(), lr=0.001)   # fresh weights, caches warm
with torch.cuda.stream(s):
    o2.zero_grad(set_to_none=True); l = crit(m2(sx), sy); l.backward(); o2.step()
torch.cuda.current_stream().wait_stream(s)
torch.cuda.synchronize()
m2b = build()
m2.load_state_dict(m2b.state_dict())
for p in m2.parameters(): torch.autograd.graph.increment_version(p)
g = torch.cuda.CUDAGraph()
o2.zero_grad(set_to_none=True)
with torch.cuda.graph(g):
    sl = crit(m2(sx), sy); sl.backward(); o2.step()
losses_g = []
for _ in range(5):
    g.replay(); losses_g.append(sl.item())
t_g = timed(lambda: g.replay())
print("batch %d: eager %.3f ms/step (CPU enqueue %.3f ms), graph replay %.3f ms/step" % (B, t_e, cpu_enqueue, t_g))
print("eager losses", ["%.5f" % v for v in losses_e])
print("graph losses", ["%.5f" % v for v in losses_g])

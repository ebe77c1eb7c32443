"""Where does the end-to-end step (host batches, H2D every step, loss.item() every step) lose time against the
device-resident step?  Variants of the loader and of the read-back, CUDA-event timed over 20 steps each."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import b200net
from oracle import torch_oracle as O

dev = torch.device("cuda:0")
torch.manual_seed(1)
m = b200net.models.resnet50()
m.fc = b200net.nn.Sequential(b200net.nn.Linear(2048, 512), b200net.nn.ReLU(), b200net.nn.Dropout(0.2),
                             b200net.nn.Linear(512, 10), b200net.nn.LogSoftmax(dim=1))
m.to(dev).train()
opt = b200net.optim.SGD(m.parameters(), lr=0.001)
crit = b200net.nn.NLLLoss()
B = 256
x, y = O.make_batch(B, 160, 7, torch.device("cpu"))
xh, yh = x.pin_memory(), y.pin_memory()
xd, yd = xh.to(dev), yh.to(dev)
K = 20


def step(inputs, labels, item=True):
    opt.zero_grad()
    loss = crit(m.forward(inputs), labels)
    loss.backward()
    opt.step()
    return loss.item() if item else loss


def timed(fn):
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record(); fn(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / K, (time.perf_counter() - t0) * 1e3 / K


for _ in range(8):
    step(xd, yd)
print("resident, no item      : %.2f ms (wall %.2f)" % timed(lambda: [step(xd, yd, False) for _ in range(K)]))
print("resident, item         : %.2f ms (wall %.2f)" % timed(lambda: [step(xd, yd, True) for _ in range(K)]))
print("verbatim .to() + item  : %.2f ms (wall %.2f)" % timed(lambda: [step(xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True)) for _ in range(K)]))


def with_prefetcher(item):
    loader = b200net.utils.DevicePrefetcher(((xh, yh) for _ in range(K)), dev)
    for inputs, labels in loader:
        step(inputs.to(dev), labels.to(dev), item)
for it in (False, True):
    print("prefetcher, item=%-5s : %.2f ms (wall %.2f)" % ((it,) + timed(lambda: with_prefetcher(it))))
    print("prefetcher, item=%-5s : %.2f ms (wall %.2f)" % ((it,) + timed(lambda: with_prefetcher(it))))
print("E2E PROBE DONE")

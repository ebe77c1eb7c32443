"""cProfile of the host side of the ResNet training step (the reference loop body with loss.item() every step):
where the ~1.3 ms per step that the GPU waits for the CPU goes.  usage: python tools/cpu_profile_step.py"""
import cProfile
import os
import pstats
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import b200net  # noqa: E402
from oracle import torch_oracle as O  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(42)
model = b200net.models.resnet50()
model.fc = b200net.nn.Sequential(b200net.nn.Linear(2048, 512), b200net.nn.ReLU(), b200net.nn.Dropout(0.2),
                                 b200net.nn.Linear(512, 10), b200net.nn.LogSoftmax(dim=1))
model.to(dev).train()
crit = b200net.nn.NLLLoss()
opt = b200net.optim.SGD(model.parameters(), lr=0.001)
x, y = O.make_batch(256, 160, 1, dev)


def step():
    opt.zero_grad()
    loss = crit(model.forward(x), y)
    loss.backward()
    opt.step()
    return loss.item()


for _ in range(8):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(30):
    step()
pr.disable()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(28)
